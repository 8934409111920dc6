set -x; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
B200_FUSE_GN=0 timeout 300 python bench.py --steps 20 --warmup 5 --no-aux --no-cpu-baseline --e2e-steps 20 > gpurun_out/bench_gn0.json 2> gpurun_out/bench_gn0.err
timeout 300 python bench.py --steps 20 --warmup 5 --no-aux --no-cpu-baseline --e2e-steps 20 > gpurun_out/bench_gn1.json 2> gpurun_out/bench_gn1.err
tail -4 gpurun_out/pytest_noalloc_alone.log; tail -30 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/bench_gn1.err
