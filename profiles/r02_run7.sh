set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ctrl_gpu.py tests/test_conv_sc_gpu.py tests/test_conv_gn_gpu.py -m gpu -q --timeout 600 > gpurun_out/pytest_ctrl.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ctrl.log
tail -15 gpurun_out/pytest_ctrl.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"
tail -5 gpurun_out/bench_n1.err
