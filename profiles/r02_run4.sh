set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_conv_gn_gpu.py tests/test_zz_no_alloc.py -m gpu -q --timeout 800 > gpurun_out/pytest_mgpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_mgpu.log
tail -30 gpurun_out/pytest_mgpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "rc=$?"
tail -5 gpurun_out/bench_n2.err
B200_PEER_XCHG=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-weak --e2e-steps 20 > gpurun_out/bench_n2_nccl.json 2> gpurun_out/bench_n2_nccl.err; echo "rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"
tail -5 gpurun_out/bench_n1.err
