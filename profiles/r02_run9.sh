set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q --timeout 800 > gpurun_out/pytest_mgpu2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_mgpu2.log
tail -25 gpurun_out/pytest_mgpu2.log
python - > gpurun_out/train_bench.txt 2>&1 <<'PY'
import sys, time, torch
sys.path.insert(0, '.')
import bench
dev = torch.device('cuda:0')
def timeit(fn, iters=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(iters): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / iters * 1e3
r = bench.other_config_measurements(dev, timeit)
print(r['C3_training_step'])
PY
cat gpurun_out/train_bench.txt | tail -5
