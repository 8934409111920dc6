set -x; mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_train_gpu.py -m gpu -q --timeout 900 > gpurun_out/pytest_train.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_train.log
tail -60 gpurun_out/pytest_train.log
