mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_train_gpu.py tests/test_input_grad_gpu.py -m gpu -q --timeout 600 > gpurun_out/pytest_train.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_train.log
tail -5 gpurun_out/pytest_train.log
python profiles/prof_train.py 128 > gpurun_out/prof_train.txt 2>&1; cat gpurun_out/prof_train.txt
