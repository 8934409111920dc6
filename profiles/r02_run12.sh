mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu_full.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_full.log
tail -12 gpurun_out/pytest_gpu_full.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -4 gpurun_out/smoke.log
