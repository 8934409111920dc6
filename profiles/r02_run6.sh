set -x; mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_conv_sc_gpu.py tests/test_conv_gn_gpu.py tests/test_conv_tc_gpu.py tests/test_score_gpu.py tests/test_input_grad_gpu.py tests/test_blocks_gpu.py -m gpu -q --timeout 900 > gpurun_out/pytest_sc.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_sc.log
tail -40 gpurun_out/pytest_sc.log
for sc in 0 1; do for b in 1024 128; do
  B200_FUSE_SC=$sc timeout 300 python bench.py --batch-per-gpu $b --steps 30 --warmup 5 --no-aux --no-cpu-baseline --e2e-steps 20 > gpurun_out/sc${sc}_b$b.json 2> gpurun_out/sc${sc}_b$b.err
done; done
python profiles/shape_breakdown.py 1024 > gpurun_out/shapes_sc1.txt 2>&1
