set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_score_gpu.py -m gpu -q --timeout 600 > gpurun_out/pytest_score.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_score.log
tail -5 gpurun_out/pytest_score.log
B200_FUSE_GN=0 python profiles/shape_breakdown.py 1024 > gpurun_out/shapes_gn0.txt 2>&1
python profiles/shape_breakdown.py 1024 > gpurun_out/shapes_gn1.txt 2>&1
B200_FUSE_GN=0 python profiles/shape_breakdown.py 128 > gpurun_out/shapes128_gn0.txt 2>&1
python profiles/shape_breakdown.py 128 > gpurun_out/shapes128_gn1.txt 2>&1
