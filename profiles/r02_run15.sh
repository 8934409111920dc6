mkdir -p gpurun_out
python profiles/shape_breakdown.py 128 > gpurun_out/shapes_b128.txt 2>&1; head -3 gpurun_out/shapes_b128.txt
python profiles/shape_breakdown.py 256 > gpurun_out/shapes_b256.txt 2>&1; head -1 gpurun_out/shapes_b256.txt
