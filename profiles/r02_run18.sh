mkdir -p gpurun_out
for pm in 48 32 8; do
B200_CONV_PAIR_MIN=$pm python profiles/shape_breakdown.py 128 > gpurun_out/shapes_b128_pm$pm.txt 2>&1; head -1 gpurun_out/shapes_b128_pm$pm.txt; grep -E "8x8|4x4" gpurun_out/shapes_b128_pm$pm.txt | grep conv | head -8
done
