mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_tc_gpu.py tests/test_conv_gn_gpu.py tests/test_conv_sc_gpu.py tests/test_score_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_conv.log 2>&1; tail -3 gpurun_out/pytest_conv.log
python profiles/shape_breakdown.py 128 > gpurun_out/shapes_b128_b.txt 2>&1; head -1 gpurun_out/shapes_b128_b.txt
python profiles/shape_breakdown.py 1024 > gpurun_out/shapes_b1024_b.txt 2>&1; head -1 gpurun_out/shapes_b1024_b.txt
python bench.py --steps 20 --warmup 5 --no-aux --no-cpu-baseline --no-ref-gpu --e2e-steps 20 > gpurun_out/bench_quick.json 2>gpurun_out/bench_quick.err; python -c "
import json; d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['clocks'], d['roofline']['achieved'])"
python bench.py --batch-per-gpu 128 --steps 40 --warmup 5 --no-aux --no-cpu-baseline --no-ref-gpu --e2e-steps 20 > gpurun_out/bench_quick128.json 2>gpurun_out/bench_quick128.err; python -c "
import json; d=json.loads(open('gpurun_out/bench_quick128.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['clocks'])"
