mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 --no-aux --no-cpu-baseline --no-ref-gpu --e2e-steps 20 > gpurun_out/bench_q1024.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/bench_q1024.json').read().strip().splitlines()[-1]); print('B=1024', d['value'], d['ms_per_step'], d['clocks']['sm_mhz'], d['forward_ms_sum_of_kernels'])"
python bench.py --batch-per-gpu 128 --steps 40 --warmup 5 --no-aux --no-cpu-baseline --no-ref-gpu --e2e-steps 20 > gpurun_out/bench_q128.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/bench_q128.json').read().strip().splitlines()[-1]); print('B=128', d['value'], d['ms_per_step'])"
