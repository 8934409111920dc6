set -x; mkdir -p gpurun_out
for b in 128 256; do for pdl in 0 1; do
  B200_PDL=$pdl timeout 300 python bench.py --batch-per-gpu $b --steps 40 --warmup 5 --no-aux --no-cpu-baseline --e2e-steps 20 > gpurun_out/pdl${pdl}_b$b.json 2> gpurun_out/pdl${pdl}_b$b.err
done; done
B200_PDL=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-aux --no-cpu-baseline --e2e-steps 20 > gpurun_out/pdl1_b1024.json 2> gpurun_out/pdl1_b1024.err
tail -3 gpurun_out/pdl1_b128.err
