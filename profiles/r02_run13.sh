mkdir -p gpurun_out
for n in 8 4 2; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 --no-aux > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "n=$n rc=$?"
done
