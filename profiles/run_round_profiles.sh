# Round-end evidence: bench lines (both arms), the ncu launch list of the bench command, per-launch DRAM bytes of every
# conv_tc launch of one forward, one `--set full` capture.   bash profiles/run_round_profiles.sh <tag>
TAG=${1:-r02f}
set -x
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; tail -c 300 gpurun_out/${TAG}_bench_n1.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>/dev/null
CMD="python bench.py --steps 2 --warmup 3 --e2e-steps 4 --no-cpu-baseline --no-aux --no-ref-gpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo list rc=$?
python profiles/summarize_ncu.py launches gpurun_out/${TAG}_launches.csv "$CMD (B=1024)" > gpurun_out/${TAG}_launches_summary.txt; gzip -f gpurun_out/${TAG}_launches.csv
CMD1="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-aux --no-ref-gpu"
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_bytes.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:conv_tc_kernel -s 210 -c 105 --csv --log-file gpurun_out/${TAG}_conv_tc_dram.csv $CMD1 > gpurun_out/ncu_dram.log 2>&1; echo dram rc=$?
ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 230 -c 6 -o gpurun_out/${TAG}_conv_tc_full $CMD1 > gpurun_out/ncu_full.log 2>&1; echo full rc=$?
python profiles/summarize_ncu.py full gpurun_out/${TAG}_conv_tc_full.ncu-rep > gpurun_out/${TAG}_conv_tc_ncu_full.txt 2>gpurun_out/summ.err
ls -la gpurun_out/*.ncu-rep; SZ=$(stat -c %s gpurun_out/${TAG}_conv_tc_full.ncu-rep); if [ "$SZ" -gt 20000000 ]; then rm gpurun_out/${TAG}_conv_tc_full.ncu-rep; fi
du -sh gpurun_out
