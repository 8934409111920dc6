mkdir -p gpurun_out
DUMP_TAGS=gpurun_out/plan_tags_b128.txt ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches_b128.csv python profiles/shape_breakdown.py 128 > gpurun_out/ncu_b128.log 2>&1; echo rc=$?
tail -2 gpurun_out/ncu_b128.log
