mkdir -p gpurun_out; python profiles/prof_train.py 128 > gpurun_out/prof_train.txt 2>&1; cat gpurun_out/prof_train.txt
