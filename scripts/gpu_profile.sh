#!/bin/bash
# ncu evidence pass (one GPU): launch list of a short bench command, then --set full captures of
# the two tcgen05 conv kernels.  Each ncu run follows a plain run of the same command that exited 0.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
CMD="python bench.py --workload c4 --batch-per-gpu 64 --micro-batch 64 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?" >> gpurun_out/rc.txt
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_gather_umma -s 60 -c 3 -f -o gpurun_out/prof_gather $CMD > gpurun_out/ncu_gather.log 2>&1
echo "gather rc=$?" >> gpurun_out/rc.txt
timeout 300 $CMD > gpurun_out/plain3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_reduce_umma -s 60 -c 3 -f -o gpurun_out/prof_reduce $CMD > gpurun_out/ncu_reduce.log 2>&1
echo "reduce rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -n 5 gpurun_out/plain.log gpurun_out/ncu_gather.log
