#!/bin/bash
# ncu launch list (durations only) of one micro-batch-256 C4 step: real per-kernel shares at the bench size
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
CMD="python bench.py --workload c4 --batch-per-gpu 256 --micro-batch 256 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline"
timeout 300 $CMD > gpurun_out/plain_l256.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches256.csv $CMD > gpurun_out/ncu_list256.log 2>&1
echo "list256 rc=$?" >> gpurun_out/rc.txt
