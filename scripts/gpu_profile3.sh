#!/bin/bash
# ncu --set full of the gather kernel: first forward launches (EPI_GN_STATS / plain) and backward launches (EPI_GN_BWD)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
CMD="python bench.py --workload c4 --batch-per-gpu 256 --micro-batch 256 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline"
timeout 300 $CMD > gpurun_out/plain_p3.log 2>&1 || { echo "plain rc=$?" >> gpurun_out/rc.txt; exit 0; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_gather_umma -s 4 -c 10 -f -o gpurun_out/prof_gf $CMD > gpurun_out/ncu_gf.log 2>&1
echo "gf rc=$?" >> gpurun_out/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_gather_umma -s 152 -c 24 -f -o gpurun_out/prof_gb $CMD > gpurun_out/ncu_gb.log 2>&1
echo "gb rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
