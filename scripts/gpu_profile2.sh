#!/bin/bash
# ncu --set full captures of the GroupNorm and wgrad kernels at the bench micro-batch (256 images of C4).
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
CMD="python bench.py --workload c4 --batch-per-gpu 256 --micro-batch 256 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline"
timeout 300 $CMD > gpurun_out/plain_p2.log 2>&1 || { echo "plain rc=$?" >> gpurun_out/rc.txt; exit 0; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gn_(stats|apply)" -c 12 -f -o gpurun_out/prof_gnf $CMD > gpurun_out/ncu_gnf.log 2>&1
echo "gnf rc=$?" >> gpurun_out/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gn_bwd -c 12 -f -o gpurun_out/prof_gnb $CMD > gpurun_out/ncu_gnb.log 2>&1
echo "gnb rc=$?" >> gpurun_out/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_reduce_umma -s 2 -c 8 -f -o gpurun_out/prof_wg $CMD > gpurun_out/ncu_wg.log 2>&1
echo "wg rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
