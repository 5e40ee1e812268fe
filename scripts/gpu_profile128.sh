#!/bin/bash
# ncu --set full of the 128-channel-tile launches: first forward gather launches (conv_in, encoder level-0 convs),
# first backward gather launches (decoder level-0 dgrads with the GroupNorm-backward epilogue) and the first wgrad
# launches (decoder level 0: <1,128,3>) of one C4 micro-batch of 256.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
CMD="python bench.py --workload c4 --batch-per-gpu 256 --micro-batch 256 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline $BENCH_ARGS"
timeout 300 $CMD > gpurun_out/plain_p128.log 2>&1 || { echo "plain rc=$?" >> gpurun_out/rc.txt; exit 0; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_gather_umma -s 0 -c 6 -f -o gpurun_out/prof_g128f $CMD > gpurun_out/ncu_g128f.log 2>&1
echo "g128f rc=$?" >> gpurun_out/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_gather_umma -s 88 -c 8 -f -o gpurun_out/prof_g128b $CMD > gpurun_out/ncu_g128b.log 2>&1
echo "g128b rc=$?" >> gpurun_out/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_reduce_umma -s 0 -c 6 -f -o gpurun_out/prof_w128 $CMD > gpurun_out/ncu_w128.log 2>&1
echo "w128 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
