#!/bin/bash
# One full evidence trip (one GPU): the -m gpu suite, smoke, the default bench line, then the ncu launch
# list and --set full captures of the dominant kernels at the bench micro-batch (256 images of C4).
# Every ncu pass follows a plain run of the same command that exited 0.  usage: scripts/gpu_round.sh [stages...]
mkdir -p gpurun_out
rm -f gpurun_out/rc.txt
export PYTHONUNBUFFERED=1
STAGES=${@:-tests smoke bench list full}
CMD="python bench.py --workload c4 --batch-per-gpu 256 --micro-batch 256 --steps 1 --warmup 1 --no-cpu-baseline --no-roofline"
for stage in $STAGES; do
  case $stage in
    tests) timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?" >> gpurun_out/rc.txt;;
    smoke) timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt;;
    bench) timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.txt;;
    benchref) timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "benchref rc=$?" >> gpurun_out/rc.txt;;
    bench_c5) timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1; echo "bench_c5 rc=$?" >> gpurun_out/rc.txt;;
    bench_c3) timeout 600 python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3.log 2>&1; echo "bench_c3 rc=$?" >> gpurun_out/rc.txt;;
    bench_c2) timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2.log 2>&1; echo "bench_c2 rc=$?" >> gpurun_out/rc.txt;;
    bench_c1) timeout 600 python bench.py --workload c1 --precision fp32 --steps 10 --warmup 3 > gpurun_out/bench_c1.log 2>&1; echo "bench_c1 rc=$?" >> gpurun_out/rc.txt;;
    torchbase)   # the GPU baseline to beat: the reference's arithmetic on stock cuDNN / cuBLAS / ATen kernels
      timeout 900 python tests/gpu_torch_baseline.py --workload c4 --steps 5 --warmup 2 > gpurun_out/torch_baseline_c4.log 2>&1; echo "torchbase rc=$?" >> gpurun_out/rc.txt;;
    list)
      timeout 300 $CMD > gpurun_out/plain_l256.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
      echo "list rc=$?" >> gpurun_out/rc.txt;;
    gbonly)   # only the dgrad + GroupNorm-backward launches of the backward pass (short trips)
      timeout 300 $CMD > gpurun_out/plain_full.log 2>&1 || { echo "plain rc=$?" >> gpurun_out/rc.txt; continue; }
      timeout 600 ncu --set full --clock-control none -k regex:conv_gather_umma -s 152 -c 8 -f -o gpurun_out/prof_gb $CMD > gpurun_out/ncu_gb.log 2>&1
      echo "gb rc=$?" >> gpurun_out/rc.txt;;
    full)
      timeout 300 $CMD > gpurun_out/plain_full.log 2>&1 || { echo "plain rc=$?" >> gpurun_out/rc.txt; continue; }
      # (merged results must stay under 64 MiB: no --import-source here, few launches per report)
      timeout 600 ncu --set full --clock-control none -k regex:conv_gather_umma -s 4 -c 6 -f -o gpurun_out/prof_gf $CMD > gpurun_out/ncu_gf.log 2>&1
      echo "gf rc=$?" >> gpurun_out/rc.txt
      timeout 600 ncu --set full --clock-control none -k regex:conv_gather_umma -s 152 -c 8 -f -o gpurun_out/prof_gb $CMD > gpurun_out/ncu_gb.log 2>&1
      echo "gb rc=$?" >> gpurun_out/rc.txt
      timeout 600 ncu --set full --clock-control none -k regex:conv_reduce_umma -s 2 -c 4 -f -o gpurun_out/prof_wg $CMD > gpurun_out/ncu_wg.log 2>&1
      echo "wg rc=$?" >> gpurun_out/rc.txt
      timeout 600 ncu --set full --clock-control none -k regex:"gn_(apply_pre|bwd_pre)" -s 4 -c 4 -f -o gpurun_out/prof_gn $CMD > gpurun_out/ncu_gn.log 2>&1
      echo "gn rc=$?" >> gpurun_out/rc.txt;;
  esac
done
# gpurun merges at most 64 MiB back (and nothing at all beyond that): drop the largest reports until it fits
while [ "$(du -sm gpurun_out | cut -f1)" -ge 60 ]; do
  big=$(ls -S gpurun_out/*.ncu-rep 2>/dev/null | head -1); [ -z "$big" ] && break
  echo "dropping $big ($(du -m "$big" | cut -f1) MiB) to stay under the merge limit" >> gpurun_out/rc.txt; rm -f "$big"
done
cat gpurun_out/rc.txt
tail -n 12 gpurun_out/t_gpu.log gpurun_out/smoke.log gpurun_out/bench*.log 2>/dev/null
