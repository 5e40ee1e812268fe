#!/bin/bash
# Round-2 evidence trip A (one GPU): the whole -m gpu suite (incl. the real-width config parity and Appendix-B kernel
# cases), smoke, the reference on stock torch kernels (the GPU baseline to beat), the bench lines of every workload, then
# the ncu launch list and --set full captures at the BENCHMARKED launch shape (1024-image pass of C4).
# Every ncu pass follows a plain run of the same command that exited 0.   usage: scripts/gpu_trip_r2a.sh [stages...]
mkdir -p gpurun_out
rm -f gpurun_out/rc.txt
export PYTHONUNBUFFERED=1
STAGES=${@:-tests smoke torchbase bench bench_c3 bench_c2 bench_c5 list full}
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline --no-roofline --no-parity-check"
rep2csv() {  # keep the counters, drop the (large) report
  for r in gpurun_out/prof_*.ncu-rep; do
    [ -f "$r" ] || continue
    ncu -i "$r" --page raw --csv > "${r%.ncu-rep}.csv" 2>/dev/null && rm -f "$r"
  done
}
for stage in $STAGES; do
  case $stage in
    tests) timeout 2400 python -m pytest tests -q -m gpu --timeout 900 -rf > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?" >> gpurun_out/rc.txt;;
    tests_new) timeout 2400 python -m pytest tests/test_configs_gpu.py tests/test_appendix_b_gpu.py -q -m gpu --timeout 900 -rf -s > gpurun_out/t_new.log 2>&1; echo "tests_new rc=$?" >> gpurun_out/rc.txt;;
    attn) timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_appendix_b_gpu.py -q -m gpu -k "attention" --timeout 300 -rf > gpurun_out/t_attn.log 2>&1; echo "attn rc=$?" >> gpurun_out/rc.txt; tail -n 25 gpurun_out/t_attn.log;;
    smoke) timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt;;
    torchbase)
      timeout 900 python tests/gpu_torch_baseline.py --workload c4 --steps 4 --warmup 3 > gpurun_out/torch_baseline_c4.log 2>&1; echo "torchbase c4 rc=$?" >> gpurun_out/rc.txt
      timeout 600 python tests/gpu_torch_baseline.py --workload c3 --steps 6 --warmup 3 > gpurun_out/torch_baseline_c3.log 2>&1; echo "torchbase c3 rc=$?" >> gpurun_out/rc.txt;;
    bench) timeout 1200 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.txt;;
    benchref) timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "benchref rc=$?" >> gpurun_out/rc.txt;;
    bench_c5) timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.log 2>&1; echo "bench_c5 rc=$?" >> gpurun_out/rc.txt;;
    bench_c3) timeout 600 python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3.log 2>&1; echo "bench_c3 rc=$?" >> gpurun_out/rc.txt;;
    bench_c2) timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_c2.log 2>&1; echo "bench_c2 rc=$?" >> gpurun_out/rc.txt;;
    bench_c1) timeout 600 python bench.py --workload c1 --precision fp32 --steps 10 --warmup 3 --no-gpu-baseline > gpurun_out/bench_c1.log 2>&1; echo "bench_c1 rc=$?" >> gpurun_out/rc.txt;;
    ab128)  # A/B on one box: CTA pairs for the 128-channel tiles (knob 13 = 1 switches them off)
      timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_pairs128_on.log 2>&1; echo "ab on rc=$?" >> gpurun_out/rc.txt
      timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline --knob 13=1 > gpurun_out/bench_pairs128_off.log 2>&1; echo "ab off rc=$?" >> gpurun_out/rc.txt;;
    ab_ts)  # A/B on one box: attention with P / dS in tensor memory (knob 15 = 1) against the shared-memory tiles
      timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_attn_ts_on.log 2>&1; echo "ts on rc=$?" >> gpurun_out/rc.txt
      timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline --knob 15=1 > gpurun_out/bench_attn_ts_off.log 2>&1; echo "ts off rc=$?" >> gpurun_out/rc.txt;;
    ab_cs)  # A/B on one box: 16 epilogue warps for the GN-statistics fprop on 128-channel tiles (knob 16 = 1: off)
      timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_stats16_on.log 2>&1; echo "cs on rc=$?" >> gpurun_out/rc.txt
      timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline --knob 16=1 > gpurun_out/bench_stats16_off.log 2>&1; echo "cs off rc=$?" >> gpurun_out/rc.txt;;
    halo)   # probe of the halo variant of the 128-channel gather kernel, then an A/B on this box if it is correct
      timeout 600 python scripts/halo_probe.py > gpurun_out/halo_probe.log 2>&1; echo "halo probe rc=$?" >> gpurun_out/rc.txt
      cat gpurun_out/halo_probe.log
      HK=$(python - <<'PY'
import re
ok = {"halo, base offset 0": True, "halo, base offset = (addr>>7)&7": True}
seen = set()
for line in open("gpurun_out/halo_probe.log"):
    for k in ok:
        if line.startswith(k):
            seen.add(k)
            m = re.search(r"fprop ([0-9.e+-]+) dgrad ([0-9.e+-]+)", line)
            if not m or float(m.group(1)) > 2e-2 or float(m.group(2)) > 2e-2: ok[k] = False
if "halo, base offset 0" in seen and ok["halo, base offset 0"]: print("18=0")
elif "halo, base offset = (addr>>7)&7" in seen and ok["halo, base offset = (addr>>7)&7"]: print("18=1")
else: print("none")
PY
)
      echo "halo mode: $HK" >> gpurun_out/rc.txt
      if [ "$HK" != "none" ]; then
        timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline --knob $HK > gpurun_out/bench_halo_on.log 2>&1; echo "halo on rc=$?" >> gpurun_out/rc.txt
        timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline --knob 17=1 > gpurun_out/bench_halo_off.log 2>&1; echo "halo off rc=$?" >> gpurun_out/rc.txt
      fi;;
    attnlist)
      timeout 300 $CMD > gpurun_out/plain_attn.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"attn_|bgemm" -c 200 --csv --log-file gpurun_out/attn_launches.csv $CMD > gpurun_out/ncu_attnlist.log 2>&1
      echo "attnlist rc=$?" >> gpurun_out/rc.txt;;
    attnsrc)   # source-level stall samples of the fused attention kernel: one backward launch at T = 256 and one at T = 64
      timeout 300 $CMD > gpurun_out/plain_attnsrc.log 2>&1 || { echo "plain rc=$?" >> gpurun_out/rc.txt; continue; }
      for spec in "b256 12" "b64 15" "f256 0"; do
        set -- $spec
        timeout 600 ncu --set full --import-source on --clock-control none -k regex:attn_fused -s $2 -c 1 -f -o gpurun_out/attnsrc_$1 $CMD > gpurun_out/ncu_attnsrc_$1.log 2>&1
        echo "attnsrc $1 rc=$?" >> gpurun_out/rc.txt
        ncu -i gpurun_out/attnsrc_$1.ncu-rep --page source --csv > gpurun_out/attnsrc_$1_source.csv 2>/dev/null
        ncu -i gpurun_out/attnsrc_$1.ncu-rep --page raw --csv > gpurun_out/attnsrc_$1_raw.csv 2>/dev/null
        rm -f gpurun_out/attnsrc_$1.ncu-rep
      done;;
    list_c3)   # launch list of the 28x28 workload (eager launches: --graph off), one 256-image step
      C3CMD="python bench.py --workload c3 --graph off --steps 1 --warmup 1 --no-cpu-baseline --no-gpu-baseline --no-roofline --no-parity-check"
      timeout 300 $C3CMD > gpurun_out/plain_list_c3.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/launches_c3.csv $C3CMD > gpurun_out/ncu_list_c3.log 2>&1
      echo "list_c3 rc=$?" >> gpurun_out/rc.txt;;
    list)
      timeout 300 $CMD > gpurun_out/plain_list.log 2>&1 &&
      timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 14000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
      echo "list rc=$?" >> gpurun_out/rc.txt;;
    full)
      timeout 300 $CMD > gpurun_out/plain_full.log 2>&1 || { echo "plain rc=$?" >> gpurun_out/rc.txt; continue; }
      cap() { name=$1; shift; timeout 900 ncu --set full --clock-control none "$@" -f -o gpurun_out/prof_$name $CMD > gpurun_out/ncu_$name.log 2>&1; echo "$name rc=$?" >> gpurun_out/rc.txt; rep2csv; }
      cap gf -k regex:conv_gather_umma -s 2 -c 16       # forward gather launches (fprop incl. the GN-statistics epilogue)
      cap gb -k regex:conv_gather_umma -s 150 -c 14     # backward gather launches (dgrad incl. the GN-backward epilogue)
      cap wg -k regex:conv_reduce_umma -s 2 -c 10
      cap gn -k regex:"gn_apply_pre" -s 4 -c 6
      cap gnb -k regex:"gn_bwd_pre" -s 4 -c 6
      cap at -k regex:"bgemm_umma|attn_" -s 0 -c 36
      cap opt -k regex:"adamw_chunked|grad_sqnorm_chunked" -s 0 -c 2
      cap loss -k regex:"reparam_kl|recon_loss|kl_bwd|scale_kernel" -s 0 -c 5
      cap misc -k regex:"pack_weights|wgrad_finalize|colsum_partial" -s 20 -c 8
      ;;
  esac
done
cat gpurun_out/rc.txt
tail -n 15 gpurun_out/t_gpu.log gpurun_out/smoke.log 2>/dev/null
tail -n 3 gpurun_out/bench*.log gpurun_out/torch_baseline*.log 2>/dev/null | cut -c1-1500
