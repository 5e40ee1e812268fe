#!/bin/bash
# Round-2 multi-GPU trip: bench lines at N GPUs (weak scaling) for C4 / C3 / C5 and, with "ab", the A/B of the SM
# reservation for the overlapped gradient all-reduce.   usage: scripts/gpu_trip_r2_multi.sh <N> [stages...]
N=$1; shift
mkdir -p gpurun_out
rm -f gpurun_out/rc_multi.txt
export PYTHONUNBUFFERED=1
STAGES=${@:-c4 c3 c5}
run() {  # name, env assignments..., -- bench args
  name=$1; shift
  envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29500 + RANDOM % 200)) bench.py --gpus $N "$@" > gpurun_out/bench_${name}_${N}gpu.log 2> gpurun_out/bench_${name}_${N}gpu.err
  echo "$name rc=$?" >> gpurun_out/rc_multi.txt
}
for stage in $STAGES; do
  case $stage in
    c4) run c4 MVAE_X=0 -- --steps 5 --warmup 3;;
    c3) run c3 MVAE_X=0 -- --workload c3 --steps 20 --warmup 5;;
    c5) run c5 MVAE_X=0 -- --workload c5 --steps 3 --warmup 3;;
    c2) run c2 MVAE_X=0 -- --workload c2 --steps 20 --warmup 5;;
    ab)  # SMs left to NCCL / NCCL's CTA count: which combination exposes the least communication
      run c4_res0 MVAE_RESERVE_SMS=0 -- --steps 4 --warmup 3 --no-roofline
      run c4_res4_cta4 MVAE_RESERVE_SMS=4 NCCL_MAX_CTAS=4 -- --steps 4 --warmup 3 --no-roofline
      run c4_res8_cta8 MVAE_RESERVE_SMS=8 NCCL_MAX_CTAS=8 -- --steps 4 --warmup 3 --no-roofline
      run c4_res0_cta4 MVAE_RESERVE_SMS=0 NCCL_MAX_CTAS=4 -- --steps 4 --warmup 3 --no-roofline
      run c3_res0 MVAE_RESERVE_SMS=0 -- --workload c3 --steps 20 --warmup 5 --no-roofline
      run c3_res4_cta4 MVAE_RESERVE_SMS=4 NCCL_MAX_CTAS=4 -- --workload c3 --steps 20 --warmup 5 --no-roofline
      ;;
    single)   # the 1-GPU lines of the same box, for the scaling ratio
      timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_c4_1gpu.log 2>&1; echo "c4 1gpu rc=$?" >> gpurun_out/rc_multi.txt
      timeout 600 python bench.py --workload c3 --steps 20 --warmup 5 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_c3_1gpu.log 2>&1; echo "c3 1gpu rc=$?" >> gpurun_out/rc_multi.txt
      timeout 600 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_c5_1gpu.log 2>&1; echo "c5 1gpu rc=$?" >> gpurun_out/rc_multi.txt
      ;;
  esac
done
cat gpurun_out/rc_multi.txt
for f in gpurun_out/bench_*gpu.log; do echo "== $f"; tail -n 1 $f | cut -c1-400; done
