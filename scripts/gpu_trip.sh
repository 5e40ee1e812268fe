#!/bin/bash
# One GPU visit: each stage under its own timeout so a hung kernel cannot eat the whole lease.
# usage: scripts/gpu_trip.sh <stage>...   (stages: simt umma model bench)
mkdir -p gpurun_out
rm -f gpurun_out/rc.txt
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for stage in "$@"; do
  case $stage in
    simt)  timeout 600 python -m pytest tests/test_kernels_gpu.py -q -m "gpu and not umma" -x --timeout 300 > gpurun_out/t_simt.log 2>&1; echo "simt rc=$?" >> gpurun_out/rc.txt;;
    simt_all) timeout 900 python -m pytest tests/test_kernels_gpu.py -q -m "gpu and not umma" --timeout 300 > gpurun_out/t_simt.log 2>&1; echo "simt rc=$?" >> gpurun_out/rc.txt;;
    umma)  timeout 600 python -m pytest tests/test_kernels_gpu.py -q -m "umma" --timeout 120 > gpurun_out/t_umma.log 2>&1; echo "umma rc=$?" >> gpurun_out/rc.txt;;
    model) timeout 900 python -m pytest tests/test_model_gpu.py -q --timeout 300 > gpurun_out/t_model.log 2>&1; echo "model rc=$?" >> gpurun_out/rc.txt;;
    train) timeout 900 python -m pytest tests/test_train_gpu.py -q --timeout 600 > gpurun_out/t_train.log 2>&1; echo "train rc=$?" >> gpurun_out/rc.txt;;
    probe) timeout 300 python scripts/umma_probe.py > gpurun_out/probe.log 2>&1; echo "probe rc=$?" >> gpurun_out/rc.txt;;
    smoke) timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt;;
    bench) timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.txt;;
    bench_c3) timeout 600 python bench.py --workload c3 --steps 5 --warmup 3 > gpurun_out/bench_c3.log 2>&1; echo "bench_c3 rc=$?" >> gpurun_out/rc.txt;;
    bench_c4s) timeout 600 python bench.py --workload c4 --batch-per-gpu 128 --micro-batch 64 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4s.log 2>&1; echo "bench_c4s rc=$?" >> gpurun_out/rc.txt;;
    bench_c4m) timeout 600 python bench.py --workload c4 --batch-per-gpu 256 --micro-batch 256 --steps 2 --warmup 3 --no-cpu-baseline $BENCH_ARGS > gpurun_out/bench_c4m$TAG.log 2>&1; echo "bench_c4m rc=$?" >> gpurun_out/rc.txt;;
    kernels) timeout 900 python -m pytest tests/test_kernels_gpu.py -q -m gpu --timeout 300 > gpurun_out/t_kernels.log 2>&1; echo "kernels rc=$?" >> gpurun_out/rc.txt;;
  esac
done
cat gpurun_out/rc.txt
tail -n 30 gpurun_out/*.log
