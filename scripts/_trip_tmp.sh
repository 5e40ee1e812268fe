mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 1200 python -m pytest tests/test_kernels_gpu.py tests/test_appendix_b_gpu.py -q -m gpu --timeout 600 -x -rf > gpurun_out/t_kern.log 2>&1; echo "kern rc=$?" >> gpurun_out/rc.txt
timeout 300 python scripts/gather_stall_probe.py > gpurun_out/gather_stall_probe4.txt 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_elect.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_elect_c3.log 2>&1; echo "bench c3 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -5 gpurun_out/t_kern.log
