mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 1200 python -m pytest tests/test_kernels_gpu.py tests/test_appendix_b_gpu.py -q -m gpu --timeout 300 -rf > gpurun_out/t_kern.log 2>&1; echo "kern rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_attn2.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-baseline > gpurun_out/bench_attn2_c3.log 2>&1; echo "bench c3 rc=$?" >> gpurun_out/rc.txt
