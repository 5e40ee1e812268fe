mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_appendix_b_gpu.py -q -m gpu -k "attention" --timeout 300 -rf > gpurun_out/t_attn.log 2>&1; echo "attn rc=$?" >> gpurun_out/rc.txt
timeout 300 python scripts/attn_stall_probe.py > gpurun_out/attn_stall_probe8.txt 2>&1
