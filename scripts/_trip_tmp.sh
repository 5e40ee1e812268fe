mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_appendix_b_gpu.py -q -m gpu -k "conv" --timeout 300 -rf > gpurun_out/t_conv.log 2>&1; echo "conv rc=$?" >> gpurun_out/rc.txt
timeout 300 python scripts/gather_stall_probe.py > gpurun_out/gather_stall_probe5.txt 2>&1
