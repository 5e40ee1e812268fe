#!/usr/bin/env python
"""One 256->256 3x3 conv fprop at the C4 level-1 shape (B=256, 32x32), CTA-pair variant off / on (knob 8):
times with CUDA events; run under ncu for the counters.  usage: python scripts/gpu_pair_probe.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medvae_disentangled_multimodal_b200 import _lib as L  # noqa: E402
from medvae_disentangled_multimodal_b200 import ops  # noqa: E402

dev = "cuda"
B, H, W, Cin, Cout = 256, 32, 32, 256, 256
d = ops.conv_desc(B, H, W, Cin, Cout, 3, 0, L.BF16)
x = torch.randn(B, H, W, Cin, device=dev).bfloat16()
w = torch.randn(Cout, Cin, 3, 3, device=dev) / 48
wf = torch.empty(9, Cout, Cin, dtype=torch.bfloat16, device=dev)
ops.conv_pack_weights(d, w, wf, None, Cout, 0)
b = torch.zeros(Cout, device=dev)
flops = ops.conv_flops(d)
for knob in (0, 1, 0, 1):
    L.lib().mvae_set_debug(8, knob)
    for _ in range(3):
        y = ops.conv_fprop(d, x, wf, b, None)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        y = ops.conv_fprop(d, x, wf, b, None)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"pair={knob}: {ms * 1e3:.1f} us  {flops / ms / 1e9:.1f} TF/s  checksum {float(y.float().sum()):.3f}")
