#!/usr/bin/env python
"""Per-layer timing of the implicit-GEMM conv kernels at the C4 shapes (3x3, stride 1): fprop plain / with the
GroupNorm-statistics epilogue, dgrad plain / with the GroupNorm-backward epilogue on 8 and on 16 epilogue warps (knob 9).  CUDA events, 10 launches each
after 3 warm-up launches; prints TFLOP/s so the epilogue variants can be compared with the plain kernels layer by
layer.  usage: python scripts/gpu_layer_probe.py [batch] [shape indices, e.g. 0,9]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from medvae_disentangled_multimodal_b200 import _lib as L  # noqa: E402
from medvae_disentangled_multimodal_b200 import ops  # noqa: E402

dev = "cuda"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
G = 32
SHAPES = [(64, 128, 128), (32, 128, 256), (32, 256, 256), (16, 256, 512), (16, 512, 512), (8, 512, 1024), (8, 1024, 1024),
          (16, 1024, 512), (32, 512, 256), (64, 256, 128)]


if len(sys.argv) > 2:
    SHAPES = [SHAPES[int(i)] for i in sys.argv[2].split(",")]


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print(f"batch {B}: ms (TFLOP/s) per launch")
print(f"{'HxW':>5} {'Cin':>5} {'Cout':>5} | {'fprop':>14} {'fprop+stats':>14} | {'dgrad':>14} {'+gnbwd 8 warps':>14} {'+gnbwd 16 warps':>14}")


def knob9(v, fn):
    def run():
        L.lib().mvae_set_debug(9, v)
        fn()
        L.lib().mvae_set_debug(9, 0)
    return run


for res, cin, cout in SHAPES:
    d = ops.conv_desc(B, res, res, cin, cout, 3, 0, L.BF16)
    x = torch.randn(B, res, res, cin, device=dev).bfloat16()
    dy = torch.randn(B, res, res, cout, device=dev).bfloat16()
    w = torch.randn(cout, cin, 3, 3, device=dev) / (3 * cin ** 0.5)
    wf = torch.empty(9, cout, cin, dtype=torch.bfloat16, device=dev)
    wd = torch.empty(9, cin, cout, dtype=torch.bfloat16, device=dev)
    ops.conv_pack_weights(d, w, wf, wd, cout, 0)
    b = torch.zeros(cout, device=dev)
    gam, bet = torch.ones(cin, device=dev), torch.zeros(cin, device=dev)
    _a, stats = ops.gn_fwd(x, gam, bet, G, 1e-6, True)
    fl = ops.conv_flops(d)
    t = [timed(lambda: ops.conv_fprop(d, x, wf, b, None)),
         timed(lambda: ops.conv_fprop_gn(d, x, wf, b, None, G)),
         timed(lambda: ops.conv_dgrad(d, dy, wd)),
         timed(knob9(1, lambda: ops.conv_dgrad_gn(d, dy, wd, x, stats, gam, bet, G, True))),
         timed(knob9(2, lambda: ops.conv_dgrad_gn(d, dy, wd, x, stats, gam, bet, G, True)))]
    print(f"{res:>5} {cin:>5} {cout:>5} | " + " ".join(f"{m:6.3f} ({fl / m / 1e9:5.0f})" for m in t[:2]) + " | " +
          " ".join(f"{m:6.3f} ({fl / m / 1e9:5.0f})" for m in t[2:]), flush=True)
    del x, dy, w, wf, wd
