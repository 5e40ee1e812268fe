"""GPU probe: where the MMA-issuing thread of the gather kernel spends its cycles (knob 19) and CUDA-event times of the
benchmarked 128-channel layer (1024 x 64 x 64, 128 -> 128), halo tiles on / off, plain fprop and both GroupNorm epilogues."""
import math, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from medvae_disentangled_multimodal_b200 import _lib as L, ops

DEV = "cuda"
def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

def layer(B, H, W, Cin, Cout, G=32):
    g = torch.Generator().manual_seed(0)
    d = ops.conv_desc(B, H, W, Cin, Cout, 3, 0, L.BF16, L.ENGINE_UMMA)
    ns = ops.conv_num_slabs(d)
    w = torch.randn(Cout, Cin, 3, 3, generator=g) / math.sqrt(Cin * 9)
    wf = torch.empty(ns, Cout, Cin, dtype=torch.bfloat16, device=DEV)
    wd = torch.empty(ns, Cin, Cout, dtype=torch.bfloat16, device=DEV)
    ops.conv_pack_weights(d, w.to(DEV), wf, wd, Cout, 0)
    x = torch.randn(B, H, W, Cin, device=DEV).bfloat16()
    dy = torch.randn(B, H, W, Cout, device=DEV).bfloat16()
    b = torch.randn(Cout, device=DEV)
    gam, bet = torch.ones(Cin, device=DEV), torch.zeros(Cin, device=DEV)
    _a, stats = ops.gn_fwd(x, gam, bet, G, 1e-6, True)
    flop = 2.0 * B * H * W * Cin * Cout * 9
    return {
        "fprop plain": (lambda: ops.conv_fprop(d, x, wf, b, None), flop),
        "fprop + residual": (lambda: ops.conv_fprop(d, x, wf, b, dy), flop),
        "fprop + GN statistics": (lambda: ops.conv_fprop_gn(d, x, wf, b, None, G), flop),
        "dgrad plain": (lambda: ops.conv_dgrad(d, dy, wd), flop),
        "dgrad + GN backward": (lambda: ops.conv_dgrad_gn(d, dy, wd, x, stats, gam, bet, G, True), flop),
    }

for shape in [(1024, 64, 64, 128, 128), (1024, 32, 32, 256, 256)]:
    fns = layer(*shape)
    for halo in (0, 1):
        L.lib().mvae_set_debug(17, halo)
        for name, (fn, flop) in fns.items():
            L.lib().mvae_set_debug(19, 0)
            ms = timeit(fn)
            print(f"{shape} {'per-tap' if halo else 'halo   '} {name:24s} {ms:7.3f} ms  {flop / ms / 1e9:7.1f} TFLOP/s", flush=True)
            L.lib().mvae_set_debug(19, 1)
            fn(); torch.cuda.synchronize()
            L.lib().mvae_set_debug(19, 0)
L.lib().mvae_set_debug(17, 0)
