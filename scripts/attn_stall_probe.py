"""GPU probe: CUDA-event times of the fused attention kernel at the benchmarked shapes and where its MMA-issuing thread
waits (knob 19)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from medvae_disentangled_multimodal_b200 import _lib as L, ops

def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for (B, H, C) in [(1024, 16, 512), (1024, 8, 1024), (256, 14, 256), (256, 7, 512)]:
    qkv = (torch.randn(B, H, H, 3 * C, device="cuda") * 0.5).bfloat16()
    dout = torch.randn(B, H, H, C, device="cuda").bfloat16()
    out, saved = ops.attn_fwd(qkv)
    f = lambda: ops.attn_fwd(qkv)
    b = lambda: ops.attn_bwd(qkv, saved, dout)
    T = H * H
    print(f"B {B} T {T} C {C}: fwd {timeit(f):.3f} ms  bwd {timeit(b):.3f} ms  (20 in a row: fwd {timeit(f, 20):.3f} bwd {timeit(b, 20):.3f})", flush=True)
    L.lib().mvae_set_debug(19, 1)
    f(); torch.cuda.synchronize(); b(); torch.cuda.synchronize()
    L.lib().mvae_set_debug(19, 0)
