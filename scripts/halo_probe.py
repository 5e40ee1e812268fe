"""GPU probe for the halo variant of the 128-channel gather kernel: relative error of fprop / dgrad against fp64
torch for both settings of the descriptor base-offset field (knob 18) and with the variant off (knob 17 = 1)."""
import math, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from medvae_disentangled_multimodal_b200 import _lib as L, ops

DEV = "cuda"
def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))

def run(case, seed=0):
    B, H, W, Cin, Cout = case
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, Cin, H, W, generator=g).bfloat16()
    w = (torch.randn(Cout, Cin, 3, 3, generator=g) / math.sqrt(Cin * 9))
    b = torch.randn(Cout, generator=g)
    dy = torch.randn(B, Cout, H, W, generator=g).bfloat16()
    wb = w.bfloat16()
    xd, wd_, dyd = x.to(DEV).double(), wb.to(DEV).double(), dy.to(DEV).double()
    y_ref = F.conv2d(xd, wd_, b.to(DEV).double(), padding=1)
    dx_ref = F.conv_transpose2d(dyd, wd_, padding=1)
    d = ops.conv_desc(B, H, W, Cin, Cout, 3, 0, L.BF16, L.ENGINE_UMMA)
    ns = ops.conv_num_slabs(d)
    wf = torch.empty(ns, Cout, Cin, dtype=torch.bfloat16, device=DEV)
    wdd = torch.empty(ns, Cin, Cout, dtype=torch.bfloat16, device=DEV)
    ops.conv_pack_weights(d, w.to(DEV), wf, wdd, Cout, 0)
    xg = x.to(DEV).permute(0, 2, 3, 1).contiguous()
    dyg = dy.to(DEV).permute(0, 2, 3, 1).contiguous()
    y = ops.conv_fprop(d, xg, wf, b.to(DEV), None)
    dx = ops.conv_dgrad(d, dyg, wdd)
    torch.cuda.synchronize()
    return rel(y.permute(0, 3, 1, 2), y_ref), rel(dx.permute(0, 3, 1, 2), dx_ref)

cases = [(12, 64, 64, 128, 128), (80, 32, 32, 128, 128), (300, 16, 16, 128, 128), (40, 32, 32, 256, 128),
         (40, 32, 32, 64, 128), (75, 32, 48, 128, 128)]
for name, k17, k18 in [("halo off", 1, 0), ("halo, base offset 0", 0, 0), ("halo, base offset = (addr>>7)&7", 0, 1)]:
    L.lib().mvae_set_debug(17, k17); L.lib().mvae_set_debug(18, k18)
    for c in cases:
        try:
            e = run(c)
            print(f"{name:34s} {str(c):28s} fprop {e[0]:.3e} dgrad {e[1]:.3e}", flush=True)
        except Exception as ex:   # noqa
            print(f"{name:34s} {str(c):28s} FAILED {type(ex).__name__}: {str(ex)[:200]}", flush=True)
            break
L.lib().mvae_set_debug(17, 0); L.lib().mvae_set_debug(18, 0)
