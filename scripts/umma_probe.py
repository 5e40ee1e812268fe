"""Bring-up probe for the tcgen05 conv kernels: per case and per direction, error against the
SIMT kernels (same operands, same device) and a structural hint on failure.  Each (direction,
variant) runs in its own process so a trapped kernel does not poison the next probe."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [(1, 8, 8, 64, 64, 1, 0), (2, 8, 8, 64, 64, 1, 0), (2, 8, 8, 128, 128, 1, 0), (2, 8, 8, 64, 64, 3, 0),
         (2, 16, 16, 128, 256, 3, 0), (2, 64, 64, 64, 64, 3, 0), (3, 7, 7, 64, 64, 3, 0), (2, 14, 14, 64, 128, 3, 0),
         (2, 16, 16, 64, 64, 3, 1), (2, 8, 8, 64, 64, 3, 2), (2, 28, 28, 64, 64, 3, 1), (3, 7, 7, 64, 64, 3, 2),
         (8, 32, 32, 128, 128, 3, 0), (2, 8, 8, 256, 512, 1, 0)]


def child(direction, variant):
    import torch
    from medvae_disentangled_multimodal_b200 import _lib as L, ops
    L.lib().mvae_set_debug(1, int(variant))
    dev = "cuda"
    for case in CASES:
        B, H, W, Cin, Cout, k, mode = case
        if direction == "wgrad" and (Cin % 128 or Cout % 128):
            continue
        g = torch.Generator().manual_seed(0)
        ds = ops.conv_desc(B, H, W, Cin, Cout, k, mode, L.BF16, L.ENGINE_SIMT)
        du = ops.conv_desc(B, H, W, Cin, Cout, k, mode, L.BF16, L.ENGINE_UMMA)
        oh, ow = ops.conv_out_hw(ds)
        x = torch.randn(B, H, W, Cin, generator=g).to(dev, torch.bfloat16)
        w = (torch.randn(Cout, Cin, k, k, generator=g) / (Cin * k * k) ** 0.5).to(dev)
        b = torch.randn(Cout, generator=g).to(dev)
        dy = torch.randn(B, oh, ow, Cout, generator=g).to(dev, torch.bfloat16)
        ns = ops.conv_num_slabs(ds)
        wf = torch.empty(ns, Cout, Cin, dtype=torch.bfloat16, device=dev)
        wd = torch.empty(ns, Cin, Cout, dtype=torch.bfloat16, device=dev)
        ops.conv_pack_weights(ds, w, wf, wd, Cout, 0)

        def err(a, r):
            a, r = a.double(), r.double()
            return float((a - r).norm() / (r.norm() + 1e-30))
        if direction == "fprop":
            ref = ops.conv_fprop(ds, x, wf, b, None)
            got = ops.conv_fprop(du, x, wf, b, None)
        elif direction == "dgrad":
            ref = ops.conv_dgrad(ds, dy, wd)
            got = ops.conv_dgrad(du, dy, wd)
        else:
            ref, _ = ops.conv_wgrad(ds, x, dy, False)
            got, _ = ops.conv_wgrad(du, x, dy, False)
        torch.cuda.synchronize()
        e = err(got, ref)
        msg = f"{direction} v{variant} {case}: rel err {e:.3e} {'OK' if e < 1e-2 else 'MISMATCH'}"
        if e >= 1e-2:
            gf, rf = got.float().flatten(), ref.float().flatten()
            msg += (f" | got |max| {float(gf.abs().max()):.3g} ref |max| {float(rf.abs().max()):.3g}"
                    f" nan {int(torch.isnan(gf).sum())} zero-frac {float((gf == 0).float().mean()):.3f}"
                    f" cos {float((gf.double() @ rf.double()) / (gf.double().norm() * rf.double().norm() + 1e-30)):.4f}")
        print(msg, flush=True)


if __name__ == "__main__":
    if len(sys.argv) >= 3:
        child(sys.argv[1], sys.argv[2])
        sys.exit(0)
    for direction, variants in (("fprop", [0]), ("dgrad", [0]), ("wgrad", [0, 1])):
        for v in variants:
            try:
                r = subprocess.run([sys.executable, __file__, direction, str(v)], capture_output=True, text=True,
                                   timeout=120)
                print(r.stdout[-6000:])
                if r.returncode != 0:
                    print(f"[{direction} v{v}] exit {r.returncode}\n{r.stderr[-3000:]}")
            except subprocess.TimeoutExpired as e:
                print(f"[{direction} v{v}] TIMEOUT\n{(e.stdout or b'')[-3000:]}")
