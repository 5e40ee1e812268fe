"""The bf16 yardstick at the REAL widths of the BASELINE configurations (build container only):

    python oracle/make_bf16_yardstick_configs.py [c1 c3 c4 c5]     # -> tests/golden/bf16_yardstick_configs.json

The UNMODIFIED reference (oracle/_ref) is run twice on the inputs of tests/test_configs_gpu.py (bench.py's seeded weights
and synthetic batch, injected noise): in fp32 and under ``torch.autocast("cpu", dtype=torch.bfloat16)`` — what Lightning's
bf16 precision does to it.  Recorded: how far the autocast run moves from the fp32 run on every tapped layer output, on
mean / logvar / z / reconstruction, on the loss and on the gradients.  A ~60-layer chain that stores every activation in
bf16 accumulates the per-layer rounding (2^-9 per element) like a random walk; these numbers say how large that is for the
reference's own kernels, and tests/test_configs_gpu.py holds the CUDA path to max(floor, small multiple of them).
TEST INFRASTRUCTURE; never imported by the product package."""
from __future__ import annotations

import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402
from oracle import ref_runner  # noqa: E402
from oracle.make_bench_check import InjectEps  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "bf16_yardstick_configs.json")
BATCH = {"c1": 3, "c2": 3, "c3": 4, "c4": 2, "c5": 2}      # = tests/test_configs_gpu.py BATCH


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def run_once(wl, amp):
    cls, kw, lkw, *_ = bench.WORKLOADS[wl]
    _ch, zc, dis = bench.wl_shape(wl)
    n = BATCH[wl]
    x, cond = bench.wl_batch(wl, n)
    r = kw["resolution"] // 2 ** (len(kw["ch_mult"]) - 1)
    eps = torch.randn(n, zc, r, r, generator=torch.Generator().manual_seed(5))
    tr = ref_runner.ReferenceModuleTrainer(cls, kw, lkw, None, "cpu", autocast_dtype=torch.bfloat16 if amp else None,
                                           **bench.OPT)
    taps = {}
    hooks = []
    for name, mod in tr.model.named_modules():
        if (name.startswith("encoder") or name.startswith("decoder")) and type(mod).__name__ in (
                "ResnetBlock", "AttnBlock", "Downsample", "Upsample") or name.endswith(("conv_in", "conv_out")):
            hooks.append(mod.register_forward_hook(lambda m, i, o, name=name: taps.__setitem__(name, o.detach().float())))
    with InjectEps(eps):
        out, ld = tr.forward_loss(x, cond)
    ld["loss"].float().backward()
    for h in hooks:
        h.remove()
    grads = {k: (p.grad.detach().float() if p.grad is not None else torch.zeros_like(p)) for k, p in tr.model.named_parameters()}
    outs = {k: out[k].detach().float() for k in ("reconstruction", "mean", "logvar", "z")}
    return taps, outs, {k: float(v.detach()) for k, v in ld.items()}, grads


def main():
    wls = sys.argv[1:] or ["c1", "c3", "c4", "c5"]
    data = json.load(open(OUT)) if os.path.exists(OUT) else {"configs": {}}
    torch.set_num_threads(os.cpu_count() or 1)
    for wl in wls:
        t0 = time.time()
        t32, o32, l32, g32 = run_once(wl, False)
        try:
            t16, o16, l16, g16 = run_once(wl, True)
        except NotImplementedError as e:
            # the disentangled reference cannot run under CPU bf16 autocast at all (torch.pdist has no bfloat16 CPU
            # kernel): its trunk is the C4 trunk one level up, so C4's numbers stand in
            data["configs"][wl] = dict(data["configs"]["c4"], batch=BATCH[wl],
                                       proxy=f"c4 (the reference fails under CPU bf16 autocast: {str(e)[:80]})")
            print(wl, "-> proxy c4:", e)
            continue
        layer = {k: rel(t16[k], t32[k]) for k in t32}
        num = sum(float((g16[k].double() - g32[k].double()).pow(2).sum()) for k in g32)
        den = sum(float(g32[k].double().pow(2).sum()) for k in g32)
        total = den ** 0.5
        worst = max((rel(g16[k], g32[k]) for k in g32 if float(g32[k].norm()) > 1e-3 * total), default=0.0)
        data["configs"][wl] = {
            "batch": BATCH[wl], "layers": len(layer), "worst_layer": max(layer.values()),
            "worst_layer_name": max(layer, key=layer.get),
            **{k: rel(o16[k], o32[k]) for k in o32},
            "loss_rel": max(abs(l16[k] - l32[k]) / max(1.0, abs(l32[k])) for k in l32),
            "whole_gradient": (num / den) ** 0.5, "worst_tensor": worst,
        }
        print(wl, json.dumps(data["configs"][wl]), f"({time.time() - t0:.0f} s)")
    data["how"] = ("unmodified reference (oracle/_ref) under torch.autocast('cpu', bfloat16) vs its own fp32 run, real "
                   "BASELINE widths (oracle/make_bf16_yardstick_configs.py), torch " + torch.__version__)
    with open(OUT, "w") as f:
        json.dump(data, f, indent=1)


if __name__ == "__main__":
    main()
