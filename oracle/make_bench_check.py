"""Generate tests/golden/bench_first_step.json by running the UNMODIFIED reference (oracle/_ref copy) in fp32 on the CPU:

    python oracle/make_bench_check.py [c1 c2 c3 c4 c5]

For every bench.py workload: the model is built as bench.py builds it (reference default init under torch.manual_seed(42)),
fed the first N images of bench.py's synthetic batch and a seeded reparameterisation noise (torch.Generator(5), injected by
patching torch.randn_like while the reference's forward runs), and the loss terms, the global gradient norm and the norm of
every parameter gradient are recorded.  bench.py replays exactly this on the GPU before it times anything and refuses to
print a number when the first-step loss / gradient norm disagree; tests/test_configs_*.py use the per-parameter norms.
TEST INFRASTRUCTURE (build container only: needs /root/reference through oracle/build_ref.py).
"""
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

N_CHECK = {"c1": 4, "c2": 4, "c3": 4, "c4": 2, "c5": 6}
OUT = os.path.join(ROOT, "tests", "golden", "bench_first_step.json")


def check_inputs(wl: str):
    """-> (x, cond, eps) host tensors: the first N_CHECK images of bench.py's batch + seeded noise"""
    _cls, kw, *_ = bench.WORKLOADS[wl]
    _ch, zc, _dis = bench.wl_shape(wl)
    n = N_CHECK[wl]
    x, cond = bench.wl_batch(wl, n)
    r = kw["resolution"] // 2 ** (len(kw["ch_mult"]) - 1)
    eps = torch.randn(n, zc, r, r, generator=torch.Generator().manual_seed(5))
    return x, cond, eps


class InjectEps:
    """torch.randn_like -> the supplied eps while the reference's reparameterize runs (base_vae.py:86)"""

    def __init__(self, eps):
        self.eps = eps

    def __enter__(self):
        self._orig = torch.randn_like
        torch.randn_like = lambda t, *a, **k: self.eps.to(t.device, t.dtype)
        return self

    def __exit__(self, *exc):
        torch.randn_like = self._orig


def run(wl: str):
    cls, kw, lkw, *_ = bench.WORKLOADS[wl]
    x, cond, eps = check_inputs(wl)
    torch.set_num_threads(os.cpu_count() or 1)
    t0 = time.time()
    tr = ref_runner.ReferenceModuleTrainer(cls, kw, lkw, None, "cpu", **bench.OPT)     # seeds 42 before construction
    with InjectEps(eps):
        _out, ld = tr.forward_loss(x, cond)
    ld["loss"].backward()
    norms = {n: float(p.grad.double().norm()) if p.grad is not None else 0.0 for n, p in tr.model.named_parameters()}
    total = sum(v * v for v in norms.values()) ** 0.5
    res = {"images": N_CHECK[wl], "eps_seed": 5, "weights_seed": 42,
           "loss": {k: float(v.detach()) for k, v in ld.items()},
           "grad_norm": total, "param_grad_norms": norms, "n_params": sum(p.numel() for p in tr.model.parameters())}
    print(f"{wl}: loss {res['loss']['loss']:.6f} grad norm {total:.6f} ({time.time() - t0:.1f} s)")
    return res


def main():
    wls = sys.argv[1:] or list(N_CHECK)
    data = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for wl in wls:
        data[wl] = run(wl)
    data["_how"] = "oracle/make_bench_check.py: unmodified reference (oracle/_ref), fp32 CPU, torch " + torch.__version__
    with open(OUT, "w") as f:
        json.dump(data, f, indent=1)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
