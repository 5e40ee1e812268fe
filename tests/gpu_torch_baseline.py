"""The GPU baseline to beat: the reference's own arithmetic — the
oracle port, i.e. the same torch leaf ops the reference's modules call (F.conv2d, F.group_norm, torch.bmm,
F.softmax, F.interpolate, ...) — on one B200 through stock cuDNN / cuBLAS / ATen kernels, same step
definition as bench.py (forward + VAELoss + backward + NaN scan + clip + AdamW), same synthetic batch and
seeded weights.  Variants: fp32 with TF32 off / on, bf16 autocast, NCHW and channels_last.

NOT the product and not part of bench.py's arms: none of libmedvae_b200's kernels run here.  One JSON line
per variant:

    python tests/gpu_torch_baseline.py [--workload c4] [--batch 128] [--steps 5] [--warmup 2]

(--device cpu runs the same loop on the host for a syntax check in the build container.)"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402  (workload table, synthetic batches, seeded state_dict)
from oracle import vae_oracle as O  # noqa: E402

DEFAULT_BATCH = {"c1": 64, "c2": 512, "c3": 256, "c4": 128, "c5": 32}
VARIANTS = {
    # name: (autocast dtype or None, allow TF32, channels_last)
    "fp32": (None, False, False),
    "tf32": (None, True, False),
    "bf16_autocast": (torch.bfloat16, True, False),
    "bf16_autocast_channels_last": (torch.bfloat16, True, True),
}


def run_variant(wl: str, name: str, batch: int, steps: int, warmup: int, device: torch.device):
    amp_dtype, tf32, cl = VARIANTS[name]
    cls, kw, lkw, bpg, _conditional, gflop, _scaling = bench.WORKLOADS[wl]
    _ch, zc, dis = bench.wl_shape(wl)
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.allow_tf32 = tf32
    torch.backends.cudnn.benchmark = True
    _model, sd = bench.build_state_dict(wl)
    fmt = (lambda t: t.to(memory_format=torch.channels_last) if (cl and t.dim() == 4) else t)
    sd = {k: fmt(v.to(device)) for k, v in sd.items()}
    tr = O.ReferenceTrainer(sd, lr=bench.OPT["lr"], weight_decay=bench.OPT["weight_decay"], betas=bench.OPT["betas"],
                            max_norm=bench.OPT["max_norm"], loss_kwargs=lkw,
                            disentangled=(kw["shared_latent_dim"], kw["modality_latent_dim"]) if dis else None)
    x, cond = bench.wl_batch(wl, batch)
    x = fmt(x.to(device))
    cond = cond.to(device) if cond is not None else None
    r = kw["resolution"] // 2 ** (len(kw["ch_mult"]) - 1)
    eps = torch.randn(batch, zc, r, r, generator=torch.Generator().manual_seed(7)).to(device)
    cuda = device.type == "cuda"
    amp = torch.autocast(device.type, dtype=amp_dtype) if amp_dtype is not None else contextlib.nullcontext()

    def one():
        with amp:
            return tr.step(x, eps, cond)      # returns python floats: one device->host read per step, as bench.py's e2e

    for _ in range(warmup):
        one()
    if cuda:
        torch.cuda.synchronize()
        torch.cuda.reset_peak_memory_stats()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    loss = None
    for _ in range(steps):
        loss = one()["loss"]
    if cuda:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = 1e3 * (time.perf_counter() - t0)
    ips = batch * steps / (ms * 1e-3)
    line = {
        "impl": "torch-stock-kernels", "variant": name, "metric": "train images/sec", "value": round(ips, 2),
        "unit": "images/s", "n_gpus": 1, "steps": steps, "warmup": warmup, "ms_per_step": round(ms / steps, 3),
        "device": torch.cuda.get_device_name(device) if cuda else "cpu",
        "config": {"workload": wl, "class": cls, "batch": batch, "full_batch_per_gpu": bpg,
                   "gflop_per_image_fwd_bwd": gflop},
        "step_tflops": round(ips * gflop * 1e-3, 2),
        "hbm_peak_allocated_gb": round(torch.cuda.max_memory_allocated() / 1e9, 2) if cuda else None,
        "loss": loss,
    }
    print(json.dumps(line), flush=True)
    del tr, sd
    if cuda:
        torch.cuda.empty_cache()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c4", choices=list(bench.WORKLOADS))
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--variants", default=",".join(VARIANTS))
    ap.add_argument("--device", default="cuda:0")
    a = ap.parse_args()
    dev = torch.device(a.device)
    batch = a.batch or DEFAULT_BATCH[a.workload]
    for name in a.variants.split(","):
        try:
            run_variant(a.workload, name, batch, a.steps, a.warmup, dev)
        except torch.OutOfMemoryError as e:     # a variant that does not fit is reported, the others still run
            print(json.dumps({"impl": "torch-stock-kernels", "variant": name, "error": f"out of memory at batch {batch}: "
                              f"{str(e).splitlines()[0]}"}), flush=True)
            if dev.type == "cuda":
                torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
