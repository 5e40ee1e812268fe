"""The GPU baseline to beat: the UNMODIFIED reference modules (the oracle/_ref copy made
by oracle/build_ref.py) on one B200 through the stock cuDNN / cuBLAS / ATen kernels — what the reference's own
`L.Trainer(accelerator="gpu", precision=16)` path runs (main.py:86-92, configs/config.yaml:20-23) — with bench.py's step
definition (forward + VAELoss + backward + NaN scan + clip + AdamW), synthetic batch and seeded weights.  Variants: fp32
with TF32 off / on, bf16 autocast in NCHW and channels_last.

NOT the product: none of libmedvae_b200's kernels run here.  bench.py runs the two bf16 variants itself after its own
timed region (`gpu_baseline` in its JSON line); this script prints every variant, one JSON line each:

    python tests/gpu_torch_baseline.py [--workload c4] [--batch 128] [--steps 5] [--warmup 3]

(--device cpu runs the same loop on the host for a syntax check in the build container.)"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402  (workload table, synthetic batches)
from oracle import ref_runner  # noqa: E402

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
    x, cond = bench.wl_batch(wl, batch)
    cuda = device.type == "cuda"
    sampler = bench.ClockSampler(device.index or 0) if cuda else None
    if sampler:
        sampler.start()
    r = ref_runner.time_reference(cls, kw, lkw, x, cond, device=device, steps=steps, warmup=warmup, opt=bench.OPT,
                                  autocast_dtype=amp_dtype, channels_last=cl, tf32=tf32)
    line = {
        "impl": "reference-on-stock-torch-kernels", "variant": name, "metric": "train images/sec",
        "value": round(r["images_per_s"], 2), "unit": "images/s", "n_gpus": 1, "steps": r["steps"], "warmup": warmup,
        "ms_per_step": round(r["ms_per_step"], 3),
        "device": torch.cuda.get_device_name(device) if cuda else "cpu",
        "config": {"workload": wl, "class": cls, "batch": batch, "full_batch_per_gpu": bpg,
                   "gflop_per_image_fwd_bwd": gflop},
        "step_tflops": round(r["images_per_s"] * gflop * 1e-3, 2),
        "hbm_peak_allocated_gb": r["hbm_peak_allocated_gb"], "loss": r["loss"],
        "clocks": sampler.stop() if sampler else None,
    }
    print(json.dumps(line), flush=True)
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c4", choices=list(bench.WORKLOADS))
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--variants", default=",".join(VARIANTS))
    ap.add_argument("--device", default="cuda:0")
    a = ap.parse_args()
    dev = torch.device(a.device)
    batch = a.batch or bench.GPU_BASE_BATCH[a.workload]
    for name in a.variants.split(","):
        try:
            run_variant(a.workload, name, batch, a.steps, a.warmup, dev)
        except torch.OutOfMemoryError as e:     # a variant that does not fit is reported, the others still run
            print(json.dumps({"impl": "reference-on-stock-torch-kernels", "variant": name,
                              "error": f"out of memory at batch {batch}: {str(e).splitlines()[0]}"}), flush=True)
            if dev.type == "cuda":
                torch.cuda.empty_cache()
        except RuntimeError as e:               # the disentangled reference cannot run channels_last (z.view)
            print(json.dumps({"impl": "reference-on-stock-torch-kernels", "variant": name,
                              "error": f"the reference fails in this variant: {str(e).splitlines()[0][:160]}"}), flush=True)
            if dev.type == "cuda":
                torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
