"""Drive the UNMODIFIED reference classes (the copy under ``oracle/_ref``, see build_ref.py) through the training
step the reference's Lightning module performs.  TEST INFRASTRUCTURE: checker in tests/, timed baseline in bench.py
(``cpu_baseline.kind = "reference"``, ``gpu_baseline`` = the same code on the B200 through stock cuDNN / cuBLAS /
ATen kernels).  Never imported by the product package.

What is restated here (Lightning itself is not installed): ``VAELightningModule.training_step``
(src/lightning_module.py:98-218: forward, criterion call by model type, NaN -> 0 / 1e6 guards on the logged terms),
``on_before_optimizer_step`` (:468-477: a parameter whose gradient holds any NaN/Inf has its WHOLE gradient zeroed),
``configure_gradient_clipping`` (:452-466: clip by global norm) and ``configure_optimizers`` (:390-406: AdamW / Adam).
Everything below that — the models and the loss — is the reference's own code, imported, not restated.
"""
from __future__ import annotations

import contextlib
import hashlib
import json
import os
import sys
import time
import types
from typing import Dict, Optional

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
_loaded = None


def available() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "MANIFEST.json"))


def load_reference():
    """-> (src.models module, src.losses.vae_losses module) of the unmodified reference copy."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("oracle/_ref is missing: run `python oracle/build_ref.py` in the build container "
                           "(needs /root/reference); the copy travels to the GPU box with the snapshot")
    man = json.load(open(os.path.join(REF_DIR, "MANIFEST.json")))
    for rel, digest in man["files"].items():
        with open(os.path.join(REF_DIR, rel), "rb") as f:
            if hashlib.sha256(f.read()).hexdigest() != digest:
                raise RuntimeError(f"oracle/_ref/{rel} was modified after it was copied from the reference")
    if "src" in sys.modules and not any(os.path.abspath(p).startswith(REF_DIR)
                                        for p in getattr(sys.modules["src"], "__path__", [])):
        raise RuntimeError("another package named `src` is already imported in this process (the drop-in shim?); "
                           "the reference copy must be loaded in a process of its own")
    sys.path.insert(0, REF_DIR)
    for name in ("lpips", "open_clip"):          # imported at module top of vae_losses.py, used by neither class we call
        sys.modules.setdefault(name, types.ModuleType(name))
    import importlib
    M = importlib.import_module("src.models")
    Lm = importlib.import_module("src.losses.vae_losses")
    if not os.path.abspath(M.__file__).startswith(REF_DIR):
        raise RuntimeError(f"`src.models` resolved to {M.__file__}, not to the reference copy")
    _loaded = (M, Lm)
    return _loaded


class ReferenceModuleTrainer:
    """One optimisation step of the reference model + criterion, as Lightning would run it."""

    def __init__(self, cls: str, kw: dict, loss_kw: dict, sd: Optional[Dict[str, torch.Tensor]] = None,
                 device="cpu", lr=1e-4, weight_decay=1e-4, betas=(0.9, 0.999), max_norm=1.0, decoupled=True,
                 autocast_dtype=None, channels_last=False, seed=42):
        M, Lm = load_reference()
        torch.manual_seed(seed)
        self.model = getattr(M, cls)(**kw)
        if sd is not None:
            self.model.load_state_dict(sd)
        self.device = torch.device(device)
        self.model.to(self.device)
        if channels_last:
            self.model.to(memory_format=torch.channels_last)
        self.disentangled = cls == "DisentangledConditionalVAE"
        self.conditional = cls in ("ConditionalVAE",)
        self.criterion = (M.DisentangledVAELoss if self.disentangled else Lm.VAELoss)(**loss_kw)
        opt = torch.optim.AdamW if decoupled else torch.optim.Adam
        self.opt = opt(self.model.parameters(), lr=lr, weight_decay=weight_decay, betas=betas)
        self.max_norm = max_norm
        self.autocast_dtype = autocast_dtype
        self.channels_last = channels_last
        self.last_update_s = 0.0

    def _amp(self):
        if self.autocast_dtype is None:
            return contextlib.nullcontext()
        return torch.autocast(self.device.type, dtype=self.autocast_dtype)

    def forward_loss(self, x, cond):
        """lightning_module.py:110-196"""
        with self._amp():
            if self.disentangled:
                out = self.model(x, cond)                   # cond = int64 modality indices (:116-118)
                ld = self.criterion(out, x)                 # :191-196
            else:
                out = self.model(x, cond) if self.conditional else self.model(x)
                ld = self.criterion(inputs=x, reconstructions=out["reconstruction"], posteriors=out["posterior"],
                                    priors=out["prior"])
        return out, ld

    def step(self, x, cond=None):
        self.opt.zero_grad(set_to_none=True)
        out, ld = self.forward_loss(x, cond)
        ld["loss"].backward()
        t0 = time.perf_counter()
        for p in self.model.parameters():                   # :468-477
            if p.grad is not None and (torch.isnan(p.grad).any() or torch.isinf(p.grad).any()):
                p.grad.zero_()
        if self.max_norm and self.max_norm > 0:             # :452-466
            torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.max_norm)
        self.opt.step()
        self.last_update_s = time.perf_counter() - t0
        return {k: float(v.detach()) for k, v in ld.items()}


def time_reference(cls: str, kw: dict, loss_kw: dict, x, cond, *, device, steps: int, warmup: int, opt: dict,
                   autocast_dtype=None, channels_last=False, tf32=True, budget_s: Optional[float] = None):
    """Time `steps` optimisation steps of the unmodified reference on `device` (host cores or one GPU through the stock
    cuDNN / cuBLAS / ATen kernels).  x / cond: host tensors of one batch (re-used every step, as bench.py does for its own
    arm).  -> dict(images_per_s, ms_per_step, steps, loss, update_ms_per_step, hbm_peak_allocated_gb)"""
    device = torch.device(device)
    cuda = device.type == "cuda"
    if cuda:
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cudnn.benchmark = True
    tr = ReferenceModuleTrainer(cls, kw, loss_kw, None, device, autocast_dtype=autocast_dtype,
                                channels_last=channels_last, **opt)
    xd = x.to(device)
    if channels_last and xd.dim() == 4:
        xd = xd.contiguous(memory_format=torch.channels_last)
    cd = cond.to(device) if cond is not None else None
    for _ in range(warmup):
        tr.step(xd, cd)
    if cuda:
        torch.cuda.synchronize(device)
        torch.cuda.reset_peak_memory_stats(device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    done, upd, loss = 0, 0.0, None
    for _ in range(steps):
        loss = tr.step(xd, cd)["loss"]
        done += 1
        upd += tr.last_update_s
        if budget_s is not None and time.perf_counter() - t0 >= budget_s:
            break
    if cuda:
        e1.record()
        torch.cuda.synchronize(device)
        ms = e0.elapsed_time(e1)
    else:
        ms = 1e3 * (time.perf_counter() - t0)
    res = {"images_per_s": x.shape[0] * done / (ms * 1e-3), "ms_per_step": ms / done, "steps": done, "loss": loss,
           "update_ms_per_step": 1e3 * upd / done, "total_s": ms * 1e-3,
           "hbm_peak_allocated_gb": round(torch.cuda.max_memory_allocated(device) / 1e9, 2) if cuda else None}
    del tr, xd, cd
    if cuda:
        torch.cuda.empty_cache()
    return res
