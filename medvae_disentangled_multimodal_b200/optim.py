"""Optimizer side of the training step on the fused multi-tensor kernels.

Replaces, for the hot path, what the reference's Lightning module does around ``loss.backward()``
(src/lightning_module.py:390-477): per-parameter NaN/Inf gradient zeroing
(``on_before_optimizer_step``), global-norm clipping (``configure_gradient_clipping``,
gradient_clip_val 1.0 by default) and ``torch.optim.Adam`` / ``AdamW`` (``configure_optimizers``).
Two kernel launches per step (per-tensor sums of squares + non-finite flags; clip + Adam update), no host
synchronisation.  Per-tensor semantics follow the reference: a parameter whose gradient holds ANY NaN / Inf has its whole
gradient treated as zero before the norm is taken (``param.grad.zero_()``, :468-477), and a parameter that produced no
gradient in the step (``skip``: torch.optim skips ``p.grad is None``) is left untouched — no weight decay, no moments.  The arithmetic follows torch.optim.Adam/AdamW (single-tensor path):
bias corrections 1-beta^t, denom = sqrt(v)/sqrt(bc2) + eps, step size lr/bc1; AdamW decays the
parameter by (1 - lr*wd) first, Adam adds wd*p to the gradient.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, List, Optional

import torch

from . import _lib as L
from ._lib import check, lib, ptr, stream


class FusedAdamW:
    """Adam / AdamW with fused global-norm clipping over a fixed list of fp32 parameters.

    ``step()`` consumes ``p.grad`` of every parameter (a missing grad is treated as zeros by
    skipping the tensor).  ``max_norm <= 0`` disables clipping.  State tensors are named as in
    torch.optim (``exp_avg``, ``exp_avg_sq``) so ``state_dict()`` can be handed to a torch optimizer.
    """

    CHUNK = 1 << 16

    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-4, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-4, decoupled: bool = True, max_norm: float = 1.0):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("FusedAdamW got an empty parameter list")
        for p in self.params:
            if p.dtype != torch.float32 or not p.is_cuda or not p.is_contiguous():
                raise L.MvaeError("FusedAdamW needs contiguous fp32 CUDA parameters (there is no CPU fallback)")
        self.step_state: Optional[torch.Tensor] = None
        self.lr, self.betas, self.eps = float(lr), (float(betas[0]), float(betas[1])), float(eps)
        self.weight_decay, self.decoupled, self.max_norm = float(weight_decay), bool(decoupled), float(max_norm)
        self.step_count = 0
        dev = self.params[0].device
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        self._sqnorm = torch.zeros((), dtype=torch.float64, device=dev)
        self._skip_key, self._skip_dev = None, None
        self._tables = None   # (grad data_ptrs tuple) -> device pointer tables
        self._dev = dev
        # step_state: device-resident {step, lr, Philox position} (train.TrainStep in CUDA-graph mode): when set, the
        # kernel reads step / lr from it instead of the by-value arguments

    @property
    def lr(self) -> float:
        return self._lr

    @lr.setter
    def lr(self, value: float):
        """schedulers assign this; in graph mode the captured kernel reads the device copy"""
        self._lr = float(value)
        if self.step_state is not None:
            self.step_state[1:2].fill_(self._lr)

    # ------------------------------------------------------------------ pointer tables
    def _ptr_table(self, tensors) -> torch.Tensor:
        return torch.tensor([t.data_ptr() for t in tensors], dtype=torch.int64).to(self._dev, non_blocking=False)

    def _ensure_tables(self, plist, grads):
        key = (tuple(p.data_ptr() for p in plist), tuple(g.data_ptr() for g in grads))
        if self._tables is None or self._tables["key"] != key:
            self._tables = {
                "key": key,
                "p": self._ptr_table(plist),
                "g": self._ptr_table(grads),
                "m": self._ptr_table([self._state_for(p)[0] for p in plist]),
                "v": self._ptr_table([self._state_for(p)[1] for p in plist]),
                "n": torch.tensor([p.numel() for p in plist], dtype=torch.int64).to(self._dev),
                "count": len(plist),
            }
            # chunk table: <= CHUNK elements per entry so every CTA gets the same amount of work
            ct, co, cl = [], [], []
            for i, p in enumerate(plist):
                n = p.numel()
                for off in range(0, n, self.CHUNK):
                    ct.append(i); co.append(off); cl.append(min(self.CHUNK, n - off))
            self._tables.update({
                "ct": torch.tensor(ct, dtype=torch.int32).to(self._dev),
                "co": torch.tensor(co, dtype=torch.int64).to(self._dev),
                "cl": torch.tensor(cl, dtype=torch.int32).to(self._dev),
                "chunks": len(ct),
            })
            # per-tensor pass-1 results, zeroed every step by ONE fill: [n] fp64 sums of squares | [n] int32 flags
            n = len(plist)
            st = torch.zeros(n * 12, dtype=torch.uint8, device=self._dev)
            self._tables.update({"state": st, "tsq": st[:n * 8].view(torch.float64), "tbad": st[n * 8:].view(torch.int32)})
        return self._tables

    def _state_for(self, p):
        if not hasattr(self, "_index"):
            self._index = {id(q): i for i, q in enumerate(self.params)}
        i = self._index[id(p)]
        return self.exp_avg[i], self.exp_avg_sq[i]

    # ------------------------------------------------------------------ API
    def zero_grad(self, set_to_none: bool = False):
        for p in self.params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()

    def grad_norm(self) -> torch.Tensor:
        """global L2 norm (fp64, 0-d device tensor) of the gradients that took part in the last step() (poisoned and
        skipped tensors excluded)."""
        return self._sqnorm.sqrt()

    @torch.no_grad()
    def step(self, lr: Optional[float] = None, skip=None):
        """skip: iterable of parameters that produced no gradient in this step (their ``.grad`` is a zero view of a
        flat buffer rather than None): they are left untouched, as torch.optim leaves ``p.grad is None``"""
        L.require_device()
        plist = [p for p in self.params if p.grad is not None]
        if not plist:
            return
        grads = []
        for p in plist:
            g = p.grad
            if g.dtype != torch.float32 or not g.is_contiguous():
                g = g.float().contiguous()
                p.grad = g
            grads.append(g)
        t = self._ensure_tables(plist, grads)
        self.step_count += 1
        skip_ids = frozenset(id(p) for p in skip) if skip else frozenset()
        if skip_ids != self._skip_key or (self._skip_dev is not None and self._skip_dev.numel() != len(plist)):
            self._skip_key = skip_ids
            self._skip_dev = (torch.tensor([int(id(p) in skip_ids) for p in plist], dtype=torch.int32).to(self._dev)
                              if skip_ids else None)
        t["state"].zero_()
        from . import profile as P
        nel = sum(p.numel() for p in plist)
        # 32 B / parameter: pass 1 reads g (4); the update reads p, g, m, v (16) and writes p, m, v (12)
        with P.region("adamw", 0.0, 32.0 * nel):
            self._launch(t, lr)
        # parameters changed in place behind autograd's back: bump versions so the packed bf16 operand caches
        # (blocks.ConvOp.packed) are refreshed on the next forward
        torch._C._increment_version(plist)

    def _launch(self, t, lr):
        check(lib().mvae_grad_sqnorm_chunked(ptr(t["g"]), ptr(t["ct"]), ptr(t["co"]), ptr(t["cl"]), t["chunks"],
                                             ptr(t["tsq"]), ptr(t["tbad"]), stream()), "grad_sqnorm")
        check(lib().mvae_adamw_clip_step_chunked(ptr(t["p"]), ptr(t["g"]), ptr(t["m"]), ptr(t["v"]), ptr(t["ct"]),
                                                 ptr(t["co"]), ptr(t["cl"]), t["chunks"], ptr(t["tsq"]), ptr(t["tbad"]),
                                                 ptr(self._skip_dev), t["count"], ptr(self._sqnorm),
                                                 self.max_norm, float(self.lr if lr is None else lr), self.betas[0],
                                                 self.betas[1], self.eps, self.weight_decay, int(self.decoupled),
                                                 self.step_count, ptr(self.step_state), stream()), "adamw_clip_step")

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq,
                "lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay,
                "decoupled": self.decoupled, "max_norm": self.max_norm}

    def load_state_dict(self, sd):
        """moments are copied IN PLACE (the kernel's pointer tables and a captured graph keep pointing at them)"""
        if len(sd["exp_avg"]) != len(self.exp_avg):
            raise ValueError(f"optimizer state has {len(sd['exp_avg'])} tensors, this optimizer {len(self.exp_avg)}")
        with torch.no_grad():
            for dst, src in zip(self.exp_avg, sd["exp_avg"]):
                dst.copy_(src)
            for dst, src in zip(self.exp_avg_sq, sd["exp_avg_sq"]):
                dst.copy_(src)
        self.step_count = int(sd["step"])
        self.betas, self.eps = (float(sd["betas"][0]), float(sd["betas"][1])), float(sd["eps"])
        self.weight_decay, self.decoupled, self.max_norm = float(sd["weight_decay"]), bool(sd["decoupled"]), float(sd["max_norm"])
        self.lr = float(sd["lr"])
