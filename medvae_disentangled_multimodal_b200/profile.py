"""Per-launch device timing of the C-ABI kernels (CUDA events on the launching stream), used by
bench.py's roofline pass.  Disabled (zero overhead beyond one ``is None`` test) unless a
``KernelTimer`` is installed with ``install()``."""
from __future__ import annotations

from collections import defaultdict
from typing import Dict, Optional

import torch

_ACTIVE: Optional["KernelTimer"] = None


class KernelTimer:
    def __init__(self):
        self.records = []   # (kind, flops, bytes, ev0, ev1)

    def add(self, kind, flops, nbytes, ev0, ev1):
        self.records.append((kind, flops, nbytes, ev0, ev1))

    def summary(self) -> Dict[str, dict]:
        torch.cuda.synchronize()
        out = defaultdict(lambda: {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
        for kind, flops, nbytes, e0, e1 in self.records:
            d = out[kind]
            d["launches"] += 1
            d["ms"] += e0.elapsed_time(e1)
            d["flops"] += flops
            d["bytes"] += nbytes
        return dict(out)


def install(t: Optional[KernelTimer]):
    global _ACTIVE
    _ACTIVE = t


def active() -> Optional[KernelTimer]:
    return _ACTIVE


class region:
    """with region("conv_fprop", flops=...): <launch>   — no-op unless a timer is installed"""
    __slots__ = ("kind", "flops", "nbytes", "e0")

    def __init__(self, kind: str, flops: float = 0.0, nbytes: float = 0.0):
        self.kind, self.flops, self.nbytes = kind, flops, nbytes
        self.e0 = None

    def __enter__(self):
        if _ACTIVE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if self.e0 is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            _ACTIVE.add(self.kind, self.flops, self.nbytes, self.e0, e1)
        return False
