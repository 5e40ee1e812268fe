#!/usr/bin/env python
"""Which Python call sites issue the small torch kernels (fill / add / copy) inside one C4 training step?
usage (GPU box): python scripts/gpu_small_launches.py > gpurun_out/small_launches.txt"""
import collections
import os
import sys
import traceback

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import medvae_disentangled_multimodal_b200 as M  # noqa: E402
from medvae_disentangled_multimodal_b200.train import TrainStep  # noqa: E402

cls, kw, lkw, bpg, conditional, gflop, scaling = bench.WORKLOADS["c4"]
dev = torch.device("cuda")
torch.manual_seed(42)
model = M.ConditionalVAE(**kw).to(dev)
ts = TrainStep(model, M.VAELoss(**lkw), micro_batch=64, **bench.OPT)
x, cond = bench.wl_batch("c4", 64)
x, cond = x.to(dev), cond.to(dev)
for _ in range(2):
    ts.step(x, cond)
torch.cuda.synchronize()

counts = collections.Counter()
from torch.utils._python_dispatch import TorchDispatchMode


class Spy(TorchDispatchMode):
    def __torch_dispatch__(self, func, types, args=(), kwargs=None):
        name = str(func)
        if any(k in name for k in ("fill", "zero", "add", "copy", "mul", "div", "cat", "ones", "full")):
            st = [f for f in traceback.extract_stack() if "medvae_disentangled" in f.filename or "bench" in f.filename]
            site = " <- ".join(f"{os.path.basename(f.filename)}:{f.lineno}" for f in st[-3:])
            counts[(name, site)] += 1
        return func(*args, **(kwargs or {}))


with Spy():
    ts.step(x, cond)
torch.cuda.synchronize()
for (name, site), n in counts.most_common(40):
    print(f"{n:5d}  {name:40s} {site}")
