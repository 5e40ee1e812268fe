"""Learning-rate schedules for FusedAdamW — the host side of SURVEY 8(f) rank 1.

Mirrors ``get_scheduler`` of the reference (src/utils/training_utils.py:12-57): the same ``type``
strings, the same config keys and defaults, the same ``ValueError`` for an unknown type, and the
same per-epoch stepping the Lightning module configures (src/lightning_module.py:408-450:
``interval="epoch"``, ``monitor="val/loss"`` for the plateau schedule).  FusedAdamW is not a
``torch.optim.Optimizer`` (its update is one multi-tensor kernel over a pointer table), so the
torch scheduler classes cannot drive it; these closed forms set ``optimizer.lr`` instead — a host
scalar that is passed to the kernel each step, no device work.  ``tests/test_schedulers.py`` checks
every schedule against the torch class the reference instantiates.
"""
from __future__ import annotations

import math
from bisect import bisect_right
from typing import Any, Dict, Optional, Sequence


class _Schedule:
    """epoch-indexed schedule: ``step()`` advances one epoch and writes ``optimizer.lr``."""

    def __init__(self, optimizer):
        self.optimizer = optimizer
        self.base_lr = float(optimizer.lr)
        self.last_epoch = 0

    def lr_at(self, epoch: int) -> float:
        raise NotImplementedError

    def step(self, metric: Optional[float] = None) -> float:
        self.last_epoch += 1
        self.optimizer.lr = self.lr_at(self.last_epoch)
        return self.optimizer.lr

    def get_last_lr(self):
        return [self.optimizer.lr]

    def state_dict(self) -> Dict[str, Any]:
        return {k: v for k, v in self.__dict__.items() if k != "optimizer"}

    def load_state_dict(self, sd: Dict[str, Any]):
        self.__dict__.update(sd)
        self.optimizer.lr = self.lr_at(self.last_epoch)


class StepLR(_Schedule):
    """lr = base * gamma^(epoch // step_size)   (torch.optim.lr_scheduler.StepLR)"""

    def __init__(self, optimizer, step_size: int = 30, gamma: float = 0.1):
        super().__init__(optimizer)
        self.step_size, self.gamma = int(step_size), float(gamma)

    def lr_at(self, epoch):
        return self.base_lr * self.gamma ** (epoch // self.step_size)


class MultiStepLR(_Schedule):
    """lr = base * gamma^(number of milestones <= epoch)"""

    def __init__(self, optimizer, milestones: Sequence[int] = (50, 100), gamma: float = 0.1):
        super().__init__(optimizer)
        self.milestones, self.gamma = sorted(int(m) for m in milestones), float(gamma)

    def lr_at(self, epoch):
        return self.base_lr * self.gamma ** bisect_right(self.milestones, epoch)


class ExponentialLR(_Schedule):
    """lr = base * gamma^epoch"""

    def __init__(self, optimizer, gamma: float = 0.95):
        super().__init__(optimizer)
        self.gamma = float(gamma)

    def lr_at(self, epoch):
        return self.base_lr * self.gamma ** epoch


class CosineAnnealingLR(_Schedule):
    """lr = eta_min + (base - eta_min) * (1 + cos(pi * epoch / T_max)) / 2 (periodic beyond T_max,
    like torch's recursive form)"""

    def __init__(self, optimizer, T_max: int = 100, eta_min: float = 0.0):
        super().__init__(optimizer)
        self.T_max, self.eta_min = int(T_max), float(eta_min)

    def lr_at(self, epoch):
        return self.eta_min + (self.base_lr - self.eta_min) * (1.0 + math.cos(math.pi * epoch / self.T_max)) / 2.0


class ReduceLROnPlateau(_Schedule):
    """torch.optim.lr_scheduler.ReduceLROnPlateau with threshold_mode="rel", cooldown 0, min_lr 0,
    eps 1e-8 (the arguments the reference leaves at their defaults)."""

    def __init__(self, optimizer, mode: str = "min", factor: float = 0.5, patience: int = 10, threshold: float = 1e-4):
        super().__init__(optimizer)
        if mode not in ("min", "max"):
            raise ValueError(f"mode {mode} is unknown!")
        if factor >= 1.0:
            raise ValueError("Factor should be < 1.0.")
        self.mode, self.factor, self.patience, self.threshold = mode, float(factor), int(patience), float(threshold)
        self.best = math.inf if mode == "min" else -math.inf
        self.num_bad_epochs = 0

    def _better(self, a: float) -> bool:
        if self.mode == "min":
            return a < self.best * (1.0 - self.threshold)
        return a > self.best * (1.0 + self.threshold)

    def lr_at(self, epoch):
        return self.optimizer.lr

    def step(self, metric: Optional[float] = None) -> float:
        if metric is None:
            raise TypeError("ReduceLROnPlateau.step() needs the monitored metric (val/loss)")
        metric = float(metric)
        self.last_epoch += 1
        if self._better(metric):
            self.best, self.num_bad_epochs = metric, 0
        else:
            self.num_bad_epochs += 1
        if self.num_bad_epochs > self.patience:
            new_lr = self.optimizer.lr * self.factor
            if self.optimizer.lr - new_lr > 1e-8:
                self.optimizer.lr = new_lr
            self.num_bad_epochs = 0
        return self.optimizer.lr


def get_scheduler(optimizer, scheduler_config: Dict[str, Any]):
    """src/utils/training_utils.py:12-57 — same ``type`` strings, keys and defaults."""
    c = scheduler_config
    t = c.get("type", "none")
    if t == "none":
        return None
    if t == "step":
        return StepLR(optimizer, step_size=c.get("step_size", 30), gamma=c.get("gamma", 0.1))
    if t == "multistep":
        return MultiStepLR(optimizer, milestones=c.get("milestones", [50, 100]), gamma=c.get("gamma", 0.1))
    if t == "exponential":
        return ExponentialLR(optimizer, gamma=c.get("gamma", 0.95))
    if t == "cosine":
        return CosineAnnealingLR(optimizer, T_max=c.get("T_max", 100), eta_min=c.get("eta_min", 0))
    if t == "reduce_on_plateau":
        return ReduceLROnPlateau(optimizer, mode=c.get("mode", "min"), factor=c.get("factor", 0.5),
                                 patience=c.get("patience", 10), threshold=c.get("threshold", 1e-4))
    raise ValueError(f"Unknown scheduler type: {t}")
