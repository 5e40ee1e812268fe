"""LR schedules (host logic, CPU): every schedule of src/utils/training_utils.py:12-57 against the torch
class the reference instantiates, epoch by epoch."""
import math
import types

import pytest
import torch

from medvae_disentangled_multimodal_b200.schedulers import get_scheduler

CASES = [
    ({"type": "step", "step_size": 7, "gamma": 0.5}, lambda o: torch.optim.lr_scheduler.StepLR(o, 7, 0.5)),
    ({"type": "step"}, lambda o: torch.optim.lr_scheduler.StepLR(o, 30, 0.1)),
    ({"type": "multistep", "milestones": [5, 11, 40], "gamma": 0.3},
     lambda o: torch.optim.lr_scheduler.MultiStepLR(o, [5, 11, 40], 0.3)),
    ({"type": "exponential"}, lambda o: torch.optim.lr_scheduler.ExponentialLR(o, 0.95)),
    ({"type": "cosine", "T_max": 25, "eta_min": 1e-6}, lambda o: torch.optim.lr_scheduler.CosineAnnealingLR(o, 25, 1e-6)),
    ({"type": "cosine"}, lambda o: torch.optim.lr_scheduler.CosineAnnealingLR(o, 100, 0)),
]


@pytest.mark.parametrize("cfg,make", CASES)
def test_epoch_schedules_match_torch(cfg, make):
    opt_t = torch.optim.AdamW([torch.nn.Parameter(torch.zeros(1))], lr=1e-4)
    ref = make(opt_t)
    mine_opt = types.SimpleNamespace(lr=1e-4)
    mine = get_scheduler(mine_opt, cfg)
    for _epoch in range(70):
        opt_t.step()
        ref.step()
        mine.step()
        assert math.isclose(mine_opt.lr, opt_t.param_groups[0]["lr"], rel_tol=1e-9, abs_tol=1e-15)


def test_plateau_matches_torch():
    opt_t = torch.optim.AdamW([torch.nn.Parameter(torch.zeros(1))], lr=1e-3)
    ref = torch.optim.lr_scheduler.ReduceLROnPlateau(opt_t, mode="min", factor=0.5, patience=3, threshold=1e-4)
    mine_opt = types.SimpleNamespace(lr=1e-3)
    mine = get_scheduler(mine_opt, {"type": "reduce_on_plateau", "patience": 3})
    g = torch.Generator().manual_seed(0)
    metric = 1.0
    for epoch in range(80):
        metric = metric * (0.97 if epoch < 15 else 1.0) + 1e-3 * float(torch.rand((), generator=g))
        ref.step(metric)
        mine.step(metric)
        assert math.isclose(mine_opt.lr, opt_t.param_groups[0]["lr"], rel_tol=1e-12)
    assert mine_opt.lr < 1e-3


def test_none_and_unknown():
    o = types.SimpleNamespace(lr=1.0)
    assert get_scheduler(o, {}) is None
    assert get_scheduler(o, {"type": "none"}) is None
    with pytest.raises(ValueError, match="Unknown scheduler type"):
        get_scheduler(o, {"type": "warmup"})


def test_state_dict_round_trip():
    o = types.SimpleNamespace(lr=1e-4)
    s = get_scheduler(o, {"type": "cosine", "T_max": 10})
    for _ in range(4):
        s.step()
    sd = s.state_dict()
    o2 = types.SimpleNamespace(lr=1e-4)
    s2 = get_scheduler(o2, {"type": "cosine", "T_max": 10})
    s2.load_state_dict(sd)
    assert o2.lr == o.lr and s2.step() == s.step()
