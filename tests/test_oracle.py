"""Pins oracle/vae_oracle.py against golden vectors produced by the unmodified reference
(oracle/make_golden.py).  CPU only."""
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import vae_oracle as O

TOL = 2e-5  # fp32 CPU vs fp32 CPU, different op order only


@pytest.mark.parametrize("name,loss_kwargs,conditional", [
    ("base_vae_28", dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0), False),
    ("beta_vae_28", dict(recon_loss_type="l1", kl_weight=4.0, recon_weight=1.0), False),
    ("cvae_16", dict(recon_loss_type="bce", kl_weight=0.5, recon_weight=2.0), True),
    ("cvae_32", dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0), True),
])
def test_vae_step_matches_reference(name, loss_kwargs, conditional):
    g = load_golden(name)
    taps = {}
    ld, out, grads = O.training_step_grads(g["sd"], g["x"], g["eps"],
                                           g["cond"] if conditional else None, loss_kwargs, taps)
    for k in ("reconstruction", "mean", "logvar", "z"):
        assert rel_err(out[k], g["out"][k]) < TOL, k
    for k in ("loss", "recon_loss", "kl_loss"):
        assert abs(float(ld[k]) - float(g["loss"][k])) <= TOL * max(1.0, abs(float(g["loss"][k]))), k
    for k, v in g["taps"].items():
        assert k in taps, f"oracle has no activation named {k}"
        assert rel_err(taps[k], v) < TOL, k
    assert set(grads) == set(g["grads"])
    for k, v in g["grads"].items():
        assert rel_err(grads[k], v) < 1e-4 or float((grads[k] - v).abs().max()) < 1e-7, k


@pytest.mark.parametrize("name", ["disentangled_28_mixed", "disentangled_28_gray"])
def test_disentangled_matches_reference(name):
    g = load_golden(name)
    sd = {k: v.clone().requires_grad_(True) for k, v in g["sd"].items()}
    out = O.dis_forward(sd, g["x"], g["idx"], g["eps"], shared=4, modal=4)
    ld = O.dis_loss(out, g["x"])
    for k in ("reconstruction", "mean", "logvar", "z", "separation_loss", "contrastive_loss"):
        assert out[k].shape == g["out"][k].shape, k
        assert rel_err(out[k], g["out"][k]) < TOL, k
    for k, v in g["loss"].items():
        assert abs(float(ld[k]) - float(v)) <= 5e-5 * max(1.0, abs(float(v))), k
    names = [n for n in sd if sd[n].is_floating_point()]
    gr = torch.autograd.grad(ld["loss"], [sd[n] for n in names], allow_unused=True)
    for n, gi in zip(names, gr):
        if n not in g["grads"]:
            continue
        gi = gi if gi is not None else torch.zeros_like(sd[n])
        assert rel_err(gi, g["grads"][n]) < 2e-4 or float((gi - g["grads"][n]).abs().max()) < 1e-6, n


def test_reference_shape_assertions():
    """test_modality_channels.py:83-110 / quick_test.py:63-70: gray batch -> 1 channel."""
    g = load_golden("disentangled_28_gray")
    assert tuple(g["out"]["reconstruction"].shape) == (2, 1, 28, 28)
    g = load_golden("disentangled_28_mixed")
    assert tuple(g["out"]["reconstruction"].shape) == (6, 3, 28, 28)
