"""CPU: the oracle's restatements of the rows either side of the training step (linear attention, dropout,
FiLM, validation metrics, batch feed) against fixtures produced by the reference's own code
(oracle/make_golden_widen.py), plus the known answers torchmetrics publishes for PSNR / SSIM."""
import pytest
import torch

from conftest import grad_rel_err, load_golden, rel_err
from oracle import feed_oracle as FO
from oracle import metrics_oracle as MO
from oracle import vae_oracle as O

LKW = dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0)


def _check_step(g, ld, out, grads, tol=2e-5):
    for k in ("reconstruction", "mean", "logvar", "z"):
        assert rel_err(out[k], g["out"][k]) < tol, k
    for k, v in g["loss"].items():
        assert abs(float(ld[k]) - float(v)) <= 5e-5 * max(1.0, abs(float(v))), k
    for n, gr in g["grads"].items():
        assert grad_rel_err(n, grads[n], g["grads"]) < 1e-3, n


def test_linear_attention_vae_matches_reference():
    g = load_golden("linattn_vae_16")
    assert any(k.endswith("to_qkv.weight") for k in g["sd"])
    ld, out, grads = O.training_step_grads(g["sd"], g["x"], g["eps"], None, LKW)
    _check_step(g, ld, out, grads)


def test_dropout_vae_matches_reference_on_its_masks():
    g = load_golden("dropout_vae_16")
    masks = g["masks"]
    assert len(masks) >= 6 and all(0.5 < float(m.float().mean()) < 0.95 for m in masks.values())
    with O.dropout_masks(0.25, masks):
        ld, out, grads = O.training_step_grads(g["sd"], g["x"], g["eps"], None, LKW)
    _check_step(g, ld, out, grads)


def test_film_matches_reference():
    g = load_golden("film")
    f = g["features"].clone().requires_grad_(True)
    c = g["cond"].clone().requires_grad_(True)
    sd = {k: v.clone().requires_grad_(True) for k, v in g["sd"].items()}
    out = FO.film(f, c, sd)
    assert rel_err(out, g["out"]) < 1e-6
    (out * g["w"]).sum().backward()
    assert rel_err(f.grad, g["d_features"]) < 1e-6 and rel_err(c.grad, g["d_cond"]) < 1e-5
    for n, gr in g["grads"].items():
        assert rel_err(sd[n].grad, gr) < 1e-5, n


def test_kl_and_latent_metrics_match_reference():
    g = load_golden("metrics")
    for got, want in ((MO.compute_kl_metrics(g["mean"], g["logvar"]), g["kl"]),
                      (MO.compute_kl_metrics(g["mean2"], g["logvar2"]), g["kl2"]),
                      (MO.compute_latent_metrics(g["z"]), g["latent"])):
        for k, v in want.items():
            assert abs(got[k] - float(v)) <= 1e-6 * max(1.0, abs(float(v))), k


def test_psnr_ssim_known_answers():
    """the two examples in torchmetrics' docstrings (data_range inferred from the data there)"""
    pred = torch.tensor([[0.0, 1.0], [2.0, 3.0]])
    target = torch.tensor([[3.0, 2.0], [1.0, 0.0]])
    assert abs(float(MO.psnr(pred, target, 3.0)) - 2.5527) < 1e-4
    torch.manual_seed(42)
    preds = torch.rand([3, 3, 256, 256])
    target = preds * 0.75
    dr = max(float(preds.max() - preds.min()), float(target.max() - target.min()))
    assert abs(float(MO.ssim(preds, target, dr)) - 0.9219) < 1e-4
    # closed forms: identical images -> 1; two constant images -> (2ab + c1) / (a^2 + b^2 + c1)
    x = torch.rand(2, 1, 28, 28)
    assert abs(float(MO.ssim(x, x, 1.0)) - 1.0) < 1e-6
    a, b = 0.3, 0.5
    want = (2 * a * b + 1e-4) / (a * a + b * b + 1e-4)
    assert abs(float(MO.ssim(torch.full((1, 1, 20, 20), a), torch.full((1, 1, 20, 20), b), 1.0)) - want) < 1e-5


@pytest.mark.parametrize("tag", ["mixed", "gray", "rgb"])
def test_batch_feed_matches_reference_bit_for_bit(tag):
    g = load_golden("feed")
    x, onehot, idx = FO.feed_batch(g["staged"], g["src_c"], g["modality_idx"], g[f"{tag}_sel"].tolist())
    assert x.shape == g[f"{tag}_x"].shape
    assert torch.equal(x, g[f"{tag}_x"])
    assert torch.equal(onehot, g[f"{tag}_onehot"]) and torch.equal(idx, g[f"{tag}_idx"])
    assert x.shape[1] == (1 if tag == "gray" else 3)
