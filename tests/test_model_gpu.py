"""Whole-path parity tests (GPU): the drop-in models + VAELoss, forward and backward, against
(a) the golden vectors produced by the unmodified reference and (b) the CPU oracle on larger,
tensor-core-eligible shapes.  Reads like the reference's own smoke tests
(quick_test.py:38-70, test_modality_channels.py:50-112) with numeric assertions added."""
import pytest
import torch

from conftest import grad_rel_err, load_golden, rel_err
from oracle import vae_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"

GOLDEN = {
    "base_vae_28": ("BaseVAE", dict(input_channels=1, latent_dim=4, hidden_channels=8, num_res_blocks=1,
                                    ch_mult=(1, 2, 4), attn_resolutions=[14], resolution=28),
                    dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0), False),
    "beta_vae_28": ("BetaVAE", dict(beta=4.0, input_channels=3, latent_dim=4, hidden_channels=16, num_res_blocks=1,
                                    ch_mult=(1, 2), attn_resolutions=[], resolution=28),
                    dict(recon_loss_type="l1", kl_weight=4.0, recon_weight=1.0), False),
    "cvae_16": ("ConditionalVAE", dict(input_channels=3, latent_dim=4, hidden_channels=16, num_res_blocks=1,
                                       ch_mult=(1, 2), attn_resolutions=[8], resolution=16, condition_method="concat"),
                dict(recon_loss_type="bce", kl_weight=0.5, recon_weight=2.0), True),
    "cvae_32": ("ConditionalVAE", dict(input_channels=3, latent_dim=4, hidden_channels=8, num_res_blocks=2,
                                       ch_mult=(1, 2, 2, 4), attn_resolutions=[8, 4], resolution=32,
                                       condition_method="concat"),
                dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0), True),
}


def _step(model, crit, x, cond, eps):
    model.eps_override = eps
    model.zero_grad(set_to_none=True)
    out = model(x, cond) if cond is not None else model(x)
    ld = crit(inputs=x, reconstructions=out["reconstruction"], posteriors=out["posterior"], priors=out["prior"])
    ld["loss"].backward()
    torch.cuda.synchronize()
    grads = {n: p.grad for n, p in model.named_parameters()}
    return out, ld, grads


def _compare(out, ld, grads, ref_out, ref_loss, ref_grads, tol_act, tol_grad, tol_loss, tol_grad_global=None):
    """tol_grad bounds every parameter tensor's relative error; tol_grad_global (if given) bounds the
    relative error of the whole gradient vector (all parameters concatenated)."""
    worst = {}
    for k in ("mean", "logvar", "z", "reconstruction"):
        e = rel_err(out[k], ref_out[k])
        worst[k] = e
        assert e < tol_act, f"{k}: rel err {e:.3e}"
    for k in ("loss", "recon_loss", "kl_loss"):
        a, b = float(ld[k].detach()), float(ref_loss[k])
        assert abs(a - b) <= tol_loss * max(1.0, abs(b)), f"{k}: {a} vs {b}"
    bad = []
    num = den = 0.0
    for n, g in ref_grads.items():
        assert grads[n] is not None, f"no gradient for {n}"
        e = grad_rel_err(n, grads[n], ref_grads)
        num += float((grads[n].detach().double().cpu() - g.double()).pow(2).sum())
        den += float(g.double().pow(2).sum())
        if e >= tol_grad and float((grads[n].cpu().double() - g.double()).abs().max()) > 1e-7:
            bad.append((n, e))
    assert not bad, f"{len(bad)} gradient mismatches, worst: {sorted(bad, key=lambda t: -t[1])[:5]}"
    if tol_grad_global is not None:
        assert (num / den) ** 0.5 < tol_grad_global, f"whole-gradient rel err {(num / den) ** 0.5:.3e}"
    return worst


@pytest.mark.parametrize("name", list(GOLDEN))
def test_golden_fp32_mode(name):
    """fp32 verification mode vs the reference's own outputs: 1e-5 relative on activations
    (north_star), 1e-4 on parameter gradients (long fp32 reductions in a different order)."""
    import medvae_disentangled_multimodal_b200 as M
    cls, kw, lkw, conditional = GOLDEN[name]
    g = load_golden(name)
    model = getattr(M, cls)(compute_dtype="fp32", **kw).to(DEV)
    model.load_state_dict({k: v.to(DEV) for k, v in g["sd"].items()})
    crit = M.VAELoss(**lkw)
    cond = g["cond"].to(DEV) if conditional else None
    out, ld, grads = _step(model, crit, g["x"].to(DEV), cond, g["eps"].to(DEV))
    assert set(out) >= {"reconstruction", "mean", "logvar", "z", "prior", "posterior"}
    assert out["reconstruction"].shape == g["out"]["reconstruction"].shape
    _compare(out, ld, grads, g["out"], g["loss"], g["grads"], 1e-5, 1e-4, 1e-5)


@pytest.mark.parametrize("name", list(GOLDEN))
def test_golden_bf16_mode(name):
    """bf16 storage / fp32 accumulate vs the reference's fp32 results, END TO END through ~30-60
    layers of these very narrow golden models (8-32 channels).  The yardstick is the reference
    itself: tests/golden/bf16_yardstick.json holds how far the UNMODIFIED reference moves under
    torch.autocast(bfloat16) on the same inputs (oracle/make_bf16_yardstick.py; e.g. cvae_32:
    6.5e-2 on the reconstruction, 1.4e-1 on the whole gradient).  The CUDA path must stay within
    max(2e-2, 2x that) on activations, max(5e-2, 1.5x) on the whole gradient vector and max(0.25, 2x the
    reference's own worst tensor; measured worst 0.24 on a 16-element AttnBlock norm bias) on any single parameter tensor.  The per-layer 1e-2 bound of north_star is what
    tests/test_kernels_gpu.py holds for every kernel on identical inputs, and the
    tensor-core-sized models below hold 2e-2 / 6e-2 end to end."""
    import json
    import os
    import medvae_disentangled_multimodal_b200 as M
    from conftest import GOLDEN as GOLDEN_DIR
    yard = json.load(open(os.path.join(GOLDEN_DIR, "bf16_yardstick.json")))["fixtures"][name]
    cls, kw, lkw, conditional = GOLDEN[name]
    g = load_golden(name)
    model = getattr(M, cls)(compute_dtype="bf16", **kw).to(DEV)
    model.load_state_dict({k: v.to(DEV) for k, v in g["sd"].items()})
    crit = M.VAELoss(**lkw)
    cond = g["cond"].to(DEV) if conditional else None
    out, ld, grads = _step(model, crit, g["x"].to(DEV), cond, g["eps"].to(DEV))
    tol_act = max(2e-2, 2.0 * max(yard[k] for k in ("reconstruction", "mean", "logvar", "z")))
    # per parameter tensor: 2x what the reference's own bf16 run does on its worst tensor (0.04-0.27 on these fixtures:
    # gradients of 8-channel layers are sums of a few hundred bf16-rounded terms), at least 0.25
    _compare(out, ld, grads, g["out"], g["loss"], g["grads"], tol_act, min(0.4, max(0.25, 2.0 * yard["worst_tensor"])), 2e-2,
             tol_grad_global=max(5e-2, 1.5 * yard["whole_gradient"]))


BIG = [
    # name, cls, kwargs, batch, conditional — channel counts are multiples of 64 so every hidden
    # conv runs on the tcgen05 path in bf16 mode
    ("cvae64_16", "ConditionalVAE", dict(input_channels=3, latent_dim=64, hidden_channels=64, num_res_blocks=1,
                                         ch_mult=(1, 2), attn_resolutions=[8], resolution=16), 4, True),
    ("base64_28", "BaseVAE", dict(input_channels=1, latent_dim=64, hidden_channels=64, num_res_blocks=1,
                                  ch_mult=(1, 2, 2), attn_resolutions=[14], resolution=28), 3, False),
    ("cvae128_32", "ConditionalVAE", dict(input_channels=3, latent_dim=128, hidden_channels=128, num_res_blocks=1,
                                          ch_mult=(1, 2, 2), attn_resolutions=[8], resolution=32), 2, True),
]


def _big_case(name, cls, kw, batch, conditional, dtype, tols):
    import medvae_disentangled_multimodal_b200 as M
    torch.manual_seed(11)
    model = getattr(M, cls)(compute_dtype=dtype, **kw)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    x, onehot, _ = O.synthetic_batch(batch, kw["input_channels"], kw["resolution"])
    r = kw["resolution"] // 2 ** (len(kw["ch_mult"]) - 1)
    eps = torch.randn(batch, kw["latent_dim"], r, r, generator=torch.Generator().manual_seed(5))
    cond = onehot if conditional else None
    lkw = dict(recon_loss_type="mse", kl_weight=0.5, recon_weight=1.0)
    ref_loss, ref_out, ref_grads = O.training_step_grads(sd, x, eps, cond, lkw)
    model = model.to(DEV)
    crit = M.VAELoss(**lkw)
    out, ld, grads = _step(model, crit, x.to(DEV), cond.to(DEV) if conditional else None, eps.to(DEV))
    return _compare(out, ld, grads, ref_out, ref_loss, ref_grads, *tols)


@pytest.mark.parametrize("case", BIG, ids=[c[0] for c in BIG])
def test_oracle_fp32_mode(case):
    """128-channel models against the CPU oracle evaluated on this box.  The golden-vector tests above
    hold the 1e-5 bar against fixed reference outputs; here the yardstick itself is an fp32 chain of
    ~40 layers whose oneDNN kernels (and summation order) depend on the host CPU, so two fp32
    evaluations agree to a few 1e-5 at the end of the chain."""
    _big_case(*case, "fp32", (3e-5, 1e-4, 1e-5))


@pytest.mark.umma
@pytest.mark.parametrize("case", BIG, ids=[c[0] for c in BIG])
def test_oracle_bf16_tcgen05(case):
    _big_case(*case, "bf16", (2e-2, 6e-2, 2e-2))


def test_api_surface_and_shapes():
    """scripts/demo.sh:21-42 — BaseVAE(1,16,32,[1,2,4],1,[],resolution=28) on (4,1,28,28)."""
    import medvae_disentangled_multimodal_b200 as M
    model = M.BaseVAE(1, 16, 32, [1, 2, 4], 1, [], resolution=28).to(DEV)
    x = torch.randn(4, 1, 28, 28, device=DEV)
    out = model(x)
    assert out["reconstruction"].shape == (4, 1, 28, 28)
    assert out["mean"].shape == out["logvar"].shape == out["z"].shape == (4, 16, 7, 7)
    assert out["posterior"].loc.shape == (4, 16, 7, 7) and out["prior"].scale.shape == (4, 16, 7, 7)
    mean, logvar = model.encode(x)
    z = model.reparameterize(mean, logvar)
    assert model.decode(z).shape == (4, 1, 28, 28)
    with torch.no_grad():
        assert model.sample(3, torch.device(DEV)).shape == (3, 1, 28, 28)
    # standalone stages agree with the fused forward when the noise is pinned
    eps = torch.randn(4, 16, 7, 7, device=DEV)
    model.eps_override = eps
    out2 = model(x)
    m2, lv2 = model.encode(x)
    z2 = model.reparameterize(m2, lv2, eps=eps)
    # bf16 mode accumulates the fused GroupNorm statistics with fp32 atomics: two runs agree to bf16
    # rounding, not bit for bit (MVAE_GN_FUSE=0 restores the bitwise-reproducible two-pass kernels)
    assert rel_err(m2, out2["mean"]) < 2e-2 and rel_err(z2, out2["z"]) < 2e-2
    assert rel_err(model.decode(z2), out2["reconstruction"]) < 2e-2


def test_foreign_loss_through_distributions():
    """The reference's own VAELoss differentiates kl_divergence(posterior, prior); the Normal
    objects returned here must carry that gradient too."""
    import medvae_disentangled_multimodal_b200 as M
    from torch.distributions import kl_divergence
    model = M.BaseVAE(1, 4, 16, [1, 2], 1, [], resolution=16, compute_dtype="fp32").to(DEV)
    x = torch.randn(2, 1, 16, 16, device=DEV)
    model.eps_override = torch.randn(2, 4, 8, 8, device=DEV)
    out = model(x)
    (torch.nn.functional.mse_loss(out["reconstruction"], x) + kl_divergence(out["posterior"], out["prior"]).mean()).backward()
    g1 = {n: p.grad.clone() for n, p in model.named_parameters()}
    model.zero_grad()
    out = model(x)
    ld = M.VAELoss("mse", 1.0, 1.0)(inputs=x, reconstructions=out["reconstruction"], posteriors=out["posterior"],
                                    priors=out["prior"])
    ld["loss"].backward()
    for n, p in model.named_parameters():
        assert rel_err(p.grad, g1[n]) < 1e-4 or float((p.grad - g1[n]).abs().max()) < 1e-7, n


# ------------------------------------------------------------------ DisentangledConditionalVAE
DIS_KW = dict(num_modalities=5, shared_latent_dim=4, modality_latent_dim=4, resolution=28, hidden_channels=16,
              ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=[])


def _dis_step(name, dtype):
    import medvae_disentangled_multimodal_b200 as M
    g = load_golden(name)
    model = M.DisentangledConditionalVAE(compute_dtype=dtype, **DIS_KW).to(DEV)
    model.load_state_dict({k: v.to(DEV) for k, v in g["sd"].items()})
    model.eps_override = g["eps"].to(DEV)
    crit = M.DisentangledVAELoss("mse", 1.0, 1.0, 0.1, 0.05)
    x = g["x"].to(DEV)
    out = model(x, g["idx"])          # CPU index tensor: no host sync inside
    ld = crit(out, x)
    ld["loss"].backward()
    torch.cuda.synchronize()
    grads = {n: (p.grad if p.grad is not None else torch.zeros_like(p)) for n, p in model.named_parameters()}
    return g, out, ld, grads


@pytest.mark.parametrize("name", ["disentangled_28_mixed", "disentangled_28_gray"])
def test_disentangled_golden_fp32_mode(name):
    """the reference's own test shapes (test_modality_channels.py:50-112, quick_test.py:63-70) with the
    numbers pinned: outputs, the five loss terms and every parameter gradient."""
    g, out, ld, grads = _dis_step(name, "fp32")
    assert out["reconstruction"].shape == g["out"]["reconstruction"].shape
    for k in ("reconstruction", "mean", "logvar", "z", "separation_loss", "contrastive_loss"):
        assert rel_err(out[k], g["out"][k]) < 2e-5, f"{k}: {rel_err(out[k], g['out'][k]):.3e}"
    assert out["mu"] is out["mean"]
    for k, v in g["loss"].items():
        assert abs(float(ld[k].detach()) - float(v)) <= 2e-5 * max(1.0, abs(float(v))), k
    bad = []
    for n, ref in g["grads"].items():
        e = grad_rel_err(n, grads[n], g["grads"])
        tol = 5e-4 if n.endswith(".bias") else 2e-4   # bias gradients: long cancelling sums of four loss terms
        if e >= tol and float((grads[n].cpu().double() - ref.double()).abs().max()) > 1e-7:
            bad.append((n, e))
    assert not bad, sorted(bad, key=lambda t: -t[1])[:6]


@pytest.mark.parametrize("name", ["disentangled_28_mixed", "disentangled_28_gray"])
def test_disentangled_golden_bf16_mode(name):
    g, out, ld, grads = _dis_step(name, "bf16")
    for k in ("reconstruction", "mean", "logvar", "z"):
        assert rel_err(out[k], g["out"][k]) < 5e-2, f"{k}: {rel_err(out[k], g['out'][k]):.3e}"
    for k, v in g["loss"].items():
        assert abs(float(ld[k].detach()) - float(v)) <= 5e-2 * max(1.0, abs(float(v))), k
    num = sum(float((grads[n].double().cpu() - r.double()).pow(2).sum()) for n, r in g["grads"].items())
    den = sum(float(r.double().pow(2).sum()) for r in g["grads"].values())
    assert (num / den) ** 0.5 < 1e-1


def test_disentangled_api_surface():
    """API the reference's scripts use: encode/decode/partition/reconstruct/sample_conditional and the
    channel convention (all-gray batch -> 1 channel, mixed -> 3)."""
    import medvae_disentangled_multimodal_b200 as M
    model = M.DisentangledConditionalVAE(**DIS_KW).to(DEV)
    x = torch.rand(4, 3, 28, 28, device=DEV) * 2 - 1
    idx = torch.tensor([0, 1, 3, 4])
    out = model(x, idx, return_latents=True)
    assert out["reconstruction"].shape == (4, 3, 28, 28)
    assert out["z_shared"].shape == (4, 4) and out["z_modality"].shape == (4, 4)
    mu, lv = model.encode(x, idx)
    assert mu.shape == lv.shape == (4, 8, 14, 14)
    gray = model(x[:2, :1], torch.tensor([0, 3]))
    assert gray["reconstruction"].shape == (2, 1, 28, 28)
    with torch.no_grad():
        assert model.sample_conditional(3, torch.tensor([1, 2, 4]), torch.device(DEV)).shape == (3, 3, 28, 28)
        assert model.decode(out["z"]).shape == (4, 3, 28, 28)
    zs, zm = model.partition_latent(out["z"])
    assert zs.shape == zm.shape == (4, 4)
    # reconstruct_latent (:208-239) only reshapes when the partition sizes cover whole (R, R) maps
    full = model.reconstruct_latent(torch.zeros(2, 4 * 196, device=DEV), torch.zeros(2, 4 * 196, device=DEV))
    assert full.shape == (2, 8, 14, 14)
    # GPU-resident indices work too (one host sync for the channel-width decision)
    out2 = model(x, idx.to(DEV))
    assert out2["reconstruction"].shape == (4, 3, 28, 28)
    with pytest.raises(ValueError):
        M.DisentangledVAELoss("bce")


def test_vae_loss_refuses_a_prior_that_is_not_standard_normal():
    """VAELoss computes KL(posterior || N(0, 1)) in closed form: the models' own prior passes by its marker, a foreign
    N(0, 1) prior passes by value, anything else is refused instead of being silently treated as N(0, 1)"""
    import medvae_disentangled_multimodal_b200 as M
    from torch.distributions import Normal
    model = M.BaseVAE(1, 4, 16, [1, 2], 1, [], resolution=16, compute_dtype="fp32").to(DEV)
    x = torch.randn(2, 1, 16, 16, device=DEV)
    out = model(x)
    crit = M.VAELoss("mse", 1.0, 1.0)
    kw = dict(inputs=x, reconstructions=out["reconstruction"], posteriors=out["posterior"])
    a = crit(priors=out["prior"], **kw)
    b = crit(priors=Normal(torch.zeros_like(out["mean"]), torch.ones_like(out["mean"])), **kw)
    assert float(a["loss"]) == float(b["loss"])
    with pytest.raises(NotImplementedError):
        crit(priors=Normal(torch.zeros_like(out["mean"]), 2.0 * torch.ones_like(out["mean"])), **kw)
