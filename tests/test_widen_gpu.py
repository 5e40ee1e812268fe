"""GPU parity of the rows either side of the training step (8a rows a3 dropout, a16 FiLM, a22 linear
attention): the CUDA path against fixtures made by the reference's own code (tests/golden/*, see
oracle/make_golden_widen.py) and against the CPU oracle on other sizes."""
import pytest
import torch

from conftest import grad_rel_err, load_golden, rel_err
from oracle import feed_oracle as FO
from oracle import metrics_oracle as MO
from oracle import vae_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
LKW = dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0)
TINY = dict(input_channels=1, latent_dim=4, hidden_channels=8, num_res_blocks=1, ch_mult=(1, 2), resolution=16,
            attn_resolutions=[8])


def _golden_step(model, g, tol_act, tol_grad, per_tensor=True):
    import medvae_disentangled_multimodal_b200 as M
    model.load_state_dict(g["sd"])
    model = model.to(DEV)
    model.eps_override = g["eps"].to(DEV)
    x = g["x"].to(DEV)
    out = model(x)
    ld = M.VAELoss(**LKW)(inputs=x, reconstructions=out["reconstruction"], posteriors=out["posterior"], priors=out["prior"])
    ld["loss"].backward()
    torch.cuda.synchronize()
    for k in ("mean", "logvar", "z", "reconstruction"):
        assert rel_err(out[k], g["out"][k]) < tol_act, (k, rel_err(out[k], g["out"][k]))
    for k, v in g["loss"].items():
        assert abs(float(ld[k]) - float(v)) <= max(tol_act, 1e-5) * max(1.0, abs(float(v))), k
    grads = {n: p.grad for n, p in model.named_parameters()}
    num = den = 0.0
    for n in g["grads"]:
        e = grad_rel_err(n, grads[n], g["grads"])
        num += float((grads[n].double().cpu() - g["grads"][n].double()).pow(2).sum())
        den += float(g["grads"][n].double().pow(2).sum())
        # bf16 on these 8-channel models: biases whose gradient is a small remainder of cancelling paths are
        # judged inside the whole-gradient error below; every weight tensor is judged on its own
        if tol_grad < 1e-2 or (per_tensor and grads[n].dim() > 1):
            assert e < tol_grad or float((grads[n].cpu() - g["grads"][n]).abs().max()) < 1e-7, (n, e)
    assert (num / den) ** 0.5 < (tol_grad / 2 if per_tensor else tol_grad), (num / den) ** 0.5
    return model


@pytest.mark.parametrize("dtype,tol_act,tol_grad", [("fp32", 1e-5, 1e-4), ("bf16", 1e-1, 2e-1)])
def test_linear_attention_vae_golden(dtype, tol_act, tol_grad):
    """BaseVAE(attn_type="linear"): LinAttnBlock (encoder_decoder.py:36-65) in encoder level 1, both mids, decoder"""
    import medvae_disentangled_multimodal_b200 as M
    g = load_golden("linattn_vae_16")
    model = M.BaseVAE(attn_type="linear", compute_dtype=dtype, **TINY)
    assert set(model.state_dict()) == set(g["sd"])
    # bf16 on this model (8 channels, attention without any normalisation) is judged on the activations and the
    # whole gradient vector; the bf16 kernel itself is checked per element in the next test, fp32 mode per tensor here.
    # Yardstick: the reference itself under torch.autocast(bfloat16) deviates from its fp32 run on this fixture by
    # 9.3 % (reconstruction) and 13.9 % (gradient vector) — vs 1.7 % / 2.3 % for the same model with AttnBlock.
    _golden_step(model, g, tol_act, tol_grad, per_tensor=(dtype == "fp32"))


def test_linear_attention_kernel_vs_oracle_bf16():
    """the core alone on a wider head (C = 64, T = 144), bf16 storage"""
    from medvae_disentangled_multimodal_b200 import ops
    torch.manual_seed(0)
    b, h, w, c = 3, 12, 12, 64
    qkv = torch.randn(b, h, w, 3 * c)
    dout = torch.randn(b, h, w, c)
    q, k, v = [t.reshape(b, h * w, c).clone().requires_grad_(True) for t in qkv.bfloat16().float().split(c, dim=-1)]
    ks = k.softmax(dim=1)
    ref = torch.einsum("btd,bde->bte", q, torch.einsum("btd,bte->bde", ks, v))
    ref.backward(dout.bfloat16().float().reshape(b, h * w, c))
    qkv_d = qkv.to(DEV).bfloat16()
    out, ks_d, ctx = ops.linattn_fwd(qkv_d)
    dqkv = ops.linattn_bwd(qkv_d, ks_d, ctx, dout.to(DEV).bfloat16())
    torch.cuda.synchronize()
    assert rel_err(out.float().reshape(b, h * w, c), ref) < 1e-2
    want = torch.cat([q.grad, k.grad, v.grad], dim=-1)
    assert rel_err(dqkv.float().reshape(b, h * w, 3 * c), want) < 1e-2


@pytest.mark.parametrize("dtype,tol_act,tol_grad", [("fp32", 1e-5, 1e-4), ("bf16", 8e-2, 2e-1)])
def test_dropout_vae_golden_on_the_reference_masks(dtype, tol_act, tol_grad):
    """training-mode dropout 0.25 with the keep masks the reference's nn.Dropout drew"""
    import medvae_disentangled_multimodal_b200 as M
    g = load_golden("dropout_vae_16")
    model = M.BaseVAE(dropout=0.25, compute_dtype=dtype, **TINY)
    model.train()
    for stack, prefix in ((model.encoder, "encoder."), (model.decoder, "decoder.")):
        stack.dropout_mask_override = {}
        for name, mod in stack.named_modules():
            key = prefix + name
            if key in g["masks"]:
                stack.dropout_mask_override[mod] = g["masks"][key].permute(0, 2, 3, 1).contiguous().to(DEV)
        assert len(stack.dropout_mask_override) >= 3
    _golden_step(model, g, tol_act, tol_grad)


def test_dropout_kernel_statistics_and_determinism():
    from medvae_disentangled_multimodal_b200 import ops
    x = torch.ones(1 << 20, device=DEV)
    y1, m1 = ops.dropout_apply(x, 0.3, 1234, 7, want_mask=True)
    y2, m2 = ops.dropout_apply(x, 0.3, 1234, 7, want_mask=True)
    y3, m3 = ops.dropout_apply(x, 0.3, 1234, 8, want_mask=True)
    assert torch.equal(y1, y2) and torch.equal(m1, m2)              # counter based: same (seed, offset) -> same mask
    assert not torch.equal(m1, m3)
    keep = float(m1.float().mean())
    assert abs(keep - 0.7) < 3e-3                                   # 1M Bernoulli(0.7): sigma = 4.6e-4
    assert torch.equal(y1, m1.float() / 0.7) or rel_err(y1, m1.float() / 0.7) < 1e-6
    # backward = same call on the gradient
    dy = torch.randn(1 << 20, device=DEV)
    dx, _ = ops.dropout_apply(dy, 0.3, 1234, 7)
    assert rel_err(dx, dy * m1.float() / 0.7) < 1e-6
    # bf16 storage, odd length
    xb = torch.randn(1001, device=DEV).bfloat16()
    yb, mb = ops.dropout_apply(xb, 0.5, 5, 1, want_mask=True)
    assert rel_err(yb.float(), (xb.float() * 2 * mb.float()).bfloat16().float()) < 1e-6


def test_dropout_model_trains_and_eval_is_identity():
    """own Philox masks: eval mode ignores dropout; training mode differs from eval and backpropagates"""
    import medvae_disentangled_multimodal_b200 as M
    torch.manual_seed(1)
    model = M.BaseVAE(dropout=0.2, compute_dtype="fp32", **TINY).to(DEV)
    ref = M.BaseVAE(dropout=0.0, compute_dtype="fp32", **TINY).to(DEV)
    ref.load_state_dict(model.state_dict())
    x = torch.rand(4, 1, 16, 16, device=DEV) * 2 - 1
    eps = torch.randn(4, 4, 8, 8, device=DEV)
    model.eps_override = ref.eps_override = eps
    model.eval(); ref.eval()
    assert rel_err(model(x)["reconstruction"], ref(x)["reconstruction"]) < 1e-6
    model.train()
    out = model(x)
    assert rel_err(out["reconstruction"], ref(x)["reconstruction"]) > 1e-3
    out["reconstruction"].square().mean().backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())
    out2 = model(x)                                                  # next step: a new mask
    assert rel_err(out2["reconstruction"], out["reconstruction"]) > 1e-4


def test_film_layer_golden():
    import medvae_disentangled_multimodal_b200 as M
    g = load_golden("film")
    layer = M.FiLMLayer(12, 24)
    layer.load_state_dict(g["sd"])
    layer = layer.to(DEV)
    f = g["features"].to(DEV).requires_grad_(True)
    c = g["cond"].to(DEV).requires_grad_(True)
    out = layer(f, c)
    assert rel_err(out, g["out"]) < 1e-6
    (out * g["w"].to(DEV)).sum().backward()
    assert rel_err(f.grad, g["d_features"]) < 1e-6 and rel_err(c.grad, g["d_cond"]) < 1e-5
    for n, p in layer.named_parameters():
        assert rel_err(p.grad, g["grads"][n]) < 1e-5, n


def test_film_in_conditional_vae_state_dict():
    import medvae_disentangled_multimodal_b200 as M
    m = M.ConditionalVAE(condition_method="film", input_channels=3, latent_dim=4, hidden_channels=8, num_res_blocks=1,
                         ch_mult=(1, 2), attn_resolutions=[], resolution=16).to(DEV)
    assert [l.scale_transform.out_features for l in m.film_layers] == [8, 16]
    y = m.film_layers[1](torch.randn(2, 16, 4, 4, device=DEV), m.get_modality_condition("oct").to(DEV).repeat(2, 1))
    assert y.shape == (2, 16, 4, 4)


# ------------------------------------------------------------------------------------------------ metrics
def test_metrics_golden_and_oracle():
    from medvae_disentangled_multimodal_b200 import metrics as MM
    g = load_golden("metrics")
    got = MM.compute_reconstruction_metrics(g["x"].to(DEV), g["r"].to(DEV))
    for k, v in g["recon"].items():
        assert abs(got[k] - float(v)) <= 2e-5 * max(1.0, abs(float(v))), (k, got[k], float(v))
    for a, b, want in ((g["mean"], g["logvar"], g["kl"]), (g["mean2"], g["logvar2"], g["kl2"])):
        got = MM.compute_kl_metrics(a.to(DEV), b.to(DEV))
        for k, v in want.items():
            assert abs(got[k] - float(v)) <= 1e-5 * max(1.0, abs(float(v))), (k, got[k], float(v))
    got = MM.compute_latent_metrics(g["z"].to(DEV))
    for k, v in g["latent"].items():
        assert abs(got[k] - float(v)) <= 1e-5 * max(1.0, abs(float(v))), (k, got[k], float(v))


@pytest.mark.parametrize("shape", [(5, 1, 28, 28), (3, 3, 64, 64), (2, 3, 128, 128), (2, 1, 37, 53), (2, 1, 11, 11)])
def test_recon_metrics_vs_oracle_sizes(shape):
    """tile edges (sizes that are not multiples of the 32x32 tile), the smallest image SSIM is defined on (11x11)"""
    from medvae_disentangled_multimodal_b200 import metrics as MM
    g = torch.Generator().manual_seed(shape[-1])
    x = torch.rand(*shape, generator=g) * 2 - 1
    r = (0.8 * x + 0.1 * torch.randn(*shape, generator=g)).contiguous()
    want = MO.compute_reconstruction_metrics(x, r)
    got = MM.compute_reconstruction_metrics(x.to(DEV), r.to(DEV))
    for k in want:
        assert abs(got[k] - want[k]) <= 2e-5 * max(1.0, abs(want[k])), (k, got[k], want[k])


def test_recon_metrics_properties_at_full_size():
    """C4-sized batch (1024 x 3 x 64 x 64): identical images -> mse 0, ssim 1; SSIM is symmetric; a batch made of
    copies of one pair has the pair's metrics (a checksum of checksums); two runs agree bit for bit"""
    from medvae_disentangled_multimodal_b200 import metrics as MM
    g = torch.Generator().manual_seed(3)
    x1 = torch.rand(1, 3, 64, 64, generator=g) * 2 - 1
    r1 = (x1 + 0.3 * torch.randn(1, 3, 64, 64, generator=g)).contiguous()
    x, r = x1.repeat(1024, 1, 1, 1).to(DEV), r1.repeat(1024, 1, 1, 1).to(DEV)
    same = MM.reconstruction_metrics_device(x, x).tolist()
    assert same[0] == 0.0 and same[1] == 0.0 and same[2] == float("inf") and abs(same[3] - 1.0) < 1e-6
    big = MM.reconstruction_metrics_device(x, r)
    assert torch.equal(big, MM.reconstruction_metrics_device(x, r))
    one = MM.reconstruction_metrics_device(x1.to(DEV), r1.to(DEV))
    assert rel_err(big, one) < 1e-6
    assert abs(float(MM.reconstruction_metrics_device(r, x)[3]) - float(big[3])) < 1e-6


def test_validation_step_through_trainstep():
    """TrainStep.validate = validation_step (lightning_module.py:220-310): metrics of the eval-mode forward + loss"""
    import medvae_disentangled_multimodal_b200 as M
    from medvae_disentangled_multimodal_b200.train import TrainStep
    torch.manual_seed(2)
    kw = dict(input_channels=3, latent_dim=4, hidden_channels=16, num_res_blocks=1, ch_mult=(1, 2), attn_resolutions=[8],
              resolution=16, condition_method="concat")
    model = M.ConditionalVAE(compute_dtype="fp32", **kw)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    x, onehot, _ = O.synthetic_batch(6, 3, 16)
    eps = torch.randn(6, 4, 8, 8, generator=torch.Generator().manual_seed(4))
    out = O.vae_forward(sd, x, eps, onehot)
    want = {f"val/{k}": v for k, v in {**MO.compute_reconstruction_metrics(x, out["reconstruction"]),
                                       **MO.compute_kl_metrics(out["mean"], out["logvar"]),
                                       **MO.compute_latent_metrics(out["z"])}.items()}
    want["val/loss"] = float(O.vae_loss(out["reconstruction"], x, out["mean"], out["logvar"], **LKW)["loss"])
    model = model.to(DEV)
    model.eps_override = eps.to(DEV)
    ts = TrainStep(model, M.VAELoss(**LKW))
    got = ts.validate(x, onehot, with_latent=True)
    ts.close()
    assert set(got) == set(want)
    for k, v in want.items():
        assert abs(got[k] - v) <= 5e-5 * max(1.0, abs(v)), (k, got[k], v)
    assert model.training


# ------------------------------------------------------------------------------------------------ batch feed
@pytest.mark.parametrize("tag", ["mixed", "gray", "rgb"])
def test_batch_feed_golden_bit_exact(tag):
    from medvae_disentangled_multimodal_b200.data import DeviceBatchFeed
    g = load_golden("feed")
    feed = DeviceBatchFeed(g["staged"], g["src_c"], g["modality_idx"])
    x, labels, onehot, idx = feed.batch(g[f"{tag}_sel"])
    torch.cuda.synchronize()
    assert x.shape == g[f"{tag}_x"].shape and x.dtype == torch.float32
    assert torch.equal(x.cpu(), g[f"{tag}_x"])                       # bit for bit
    assert torch.equal(onehot.cpu(), g[f"{tag}_onehot"]) and torch.equal(idx.cpu(), g[f"{tag}_idx"])
    assert labels.shape == (len(g[f"{tag}_sel"]), 1)


def test_batch_feed_full_size_vs_oracle_sample_and_properties():
    """C4-sized feed (1024 x 64 x 64): every value is (k/255 - 0.5)/0.5 for the stored byte k in colour samples,
    padded channels of gray samples are exactly zero, and a strided sample of rows equals the oracle bit for bit"""
    from medvae_disentangled_multimodal_b200.data import GRAY_12, DeviceBatchFeed
    g = torch.Generator().manual_seed(8)
    n = 1024
    staged = torch.randint(0, 256, (n, 64, 64, 3), generator=g, dtype=torch.uint8)
    idx = torch.arange(n) % 12
    src_c = torch.tensor([1 if int(m) in (0, 3, 6, 9, 10, 11) else 3 for m in idx], dtype=torch.int32)   # tissue: gray source -> RGB
    feed = DeviceBatchFeed(staged, src_c, idx)
    x, _lab, onehot, idx_d = feed.batch(torch.arange(n))
    x = x.cpu()
    gray = torch.tensor([int(m) in GRAY_12 for m in idx])
    assert float(x[gray][:, 1:].abs().max()) == 0.0
    lut = (torch.arange(256, dtype=torch.float32).div(255) - 0.5) / 0.5
    rgb3 = (~gray) & (src_c == 3)
    assert torch.equal(x[rgb3], lut[staged[rgb3].long()].permute(0, 3, 1, 2))
    sel = list(range(0, n, 37))
    want, want_onehot, _ = FO.feed_batch(staged, src_c, idx, sel)
    assert torch.equal(x[sel], want) and torch.equal(onehot.cpu()[sel], want_onehot)
    assert torch.equal(idx_d.cpu(), idx)


def test_batch_feed_into_training_step():
    """the fed batch is what TrainStep consumes: one step on it runs and gives a finite loss"""
    import medvae_disentangled_multimodal_b200 as M
    from medvae_disentangled_multimodal_b200.data import DeviceBatchFeed
    from medvae_disentangled_multimodal_b200.train import TrainStep
    g = load_golden("feed")
    feed = DeviceBatchFeed(g["staged"], g["src_c"], g["modality_idx"])
    x, _labels, onehot, _idx = feed.batch(list(range(12)))
    torch.manual_seed(0)
    model = M.ConditionalVAE(input_channels=3, latent_dim=4, hidden_channels=16, num_res_blocks=1, ch_mult=(1, 2, 4),
                             attn_resolutions=[14], resolution=28, condition_method="concat").to(DEV)
    ts = TrainStep(model, M.VAELoss(**LKW))
    l0 = float(ts.step(x, onehot)["loss"])
    l1 = float(ts.step(x, onehot)["loss"])
    ts.close()
    assert l0 == l0 and l1 == l1 and l1 < l0 * 1.5
