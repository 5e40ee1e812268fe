"""The device RNG against the numpy Philox restatement (oracle/philox_oracle.py, pinned to the Random123
known-answer vectors by tests/test_philox_cpu.py): dropout keep masks bit for bit, reparameterisation
noise to the accuracy of the fast log / cos intrinsics.  Streams are addressed exactly as the model does
(seed, offset) through the C ABI."""
import numpy as np
import pytest
import torch

from oracle import philox_oracle as P

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops():
    from medvae_disentangled_multimodal_b200 import ops
    return ops


@pytest.mark.parametrize("n,p,seed,offset", [(7, 0.1, 1234, 5), (4096, 0.1, 0xDEADBEEFCAFE, 3 << 20),
                                             (100003, 0.5, 2 ** 62 + 11, 2 ** 40), (1 << 20, 0.25, 42, 0)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_dropout_mask_is_the_philox_stream_bit_for_bit(n, p, seed, offset, dtype):
    ops = _ops()
    x = torch.randn(n, generator=torch.Generator().manual_seed(3)).to(DEV).to(dtype)
    y, mask = ops.dropout_apply(x, p, seed, offset, want_mask=True)
    torch.cuda.synchronize()
    want = P.dropout_keep_mask(n, p, seed, offset)
    assert np.array_equal(mask.cpu().numpy().astype(bool), want)
    keep = torch.from_numpy(want).to(DEV)
    scale = float(np.float32(1.0) / (np.float32(1.0) - np.float32(p)))     # fp32 arithmetic, as in the kernel
    ref = torch.where(keep, (x.float() * scale).to(dtype), torch.zeros((), dtype=dtype, device=DEV))
    assert torch.equal(y, ref)


@pytest.mark.parametrize("B,zc,R,seed,offset", [(4, 8, 7, 1234, 1), (3, 5, 4, 2 ** 40 + 7, 2 ** 33), (64, 32, 8, 7, 0)])
def test_reparam_noise_is_box_muller_on_the_philox_stream(B, zc, R, seed, offset):
    ops = _ops()
    mom = torch.zeros(B, R, R, 2 * zc, device=DEV)           # mu = 0, logvar = 0: z == eps
    _, _, z, _, eps, _ = ops.reparam_kl_fwd(mom, None, zc, 0.0, seed, offset)
    torch.cuda.synchronize()
    got = eps.cpu().numpy().reshape(-1).astype(np.float64)   # NCHW element order = stream order
    assert np.isfinite(got).all() and torch.equal(z, eps)
    want = P.normal_noise(B * zc * R * R, seed, offset).astype(np.float64)
    err = np.abs(got - want)
    # __logf is good to ~4e-7 absolute near u1 = 1, which the root amplifies for the few samples with |z| << 1;
    # everywhere else the fast intrinsics are good to ~1e-5.  A wrong word, block or stream gives O(1) errors.
    assert err.max() < 5e-3, err.max()
    assert np.median(err) < 1e-5 and (err > 2e-4).mean() < 1e-2


def test_noise_of_a_sample_does_not_depend_on_the_split():
    """the Philox element counter is the element's index in the GLOBAL batch, so two half batches
    addressed by their first sample (what two ranks, or two micro-batches, do) draw the noise of the whole batch."""
    ops = _ops()
    B, zc, R, seed, offset = 8, 5, 3, 99, 7          # odd sample size: halves start on an odd element
    mom = torch.zeros(B, R, R, 2 * zc, device=DEV)
    whole = ops.reparam_kl_fwd(mom, None, zc, 0.0, seed, offset)[4]
    lo = ops.reparam_kl_fwd(mom[:3].contiguous(), None, zc, 0.0, seed, offset, sample_base=0)[4]
    hi = ops.reparam_kl_fwd(mom[3:].contiguous(), None, zc, 0.0, seed, offset, sample_base=3)[4]
    torch.cuda.synchronize()
    assert torch.equal(torch.cat([lo, hi]), whole)
    assert not torch.equal(hi[:3], lo)                # different samples, different noise
    want = P.normal_noise(5 * zc * R * R, seed, offset, elem_base=3 * zc * R * R)
    assert np.abs(hi.cpu().numpy().reshape(-1) - want).max() < 5e-3


def test_train_step_with_own_noise_is_independent_of_the_micro_batch_split():
    """TrainStep addresses the noise by (optimizer step, global sample index): with the model's own Philox noise two
    micro-batches of 4 take the same steps as one batch of 8 (fp32 verification mode)."""
    import medvae_disentangled_multimodal_b200 as M
    from medvae_disentangled_multimodal_b200.train import TrainStep
    from oracle import vae_oracle as O
    kw = dict(input_channels=3, latent_dim=8, hidden_channels=64, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=[8],
              resolution=16, condition_method="concat")
    batches = [O.synthetic_batch(8, 3, 16, seed=300 + i) for i in range(2)]
    curves = []
    for mb in (None, 4):
        torch.manual_seed(3)
        model = M.ConditionalVAE(compute_dtype="fp32", **kw).to(DEV)
        model.rng_seed = 21
        ts = TrainStep(model, M.VAELoss(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0), micro_batch=mb,
                       lr=1e-4, weight_decay=1e-4, max_norm=1.0)
        curve = []
        for i in range(4):
            x, onehot, _ = batches[i % 2]
            curve.append(float(ts.step(x.to(DEV), onehot.to(DEV))["loss"]))
        ts.close()
        assert model._rng_step == 4 and model._rng_sample_base == 0
        curves.append(curve)
    for a, b in zip(*curves):
        assert abs(a - b) <= 2e-5 * abs(a), curves
    assert len(set(curves[0])) == 4


def test_dropout_mask_of_a_sample_does_not_depend_on_the_split():
    """SURVEY 8e for the dropout masks: the Philox counter is the element's index in the GLOBAL batch (elem_base), so the
    two parts of a split batch — even one that starts on an element index that is not a multiple of 4 — draw exactly
    the masks of the whole batch, bit for bit, and match the numpy restatement."""
    ops = _ops()
    B, per = 6, 7 * 5 * 3                                 # odd elements per sample: parts start inside a Philox block
    x = torch.ones(B, 7, 5, 3, device=DEV)
    _, whole = ops.dropout_apply(x, 0.4, 77, 9, want_mask=True)
    _, lo = ops.dropout_apply(x[:1].contiguous(), 0.4, 77, 9, want_mask=True, elem_base=0)
    _, hi = ops.dropout_apply(x[1:].contiguous(), 0.4, 77, 9, want_mask=True, elem_base=1 * per)
    torch.cuda.synchronize()
    assert torch.equal(torch.cat([lo, hi]), whole)
    want = P.dropout_keep_mask(5 * per, 0.4, 77, 9, elem_base=per)
    assert np.array_equal(hi.cpu().numpy().reshape(-1).astype(bool), want)


def test_model_dropout_masks_do_not_depend_on_the_micro_batch_split():
    """through TrainStep: the same step with the batch in one pass and in micro-batches of 3 (fp32 mode, lr 0, injected
    reparameterisation noise) gives the same loss: every sample kept its dropout masks"""
    import medvae_disentangled_multimodal_b200 as M
    from medvae_disentangled_multimodal_b200.train import TrainStep
    from oracle import vae_oracle as O
    kw = dict(input_channels=3, latent_dim=8, hidden_channels=32, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=[8],
              resolution=16, dropout=0.3)
    x, onehot, _ = O.synthetic_batch(8, 3, 16, seed=5)
    eps = torch.randn(8, 8, 8, 8, generator=torch.Generator().manual_seed(2)).to(DEV)
    losses = []
    for mb in (None, 3):
        torch.manual_seed(3)
        model = M.ConditionalVAE(compute_dtype="fp32", **kw).to(DEV)
        model.eps_override = eps
        model.encoder.dropout_seed = model.decoder.dropout_seed = 11
        ts = TrainStep(model, M.VAELoss("mse", 1.0, 1.0), micro_batch=mb, lr=0.0, weight_decay=0.0)
        losses.append([float(ts.step(x.to(DEV), onehot.to(DEV))["loss"]) for _ in range(2)])
        ts.close()
    assert losses[0][0] != losses[0][1]                         # the stream advances from step to step
    for a, b in zip(*losses):
        assert abs(a - b) <= 2e-6 * abs(a), losses
