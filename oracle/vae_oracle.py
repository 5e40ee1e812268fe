"""CPU oracle for the conditional-VAE training step.  TEST INFRASTRUCTURE ONLY.

This file is a functional, state_dict-driven restatement of the reference's hot path
(parsakzr/medvae-disentangled-multimodal @ ff15cf3) on plain torch CPU ops.  Only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import it;
the product package never does (it fails loudly without its CUDA library).

Parity status: PINNED.  ``oracle/make_golden.py`` imports the unmodified reference from
``/root/reference`` (possible in the build container only), runs it on seeded inputs with the
reparameterisation noise captured, and commits inputs / weights / outputs / gradients under
``tests/golden/``.  ``tests/test_oracle.py`` checks every function below against those vectors.
The leaf arithmetic the reference uses is PyTorch's own (``F.conv2d``, ``F.group_norm``,
``torch.bmm``, ``F.softmax``, ``F.interpolate``, ``kl_divergence``); PyTorch is installed on
both boxes, so the oracle calls the same leaf ops and restates everything above them.

Every function cites the reference file:line it follows (paths relative to /root/reference).
All tensors are NCHW, dtype follows the weights (fp32 or fp64).
"""
from __future__ import annotations

import math
import time
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]


# --------------------------------------------------------------------------------------
# L0 blocks  (src/models/encoder_decoder.py)
# --------------------------------------------------------------------------------------
def swish(x):
    """encoder_decoder.py:13-15  x * sigmoid(x)."""
    return x * torch.sigmoid(x)


def group_norm(sd: SD, key: str, x):
    """encoder_decoder.py:28-33  GroupNorm(min(32, C), C, eps=1e-6, affine)."""
    c = x.shape[1]
    return F.group_norm(x, min(32, c), sd[key + ".weight"], sd[key + ".bias"], eps=1e-6)


def conv(sd: SD, key: str, x, stride=1, padding=0):
    return F.conv2d(x, sd[key + ".weight"], sd[key + ".bias"], stride=stride, padding=padding)


_DROPOUT = {"p": 0.0, "masks": None}


class dropout_masks:
    """``with dropout_masks(p, {block key: NCHW keep mask})``: training-mode nn.Dropout(p) of every ResnetBlock
    (encoder_decoder.py:164) with the keep masks supplied from outside — torch's own dropout RNG cannot be
    reproduced by another implementation, so parity is checked on identical masks."""

    def __init__(self, p: float, masks: Dict[str, torch.Tensor]):
        self.p, self.masks = p, masks

    def __enter__(self):
        self.prev = dict(_DROPOUT)
        _DROPOUT.update(p=self.p, masks=self.masks)

    def __exit__(self, *exc):
        _DROPOUT.update(self.prev)


def resnet_block(sd: SD, key: str, x, taps: Optional[dict] = None):
    """encoder_decoder.py:148-170.  GN1-swish-conv1-GN2-swish-dropout-conv2, 1x1
    nin_shortcut iff the channel count changes (key present), output x + h."""
    h = conv(sd, key + ".conv1", swish(group_norm(sd, key + ".norm1", x)), padding=1)
    if taps is not None:
        taps[key + ".conv1"] = h
    h = swish(group_norm(sd, key + ".norm2", h))
    if _DROPOUT["masks"] is not None and _DROPOUT["p"] > 0.0:
        h = h * _DROPOUT["masks"][key].to(h.dtype) / (1.0 - _DROPOUT["p"])
    h = conv(sd, key + ".conv2", h, padding=1)
    if (key + ".nin_shortcut.weight") in sd:
        x = conv(sd, key + ".nin_shortcut", x)
    out = x + h
    if taps is not None:
        taps[key] = out
    return out


def attn_block(sd: SD, key: str, x, taps: Optional[dict] = None):
    """encoder_decoder.py:83-107.  Single head, head_dim = C, scale C**-0.5, softmax over keys."""
    b, c, hh, ww = x.shape
    hn = group_norm(sd, key + ".norm", x)
    q = conv(sd, key + ".q", hn).reshape(b, c, hh * ww).permute(0, 2, 1)  # b, T, C
    k = conv(sd, key + ".k", hn).reshape(b, c, hh * ww)  # b, C, T
    v = conv(sd, key + ".v", hn).reshape(b, c, hh * ww)
    w = torch.bmm(q, k) * (int(c) ** (-0.5))  # [b, i, j]
    w = F.softmax(w, dim=2)
    h = torch.bmm(v, w.permute(0, 2, 1)).reshape(b, c, hh, ww)  # h[b,c,i] = sum_j v[b,c,j] w[b,i,j]
    out = x + conv(sd, key + ".proj_out", h)
    if taps is not None:
        taps[key] = out
    return out


def lin_attn_block(sd: SD, key: str, x, taps: Optional[dict] = None):
    """encoder_decoder.py:36-65 (LinAttnBlock: heads=1, dim_head=C).  to_qkv (no bias) -> softmax of k over the
    tokens -> context[d,e] = sum_n k[d,n] v[e,n] -> out[e,n] = sum_d context[d,e] q[d,n] -> to_out.  No norm,
    no residual."""
    b, c, hh, ww = x.shape
    qkv = F.conv2d(x, sd[key + ".to_qkv.weight"])
    q, k, v = qkv.reshape(b, 3, 1, c, hh * ww).unbind(1)          # "b (qkv heads c) h w -> qkv b heads c (h w)"
    k = k.softmax(dim=-1)
    context = torch.einsum("bhdn,bhen->bhde", k, v)
    out = torch.einsum("bhde,bhdn->bhen", context, q).reshape(b, c, hh, ww)
    out = conv(sd, key + ".to_out", out)
    if taps is not None:
        taps[key] = out
    return out


def any_attn_block(sd: SD, key: str, x, taps: Optional[dict] = None):
    """make_attn (encoder_decoder.py:18-25): which block was built is visible from the state_dict keys"""
    if (key + ".to_qkv.weight") in sd:
        return lin_attn_block(sd, key, x, taps)
    return attn_block(sd, key, x, taps)


def downsample(sd: SD, key: str, x):
    """encoder_decoder.py:184-188.  pad right/bottom by one, conv3x3 stride 2 padding 0."""
    return conv(sd, key + ".conv", F.pad(x, (0, 1, 0, 1), mode="constant", value=0), stride=2)


def upsample(sd: SD, key: str, x):
    """encoder_decoder.py:205-209.  nearest 2x then conv3x3 s1 p1."""
    return conv(sd, key + ".conv", F.interpolate(x, scale_factor=2.0, mode="nearest"), padding=1)


def _count(sd: SD, prefix: str) -> int:
    n = 0
    while any(k.startswith(f"{prefix}.{n}.") for k in sd):
        n += 1
    return n


def encoder_forward(sd: SD, x, prefix="encoder", taps: Optional[dict] = None):
    """encoder_decoder.py:303-328.  Structure is recovered from the state_dict keys, so the
    attn_resolutions / ch_mult bookkeeping of :248-275 is implied by which keys exist."""
    h = conv(sd, prefix + ".conv_in", x, padding=1)
    if taps is not None:
        taps[prefix + ".conv_in"] = h
    n_levels = _count(sd, prefix + ".down")
    for lv in range(n_levels):
        n_blocks = _count(sd, f"{prefix}.down.{lv}.block")
        n_attn = _count(sd, f"{prefix}.down.{lv}.attn")
        for ib in range(n_blocks):
            h = resnet_block(sd, f"{prefix}.down.{lv}.block.{ib}", h, taps)
            if n_attn > 0:
                h = any_attn_block(sd, f"{prefix}.down.{lv}.attn.{ib}", h, taps)
        if f"{prefix}.down.{lv}.downsample.conv.weight" in sd:
            h = downsample(sd, f"{prefix}.down.{lv}.downsample", h)
            if taps is not None:
                taps[f"{prefix}.down.{lv}.downsample"] = h
    h = resnet_block(sd, prefix + ".mid.block_1", h, taps)
    h = any_attn_block(sd, prefix + ".mid.attn_1", h, taps)
    h = resnet_block(sd, prefix + ".mid.block_2", h, taps)
    h = swish(group_norm(sd, prefix + ".norm_out", h))
    h = conv(sd, prefix + ".conv_out", h, padding=1)
    if taps is not None:
        taps[prefix + ".conv_out"] = h
    return h


def decoder_forward(sd: SD, z, prefix="decoder", taps: Optional[dict] = None):
    """encoder_decoder.py:421-451.  up.0 is the highest resolution (:415 inserts at 0)."""
    h = conv(sd, prefix + ".conv_in", z, padding=1)
    if taps is not None:
        taps[prefix + ".conv_in"] = h
    h = resnet_block(sd, prefix + ".mid.block_1", h, taps)
    h = any_attn_block(sd, prefix + ".mid.attn_1", h, taps)
    h = resnet_block(sd, prefix + ".mid.block_2", h, taps)
    n_levels = _count(sd, prefix + ".up")
    for lv in reversed(range(n_levels)):
        n_blocks = _count(sd, f"{prefix}.up.{lv}.block")
        n_attn = _count(sd, f"{prefix}.up.{lv}.attn")
        for ib in range(n_blocks):
            h = resnet_block(sd, f"{prefix}.up.{lv}.block.{ib}", h, taps)
            if n_attn > 0:
                h = any_attn_block(sd, f"{prefix}.up.{lv}.attn.{ib}", h, taps)
        if f"{prefix}.up.{lv}.upsample.conv.weight" in sd:
            h = upsample(sd, f"{prefix}.up.{lv}.upsample", h)
            if taps is not None:
                taps[f"{prefix}.up.{lv}.upsample"] = h
    h = swish(group_norm(sd, prefix + ".norm_out", h))
    h = conv(sd, prefix + ".conv_out", h, padding=1)
    if taps is not None:
        taps[prefix + ".conv_out"] = h
    return h


# --------------------------------------------------------------------------------------
# L1 models
# --------------------------------------------------------------------------------------
def condition_map(sd: SD, condition, channels: int, height: int, width: int):
    """conditional_vae.py:59-78,107-119.  Linear -> ReLU -> (C,8,8) -> bilinear resize."""
    m = F.relu(F.linear(condition, sd["condition_proj.0.weight"], sd["condition_proj.0.bias"]))
    m = m.reshape(condition.shape[0], channels, 8, 8)
    return F.interpolate(m, size=(height, width), mode="bilinear", align_corners=False)


def vae_forward(sd: SD, x, eps, condition=None, taps: Optional[dict] = None):
    """base_vae.py:89-118 / conditional_vae.py:121-164.  ``eps`` replaces randn_like (:86)."""
    xin = x
    if condition is not None and "condition_proj.0.weight" in sd:
        cmap = condition_map(sd, condition, x.shape[1], x.shape[2], x.shape[3])
        xin = torch.cat([x, cmap], dim=1)
    h = encoder_forward(sd, xin, taps=taps)
    mean, logvar = torch.chunk(h, 2, dim=1)  # base_vae.py:76
    z = mean + eps * torch.exp(0.5 * logvar)  # base_vae.py:83-87
    recon = decoder_forward(sd, z, taps=taps)
    return {"reconstruction": recon, "mean": mean, "logvar": logvar, "z": z}


def vae_loss(recon, x, mean, logvar, recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0):
    """src/losses/vae_losses.py:31-64.  kl_divergence(N(mu, sigma), N(0,1)).mean() in closed
    form (0.5 (e^lv + mu^2 - 1 - lv)), mean over all latent elements."""
    if recon_loss_type == "mse":
        rl = F.mse_loss(recon, x, reduction="mean")
    elif recon_loss_type == "l1":
        rl = F.l1_loss(recon, x, reduction="mean")
    elif recon_loss_type == "bce":
        rl = F.binary_cross_entropy_with_logits(recon, x, reduction="mean")
    else:
        raise ValueError(f"Unknown reconstruction loss type: {recon_loss_type}")
    kl = (0.5 * (torch.exp(logvar) + mean * mean - 1.0 - logvar)).mean()
    return {"loss": recon_weight * rl + kl_weight * kl, "recon_loss": rl, "kl_loss": kl}


# --------------------------------------------------------------------------------------
# Disentangled conditional VAE  (src/models/disentangled_conditional_vae.py)
# --------------------------------------------------------------------------------------
MODALITY_CHANNELS = {0: 1, 1: 3, 2: 3, 3: 1, 4: 3}  # :116-122


def dis_encode_input(sd: SD, x, idx):
    """:124-171.  1-channel modalities use channel 0 through their Conv2d(1,3,1) projector."""
    outs = []
    for i in range(x.shape[0]):
        m = min(int(idx[i]), len(MODALITY_CHANNELS) - 1)
        s = x[i : i + 1]
        if MODALITY_CHANNELS[m] == 1:
            s = F.conv2d(s[:, :1], sd[f"modality_input_projectors.{m}.weight"],
                         sd[f"modality_input_projectors.{m}.bias"])
        outs.append(s)
    return torch.cat(outs, dim=0)


def dis_decode_heads(sd: SD, r, idx):
    """:241-303.  per-modality conv3x3-ReLU-conv3x3, optional 3->1 projector, zero-pad to the
    widest sample in the batch."""
    outs, widest = [], 0
    n_dec = _count(sd, "modality_decoders")
    for i in range(r.shape[0]):
        m = min(int(idx[i]), n_dec - 1)
        s = r[i : i + 1]
        s = F.conv2d(s, sd[f"modality_decoders.{m}.0.weight"], sd[f"modality_decoders.{m}.0.bias"], padding=1)
        s = F.conv2d(F.relu(s), sd[f"modality_decoders.{m}.2.weight"], sd[f"modality_decoders.{m}.2.bias"], padding=1)
        if f"modality_output_projectors.{m}.weight" in sd:
            s = F.conv2d(s, sd[f"modality_output_projectors.{m}.weight"], sd[f"modality_output_projectors.{m}.bias"])
        outs.append(s)
        widest = max(widest, s.shape[1])
    outs = [F.pad(s, (0, 0, 0, 0, 0, widest - s.shape[1])) for s in outs]
    return torch.cat(outs, dim=0)


def dis_separation_loss(z, idx, shared: int, modal: int):
    """:305-349.  -mean(pdist(centroids of z_modality per present modality))."""
    zm = z.reshape(z.shape[0], -1)[:, shared : shared + modal]
    cents = [zm[idx == m].mean(dim=0) for m in torch.unique(idx)]
    if len(cents) < 2:
        return torch.zeros((), dtype=z.dtype)
    return -torch.pdist(torch.stack(cents), p=2).mean()


def dis_contrastive_loss(z, idx, shared: int, modal: int, temperature=0.1):
    """:351-386."""
    zm = z.reshape(z.shape[0], -1)[:, shared : shared + modal]
    zn = F.normalize(zm, p=2, dim=1)
    e = torch.exp(zn @ zn.t() / temperature)
    mask = (idx[None, :] == idx[:, None])
    mask = mask & ~torch.eye(len(idx), dtype=torch.bool)
    pos = (e * mask.to(e.dtype)).sum(dim=1)
    allsim = e.sum(dim=1) - torch.diag(e)
    l = -torch.log(pos / allsim + 1e-8)
    l = l[pos > 0]
    return l.mean() if len(l) > 0 else torch.zeros((), dtype=z.dtype)


def dis_forward(sd: SD, x, idx, eps, shared: int, modal: int):
    """:388-454 (NaN scrubs are no-ops on finite data)."""
    h = encoder_forward(sd, dis_encode_input(sd, x, idx))
    mu, logvar = torch.chunk(h, 2, dim=1)
    logvar = torch.clamp(logvar, -10.0, 10.0)
    mu = torch.clamp(mu, -10.0, 10.0)
    z = mu + eps * torch.exp(0.5 * logvar)
    recon = dis_decode_heads(sd, decoder_forward(sd, z), idx)
    return {
        "reconstruction": recon, "mean": mu, "mu": mu, "logvar": logvar, "z": z,
        "separation_loss": dis_separation_loss(z, idx, shared, modal),
        "contrastive_loss": dis_contrastive_loss(z, idx, shared, modal),
    }


def dis_loss(out, targets, recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0,
             separation_weight=0.1, contrastive_weight=0.05):
    """:510-573.  KL is a SUM normalised by targets.numel()."""
    r = out["reconstruction"]
    rl = F.mse_loss(r, targets) if recon_loss_type == "mse" else F.l1_loss(r, targets)
    mu, lv = out["mu"], out["logvar"]
    kl = -0.5 * torch.sum(1 + lv - mu.pow(2) - lv.exp()) / targets.numel()
    total = (recon_weight * rl + kl_weight * kl + separation_weight * out["separation_loss"]
             + contrastive_weight * out["contrastive_loss"])
    return {"loss": total, "recon_loss": rl, "kl_loss": kl,
            "separation_loss": out["separation_loss"], "contrastive_loss": out["contrastive_loss"]}


# --------------------------------------------------------------------------------------
# Training step (src/lightning_module.py:98-218, 390-477 without Lightning)
# --------------------------------------------------------------------------------------
def training_step_grads(sd: SD, x, eps, condition=None, loss_kwargs=None, taps=None):
    """fwd + VAELoss + bwd; returns (loss dict, outputs, {param: grad})."""
    loss_kwargs = loss_kwargs or {}
    params = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    out = vae_forward(params, x, eps, condition, taps=taps)
    ld = vae_loss(out["reconstruction"], x, out["mean"], out["logvar"], **loss_kwargs)
    names = list(params)
    grads = torch.autograd.grad(ld["loss"], [params[n] for n in names], allow_unused=True)
    g = {n: (gi if gi is not None else torch.zeros_like(params[n])) for n, gi in zip(names, grads)}
    return ({k: v.detach() for k, v in ld.items()}, {k: v.detach() for k, v in out.items()}, g)


def synthetic_batch(batch: int, channels: int, res: int, n_modalities: int = 12, seed: int = 42,
                    gray: Sequence[int] = (0, 3, 9, 10, 11)):
    """x = rand*2-1 under Generator(42); modality i mod n; gray modalities have
    channels 1,2 zeroed like mixed_modality_collate_fn (src/data/medmnist_data.py:16-72)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(batch, channels, res, res, generator=g) * 2 - 1
    idx = torch.arange(batch) % n_modalities
    onehot = F.one_hot(idx, n_modalities).float()
    if channels == 3:
        for m in gray:
            x[idx == m, 1:] = 0.0
    return x, onehot, idx


class ReferenceTrainer:
    """The reference's optimisation step without Lightning (src/lightning_module.py:98-218 forward +
    criterion, :468-477 NaN/Inf gradient zeroing, :452-466 clip_gradients(norm, 1.0), :390-406
    AdamW(lr, weight_decay) / Adam) on the functional model above.  Used by the loss-curve parity
    test and timed by bench.py as the CPU baseline ("port")."""

    def __init__(self, sd: SD, lr=1e-4, weight_decay=1e-4, betas=(0.9, 0.999), max_norm=1.0, decoupled=True,
                 loss_kwargs=None, disentangled=None):
        """disentangled = (shared_latent_dim, modality_latent_dim) switches to the
        DisentangledConditionalVAE forward + DisentangledVAELoss (``condition`` is then the int64
        modality index vector, lightning_module.py:116-118,191-196)."""
        self.params = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
        cls = torch.optim.AdamW if decoupled else torch.optim.Adam
        self.opt = cls(list(self.params.values()), lr=lr, weight_decay=weight_decay, betas=betas)
        self.max_norm = max_norm
        self.loss_kwargs = loss_kwargs or {}
        self.disentangled = disentangled

    def step(self, x, eps, condition=None):
        if self.disentangled is not None:
            out = dis_forward(self.params, x, condition, eps, *self.disentangled)
            ld = dis_loss(out, x, **self.loss_kwargs)
        else:
            out = vae_forward(self.params, x, eps, condition)
            ld = vae_loss(out["reconstruction"], x, out["mean"], out["logvar"], **self.loss_kwargs)
        self.opt.zero_grad(set_to_none=True)
        ld["loss"].backward()
        for p in self.params.values():      # :468-477 — ANY NaN / Inf zeroes the parameter's WHOLE gradient
            if p.grad is not None and not bool(torch.isfinite(p.grad).all()):
                p.grad.zero_()
        t0 = time.perf_counter()
        if self.max_norm and self.max_norm > 0:
            torch.nn.utils.clip_grad_norm_(list(self.params.values()), self.max_norm)
        self.opt.step()
        # seconds of the per-step (batch-size independent) part: clip + optimizer; bench.py reports it
        # next to the CPU baseline because a bounded CPU sample amortises it over few images
        self.last_update_s = time.perf_counter() - t0
        return {k: float(v.detach()) for k, v in ld.items()}
