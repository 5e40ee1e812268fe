"""BaseVAE / BetaVAE / ConditionalVAE — drop-in for src/models/{base_vae,beta_vae,conditional_vae}.py.

Same constructor kwargs, methods, output-dict keys and state_dict keys as the reference.  The
arithmetic runs in the sm_100a kernels (bf16 storage / fp32 accumulate by default, or the fp32
verification mode: ``compute_dtype="fp32"`` or env MVAE_PRECISION=fp32).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn
from torch.distributions import Normal

from . import _lib as L
from . import ops
from .blocks import Decoder, Encoder


def _resolve_dtype(compute_dtype) -> torch.dtype:
    v = compute_dtype if compute_dtype is not None else os.environ.get("MVAE_PRECISION", "bf16")
    if isinstance(v, torch.dtype):
        if v in (torch.float32, torch.bfloat16):
            return v
        raise ValueError(f"compute_dtype must be bf16 or fp32, got {v}")
    v = str(v).lower()
    if v in ("bf16", "bfloat16"):
        return torch.bfloat16
    if v in ("fp32", "float32", "f32"):
        return torch.float32
    raise ValueError(f"compute_dtype must be 'bf16' or 'fp32', got {v!r}")


class _EncodeSampleFn(torch.autograd.Function):
    """encoder -> chunk -> reparameterise (+ KL partial sum), fused on the NHWC moments.
    Inputs: packed NHWC encoder input (already differentiable), eps or None."""

    @staticmethod
    def forward(ctx, model: "BaseVAE", moments_nhwc, eps, clamp: float):
        seed, offset = model._next_rng()
        mean, logvar, z, std, eps_used, kl = ops.reparam_kl_fwd(
            moments_nhwc, eps, model.latent_dim, clamp, seed, offset, want_kl=True, want_std=True,
            step_state=model._step_state, sample_base=model._rng_sample_base)
        ctx.save_for_backward(moments_nhwc, logvar, eps_used, std)
        ctx.clamp = clamp
        ctx.mark_non_differentiable(eps_used, kl)
        return mean, logvar, z, std, eps_used, kl

    @staticmethod
    def backward(ctx, d_mean, d_logvar, d_z, d_std, _d_eps, _d_kl):
        moments, logvar, eps, std = ctx.saved_tensors
        c = lambda t: None if t is None else t.float().contiguous()
        if d_std is not None:
            # only reached when a foreign loss differentiates posterior.scale (e.g. torch's own
            # kl_divergence); VAELoss here uses the raw logvar instead.
            extra = 0.5 * d_std * std
            d_logvar = extra if d_logvar is None else d_logvar + extra
        d_mom = ops.reparam_kl_bwd(c(d_mean), c(d_logvar), c(d_z), logvar, eps, moments, ctx.clamp)
        return None, d_mom, None, None


class _ReparamFn(torch.autograd.Function):
    """standalone BaseVAE.reparameterize on NCHW fp32 mean / logvar (base_vae.py:83-87)."""

    @staticmethod
    def forward(ctx, model: "BaseVAE", mean, logvar, eps):
        # reuse the fused kernel: build NHWC moments from the two NCHW tensors
        mom = ops.nchw_f32_to_nhwc(torch.cat([mean, logvar], dim=1).float().contiguous(), torch.float32)
        seed, offset = model._next_rng()
        _m, _lv, z, _sd, eps_used, _kl = ops.reparam_kl_fwd(mom, eps, mean.shape[1], 0.0, seed, offset,
                                                            want_kl=False, want_std=False,
                                                            sample_base=model._rng_sample_base)
        ctx.save_for_backward(mom, logvar.float().contiguous(), eps_used)
        return z

    @staticmethod
    def backward(ctx, d_z):
        mom, logvar, eps = ctx.saved_tensors
        d_mom = ops.reparam_kl_bwd(None, None, d_z.float().contiguous(), logvar, eps, mom, 0.0)
        d = ops.nhwc_to_nchw_f32(d_mom)
        zc = d.shape[1] // 2
        return None, d[:, :zc], d[:, zc:], None


class _InputPackFn(torch.autograd.Function):
    """NCHW fp32 image (+ modality condition map, conditional_vae.py:107-127) -> NHWC encoder input."""

    @staticmethod
    def forward(ctx, x, cond, w, b, out_c: int, dtype):
        ctx.save_for_backward(cond, w, b)
        ctx.c = x.shape[1]
        return ops.input_pack_fwd(x, cond, w, b, out_c, dtype)

    @staticmethod
    def backward(ctx, d_in):
        cond, w, b = ctx.saved_tensors
        if w is None:
            return None, None, None, None, None, None
        dw, db = ops.input_pack_bwd(d_in.contiguous(), cond, w, b, ctx.c)
        return None, None, dw, db, None, None


class BaseVAE(nn.Module):
    """src/models/base_vae.py:14-153."""

    def __init__(self, input_channels: int = 1, latent_dim: int = 128, hidden_channels: int = 128,
                 ch_mult: Tuple[int, ...] = (1, 2, 4, 8), num_res_blocks: int = 2, attn_resolutions: list = [16],
                 dropout: float = 0.0, resolution: int = 224, use_linear_attn: bool = False,
                 attn_type: str = "vanilla", double_z: bool = True, compute_dtype=None, **kwargs):
        super().__init__()
        self.latent_dim = latent_dim
        self.input_channels = input_channels
        self.encoder_out_res = resolution // (2 ** (len(ch_mult) - 1))
        common = dict(ch=hidden_channels, out_ch=input_channels, ch_mult=tuple(ch_mult), num_res_blocks=num_res_blocks,
                      attn_resolutions=list(attn_resolutions), dropout=dropout, resamp_with_conv=True,
                      in_channels=input_channels, resolution=resolution, z_channels=latent_dim,
                      use_linear_attn=use_linear_attn, attn_type=attn_type)
        self.encoder = Encoder(double_z=double_z, **common)
        self.decoder = Decoder(**common)
        self.set_compute_dtype(compute_dtype)
        # verification hooks: externally supplied reparameterisation noise, Philox stream position
        self.eps_override: Optional[torch.Tensor] = None
        self.rng_seed: Optional[int] = None
        self._rng_step = 0
        # first sample of the current forward in the GLOBAL batch (TrainStep: rank * shard + micro-batch start): the
        # Philox element counter starts there, so a sample's noise does not depend on the split
        self._rng_sample_base = 0
        self._latent_clamp = 0.0
        self._prior_cache: dict = {}
        self._step_state: Optional[torch.Tensor] = None   # device-side Philox position (TrainStep graph mode)

    # ------------------------------------------------------------------ runtime knobs
    def set_compute_dtype(self, compute_dtype):
        dt = _resolve_dtype(compute_dtype)
        self.compute_dtype = dt
        self.encoder.compute_dtype = dt
        self.decoder.compute_dtype = dt
        return self

    def _next_rng(self):
        seed = self.rng_seed if self.rng_seed is not None else torch.initial_seed()
        self._rng_step += 1
        return seed & 0xFFFFFFFFFFFFFFFF, self._rng_step

    # ------------------------------------------------------------------ stages
    def _encoder_input(self, x: torch.Tensor, condition=None) -> torch.Tensor:
        return _InputPackFn.apply(x.float().contiguous(), None, None, None,
                                  self.encoder.in_channels_nhwc(self.compute_dtype), self.compute_dtype)

    def _moments_nhwc(self, x, condition=None):
        L.require_device()
        xin = self._encoder_input(x, condition)
        return self.encoder.forward_nhwc(xin, xin.requires_grad)

    def encode(self, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        h = ops_to_nchw(self._moments_nhwc(x))
        mean, logvar = torch.chunk(h, 2, dim=1)
        return mean, logvar

    def decode(self, z: torch.Tensor) -> torch.Tensor:
        return self.decoder(z)

    def reparameterize(self, mean: torch.Tensor, logvar: torch.Tensor, eps: Optional[torch.Tensor] = None):
        L.require_device()
        eps = eps if eps is not None else self.eps_override
        if eps is not None:
            eps = eps.float().contiguous()
        return _ReparamFn.apply(self, mean, logvar, eps)

    def _encode_sample(self, x, condition=None):
        mom = self._moments_nhwc(x, condition)
        eps = self.eps_override
        if eps is not None:
            if eps.shape[0] != mom.shape[0] or eps.numel() * 2 != mom.numel():
                raise ValueError(f"eps_override has shape {tuple(eps.shape)} but the latent of this batch is "
                                 f"({mom.shape[0]}, {mom.shape[3] // 2}, {mom.shape[1]}, {mom.shape[2]})")
            eps = eps.to(mom.device).float().contiguous()
        return _EncodeSampleFn.apply(self, mom, eps, float(self._latent_clamp))

    def _prior_constants(self, like: torch.Tensor):
        key = (tuple(like.shape), like.device)
        ent = self._prior_cache.get(key)
        if ent is None:
            ent = (torch.zeros_like(like), torch.ones_like(like))
            self._prior_cache = {key: ent}
        return ent

    def _make_outputs(self, reconstruction, mean, logvar, z, std, kl, return_latents, extra=None):
        zeros, ones = self._prior_constants(mean)
        prior = Normal(zeros, ones, validate_args=False)
        prior._mvae_std_normal = True      # losses._is_standard_normal
        posterior = Normal(mean, std, validate_args=False)
        # side channel for VAELoss: the raw logvar and the KL partial sum the fused kernel produced
        posterior._mvae_logvar = logvar
        posterior._mvae_kl_sum = kl
        outputs = {"reconstruction": reconstruction, "mean": mean, "logvar": logvar, "z": z,
                   "prior": prior, "posterior": posterior}
        if extra:
            outputs.update(extra)
        if return_latents:
            outputs["latents"] = z
        return outputs

    def forward(self, x: torch.Tensor, return_latents: bool = False) -> Dict[str, torch.Tensor]:
        mean, logvar, z, std, _eps, kl = self._encode_sample(x)
        reconstruction = self.decode(z)
        return self._make_outputs(reconstruction, mean, logvar, z, std, kl, return_latents)

    def sample(self, num_samples: int, device: torch.device) -> torch.Tensor:
        z = torch.randn(num_samples, self.latent_dim, self.encoder_out_res, self.encoder_out_res, device=device)
        return self.decode(z)

    def compute_loss(self, x, reconstruction, prior, posterior, **kwargs) -> Dict[str, torch.Tensor]:
        """base_vae.py:131-153 (unused by the training loop, kept for API parity)."""
        from .losses import VAELoss
        return VAELoss("mse", 1.0, 1.0)(inputs=x, reconstructions=reconstruction, posteriors=posterior, priors=prior)


def ops_to_nchw(y_nhwc):
    from .blocks import _ToNCHW
    return _ToNCHW.apply(y_nhwc, y_nhwc.shape[-1])


class BetaVAE(BaseVAE):
    """src/models/beta_vae.py:13-43.  `beta` is stored; training applies it through
    VAELoss(kl_weight=...) exactly as the reference's Lightning module does."""

    def __init__(self, beta: float = 1.0, **kwargs):
        super().__init__(**kwargs)
        self.beta = beta

    def compute_loss(self, x, reconstruction, prior, posterior, **kwargs):
        from .losses import VAELoss
        out = VAELoss("mse", self.beta, 1.0)(inputs=x, reconstructions=reconstruction, posteriors=posterior,
                                             priors=prior)
        out["weighted_kl_loss"] = self.beta * out["kl_loss"]
        return out


DEFAULT_MODALITIES = ["chest_xray", "pathology", "oct", "pneumonia", "dermatoscope", "blood_cell", "tissue",
                      "retina", "breast_ultrasound", "abdominal_ct_a", "abdominal_ct_c", "abdominal_ct_s"]


class _FiLMFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, features, cond, ws, bs, wt, bt):
        out, st = ops.film_fwd(features, cond, ws, bs, wt, bt)
        ctx.save_for_backward(features, cond, ws, wt, st)
        return out

    @staticmethod
    def backward(ctx, d_out):
        features, cond, ws, wt, st = ctx.saved_tensors
        df, dws, dbs, dwt, dbt, dcond = ops.film_bwd(d_out.float().contiguous(), features, cond, ws, wt, st,
                                                     ctx.needs_input_grad[1])
        return df, dcond, dws, dbs, dwt, dbt


class FiLMLayer(nn.Module):
    """conditional_vae.py:191-203: features * scale(c)[:, :, None, None] + shift(c)[:, :, None, None].
    Built when condition_method="film"; like the reference, no model forward path calls it — the
    op itself (one kernel forward, two backward) is provided for callers that do."""

    def __init__(self, condition_dim: int, feature_dim: int):
        super().__init__()
        self.scale_transform = nn.Linear(condition_dim, feature_dim)
        self.shift_transform = nn.Linear(condition_dim, feature_dim)

    def forward(self, features: torch.Tensor, condition: torch.Tensor) -> torch.Tensor:
        L.require_device()
        if features.dim() != 4 or condition.dim() != 2 or features.shape[1] != self.scale_transform.out_features:
            raise RuntimeError(f"FiLMLayer expects features (B, {self.scale_transform.out_features}, H, W) and condition "
                               f"(B, {self.scale_transform.in_features}), got {tuple(features.shape)} / {tuple(condition.shape)}")
        return _FiLMFn.apply(features.float().contiguous(), condition.float().contiguous(),
                             self.scale_transform.weight, self.scale_transform.bias,
                             self.shift_transform.weight, self.shift_transform.bias)


class ConditionalVAE(BaseVAE):
    """src/models/conditional_vae.py:14-188."""

    def __init__(self, modalities: List[str] = None, condition_dim: int = None, condition_method: str = "concat",
                 **kwargs):
        super().__init__(**kwargs)
        self.modalities = list(DEFAULT_MODALITIES) if modalities is None else modalities
        self.num_modalities = len(self.modalities)
        self.condition_dim = condition_dim or self.num_modalities
        self.condition_method = condition_method
        if condition_method == "concat":
            self.condition_proj = nn.Sequential(
                nn.Linear(self.condition_dim, self.input_channels * 8 * 8), nn.ReLU(),
                nn.Unflatten(1, (self.input_channels, 8, 8)))
            self.encoder.conv_in = nn.Conv2d(self.input_channels * 2, self.encoder.ch, 3, 1, 1)
        elif condition_method == "inject":
            self.condition_embedding = nn.Sequential(nn.Linear(self.condition_dim, 512), nn.ReLU(),
                                                     nn.Linear(512, 512))
        elif condition_method == "film":
            self.film_layers = nn.ModuleList(
                [FiLMLayer(self.condition_dim, self.encoder.ch * (2 ** i)) for i in range(len(self.encoder.down))])

    def encode_condition(self, condition: torch.Tensor) -> torch.Tensor:
        return condition.unsqueeze(0) if condition.dim() == 1 else condition

    def _encoder_input(self, x, condition=None):
        if self.condition_method != "concat" or condition is None:
            return super()._encoder_input(x)
        lin = self.condition_proj[0]
        cond = self.encode_condition(condition).to(x.device).float().contiguous()
        return _InputPackFn.apply(x.float().contiguous(), cond, lin.weight, lin.bias,
                                  self.encoder.in_channels_nhwc(self.compute_dtype), self.compute_dtype)

    def create_condition_map(self, condition: torch.Tensor, height: int, width: int) -> torch.Tensor:
        """(B, C, H, W) condition map, conditional_vae.py:107-119 (computed by the same kernel
        that feeds the encoder: channels [C, 2C) of its output)."""
        L.require_device()
        cond = self.encode_condition(condition).float().contiguous()
        dummy = torch.zeros(cond.shape[0], self.input_channels, height, width, device=cond.device)
        packed = self._encoder_input(dummy, cond)
        from .blocks import _ToNCHW
        return _ToNCHW.apply(packed, 2 * self.input_channels)[:, self.input_channels:]

    def encode(self, x: torch.Tensor, condition: torch.Tensor) -> tuple:
        h = ops_to_nchw(self._moments_nhwc(x, condition))
        return torch.chunk(h, 2, dim=1)

    def forward(self, x: torch.Tensor, condition: torch.Tensor, return_latents: bool = False):
        mean, logvar, z, std, _eps, kl = self._encode_sample(x, condition)
        reconstruction = self.decode(z)
        return self._make_outputs(reconstruction, mean, logvar, z, std, kl, return_latents,
                                  {"condition": condition})

    def conditional_sample(self, num_samples: int, condition: torch.Tensor, device: torch.device) -> torch.Tensor:
        z = torch.randn(num_samples, self.latent_dim, self.encoder_out_res, self.encoder_out_res, device=device)
        return self.decode(z)

    def get_modality_condition(self, modality: str) -> torch.Tensor:
        if modality not in self.modalities:
            raise ValueError(f"Unknown modality: {modality}")
        condition = torch.zeros(self.num_modalities)
        condition[self.modalities.index(modality)] = 1.0
        return condition
