"""DisentangledConditionalVAE / DisentangledVAELoss — drop-in for
src/models/disentangled_conditional_vae.py on the sm_100a kernels.

What the reference does with two per-sample Python loops (one ``.item()`` host sync and one or
three tiny convs per sample, :131-171 and :258-301) is batched here: one kernel selects the input
projector by modality index while packing the NHWC encoder input, and the per-modality decoder
heads are 3-channel stencils whose weights are indexed by modality inside the kernel.  The two
batch-coupled latent losses (:305-386) are three small kernels forward and one backward.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn
from torch.distributions import Normal

from . import _lib as L
from . import ops
from .blocks import _ToNHWC
from .losses import _KIND, _VAELossFn
from .vae import BaseVAE, _EncodeSampleFn

MODALITY_CHANNELS = {0: 1, 1: 3, 2: 3, 3: 1, 4: 3}   # disentangled_conditional_vae.py:112-122


class _DisInputFn(torch.autograd.Function):
    """x NCHW fp32 (zero-padded to 3 channels) + modality index -> NHWC encoder input; 1-channel
    modalities go through their Conv2d(1,3,1) projector (:148-158)."""

    @staticmethod
    def forward(ctx, x, idx, pw, pb, is_gray, out_c: int, dtype):
        ctx.save_for_backward(x, idx, is_gray)
        ctx.n_mod = pw.shape[0]
        return ops.dis_input_fwd(x, idx, pw.float().contiguous(), pb.float().contiguous(), is_gray, out_c, dtype)

    @staticmethod
    def backward(ctx, d_in):
        x, idx, is_gray = ctx.saved_tensors
        dpw, dpb = ops.dis_input_bwd(d_in.contiguous(), x, idx, is_gray, ctx.n_mod)
        return None, None, dpw, dpb, None, None, None


class _HeadFn(torch.autograd.Function):
    """decoder output NHWC -> per-modality heads -> NCHW fp32 reconstruction (:258-301)."""

    @staticmethod
    def forward(ctx, r, idx, w1, b1, w2, b2, pw, pb, is_gray, out_c: int):
        w1, b1, w2, b2, pw, pb = [t.float().contiguous() for t in (w1, b1, w2, b2, pw, pb)]
        out, hid, y = ops.dis_head_fwd(r, idx, w1, b1, w2, b2, pw, pb, is_gray, out_c)
        ctx.save_for_backward(r, hid, y, idx, w1, w2, pw, is_gray)
        return out

    @staticmethod
    def backward(ctx, d_out):
        r, hid, y, idx, w1, w2, pw, is_gray = ctx.saved_tensors
        d_r, g = ops.dis_head_bwd(d_out.float().contiguous(), r, hid, y, idx, w1, w2, pw, is_gray)
        n = w1.shape[0]
        return (d_r, None, g[:, 0:81].reshape(n, 3, 3, 3, 3), g[:, 81:84], g[:, 84:165].reshape(n, 3, 3, 3, 3),
                g[:, 165:168], g[:, 168:171], g[:, 171], None, None)


class _AuxLossFn(torch.autograd.Function):
    """(z NCHW fp32, idx) -> (modality_separation_loss, contrastive_loss), :305-386."""

    @staticmethod
    def forward(ctx, z, idx, shared: int, modal: int, n_mod: int, tau: float):
        zf = z.float().contiguous().view(z.shape[0], -1)
        aux = ops.dis_aux_fwd(zf, idx, shared, modal, n_mod, tau)
        ctx.save_for_backward(zf, idx, aux)
        ctx.cfg = (shared, modal, n_mod, tau, tuple(z.shape))
        return aux[0].clone(), aux[1].clone()

    @staticmethod
    def backward(ctx, g_sep, g_con):
        zf, idx, aux = ctx.saved_tensors
        shared, modal, n_mod, tau, shape = ctx.cfg
        c = lambda g: None if g is None else g.float().reshape(()).contiguous()
        dz = ops.dis_aux_bwd(zf, idx, aux, c(g_sep), c(g_con), shared, modal, n_mod, tau)
        return dz.view(shape), None, None, None, None, None


class DisentangledConditionalVAE(BaseVAE):
    """src/models/disentangled_conditional_vae.py:14-482."""

    def __init__(self, num_modalities: int = 5, shared_latent_dim: int = 8, modality_latent_dim: int = 8,
                 modality_separation_weight: float = 1.0, contrastive_weight: float = 0.5, resolution: int = 28,
                 hidden_channels: int = 128, ch_mult: Tuple[int, ...] = (1, 2, 4, 8), num_res_blocks: int = 2,
                 attn_resolutions: list = [16], dropout: float = 0.0, use_linear_attn: bool = False,
                 attn_type: str = "vanilla", **kwargs):
        self.modality_channels = dict(MODALITY_CHANNELS)
        max_channels = max(self.modality_channels.values())
        kwargs.update(dict(latent_dim=shared_latent_dim + modality_latent_dim, resolution=resolution,
                           input_channels=max_channels, hidden_channels=hidden_channels, ch_mult=ch_mult,
                           num_res_blocks=num_res_blocks, attn_resolutions=attn_resolutions, dropout=dropout,
                           use_linear_attn=use_linear_attn, attn_type=attn_type))
        super().__init__(**kwargs)
        self.num_modalities = num_modalities
        self.shared_latent_dim = shared_latent_dim
        self.modality_latent_dim = modality_latent_dim
        self.modality_separation_weight = modality_separation_weight
        self.contrastive_weight = contrastive_weight
        self.resolution = resolution
        # same construction order as the reference (:78-110) so a seeded init reproduces its weights
        self.modality_input_projectors = nn.ModuleDict()
        for m, ch in self.modality_channels.items():
            if ch != max_channels:
                self.modality_input_projectors[str(m)] = nn.Conv2d(ch, max_channels, kernel_size=1, padding=0)
        self.modality_output_projectors = nn.ModuleDict()
        for m, ch in self.modality_channels.items():
            if ch != max_channels:
                self.modality_output_projectors[str(m)] = nn.Conv2d(max_channels, ch, kernel_size=1, padding=0)
        self.modality_embedding = nn.Embedding(num_modalities, 64)   # unused by every forward path (:96)
        self.modality_decoders = nn.ModuleList([
            nn.Sequential(nn.Conv2d(max_channels, max_channels, 3, 1, 1), nn.ReLU(),
                          nn.Conv2d(max_channels, max_channels, 3, 1, 1)) for _ in range(num_modalities)])
        self._latent_clamp = 10.0    # forward clamps mu / logvar to [-10, 10] (:409-412)
        self._gray_cache: dict = {}

    # ------------------------------------------------------------------ parameter marshalling
    def _n_chan(self) -> int:
        return len(self.modality_channels)

    def _is_gray(self, n: int, device) -> torch.Tensor:
        """int32 [n]: entry m says whether modality min(m, 4) is a 1-channel one"""
        key = (n, str(device))
        t = self._gray_cache.get(key)
        if t is None:
            last = self._n_chan() - 1
            t = torch.tensor([int(self.modality_channels[min(m, last)] == 1) for m in range(n)], dtype=torch.int32,
                             device=device)
            self._gray_cache[key] = t
        return t

    def _input_proj_params(self, device):
        n = self._n_chan()
        ws, bs = [], []
        for m in range(n):
            p = self.modality_input_projectors[str(m)] if str(m) in self.modality_input_projectors else None
            ws.append(p.weight.reshape(3) if p is not None else torch.zeros(3, device=device))
            bs.append(p.bias.reshape(3) if p is not None else torch.zeros(3, device=device))
        return torch.stack(ws), torch.stack(bs)

    def _head_params(self, device):
        n = len(self.modality_decoders)
        w1 = torch.stack([d[0].weight for d in self.modality_decoders])
        b1 = torch.stack([d[0].bias for d in self.modality_decoders])
        w2 = torch.stack([d[2].weight for d in self.modality_decoders])
        b2 = torch.stack([d[2].bias for d in self.modality_decoders])
        pw, pb = [], []
        for m in range(n):
            key = str(m)   # the reference looks the projector up by the (clamped) decoder index (:274-278)
            p = self.modality_output_projectors[key] if key in self.modality_output_projectors else None
            pw.append(p.weight.reshape(3) if p is not None else torch.zeros(3, device=device))
            pb.append(p.bias.reshape(()) if p is not None else torch.zeros((), device=device))
        gray = torch.tensor([int(str(m) in self.modality_output_projectors) for m in range(n)], dtype=torch.int32,
                            device=device)
        return w1, b1, w2, b2, torch.stack(pw), torch.stack(pb), gray

    # ------------------------------------------------------------------ stages
    @staticmethod
    def _idx_pair(modality_indices: torch.Tensor, device):
        """-> (device int64 index tensor, host copy).  Pass a CPU tensor to avoid the one host sync
        the output-width decision (:283-301) otherwise costs."""
        if modality_indices.device.type == "cpu":
            host = modality_indices.to(torch.int64)
            return host.to(device, non_blocking=True), host
        dev = modality_indices.to(device=device, dtype=torch.int64)
        return dev, None

    def _encoder_input(self, x: torch.Tensor, condition=None) -> torch.Tensor:
        if condition is None:
            return super()._encoder_input(x)
        idx, _ = self._idx_pair(condition, x.device)
        pw, pb = self._input_proj_params(x.device)
        return _DisInputFn.apply(x.float().contiguous(), idx.contiguous(), pw, pb, self._is_gray(self._n_chan(), x.device),
                                 self.encoder.in_channels_nhwc(self.compute_dtype), self.compute_dtype)

    def encode(self, x: torch.Tensor, modality_indices: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """(:124-193; the NaN -> 0 scrubs are no-ops on finite data and are not reproduced)"""
        from .vae import ops_to_nchw
        h = ops_to_nchw(self._moments_nhwc(x, modality_indices))
        return torch.chunk(h, 2, dim=1)

    def partition_latent(self, z: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """:195-206 — slices of the FLATTENED (C, R, R) latent."""
        zf = z.reshape(z.shape[0], -1)
        s, m = self.shared_latent_dim, self.modality_latent_dim
        return zf[:, :s], zf[:, s:s + m]

    def reconstruct_latent(self, z_shared: torch.Tensor, z_modality: torch.Tensor) -> torch.Tensor:
        """:208-239."""
        remaining = self.latent_dim - self.shared_latent_dim - self.modality_latent_dim
        parts = [z_shared, z_modality]
        if remaining > 0:
            parts.append(torch.zeros(z_shared.shape[0], remaining, device=z_shared.device))
        z_full = torch.cat(parts, dim=1)
        r = self.encoder_out_res
        spatial = int((z_full.shape[1] / (r ** 2)) ** 0.5)
        if spatial ** 2 * (r ** 2) != z_full.shape[1]:
            return z_full.view(z_full.shape[0], -1, r, r)
        return z_full.view(z_full.shape[0], spatial, r, r)

    def decode(self, z: torch.Tensor, modality_indices: Optional[torch.Tensor] = None) -> torch.Tensor:
        if modality_indices is None:
            return super().decode(z)
        L.require_device()
        idx, host = self._idx_pair(modality_indices, z.device)
        w1, b1, w2, b2, pw, pb, gray = self._head_params(z.device)
        n = len(self.modality_decoders)
        if host is None:
            host = idx.cpu()     # the one host sync: is the whole batch 1-channel?
        keys = self.modality_output_projectors
        all_gray = all(str(min(int(m), n - 1)) in keys for m in host.tolist())
        out_c = 1 if all_gray else 3
        zin = _ToNHWC.apply(z.float().contiguous(), self.compute_dtype, z.shape[1])
        r = self.decoder.forward_nhwc(zin, z.requires_grad)
        return _HeadFn.apply(r, idx.contiguous(), w1, b1, w2, b2, pw, pb, gray, out_c)

    def _aux_losses(self, z, modality_indices, temperature: float = 0.1):
        L.require_device()
        idx, _ = self._idx_pair(modality_indices, z.device)
        return _AuxLossFn.apply(z, idx.contiguous(), self.shared_latent_dim, self.modality_latent_dim,
                                max(len(self.modality_decoders), self._n_chan()), float(temperature))

    def modality_separation_loss(self, z: torch.Tensor, modality_indices: torch.Tensor) -> torch.Tensor:
        return self._aux_losses(z, modality_indices)[0]

    def contrastive_loss(self, z: torch.Tensor, modality_indices: torch.Tensor, temperature: float = 0.1):
        return self._aux_losses(z, modality_indices, temperature)[1]

    def forward(self, x: torch.Tensor, modality_indices: torch.Tensor, return_latents: bool = False) -> Dict[str, torch.Tensor]:
        """:388-454."""
        mean, logvar, z, std, _eps, kl = self._encode_sample(x, modality_indices)
        reconstruction = self.decode(z, modality_indices)
        sep, con = self._aux_losses(z, modality_indices)
        zeros, ones = self._prior_constants(mean)
        prior = Normal(zeros, ones, validate_args=False)
        prior._mvae_std_normal = True      # losses._is_standard_normal
        posterior = Normal(mean, std, validate_args=False)    # std already clamped to [1e-6, 10] (:426-427)
        posterior._mvae_logvar = logvar
        posterior._mvae_kl_sum = kl
        out = {"reconstruction": reconstruction, "mean": mean, "logvar": logvar, "mu": mean, "z": z, "prior": prior,
               "posterior": posterior, "separation_loss": sep, "contrastive_loss": con}
        if return_latents:
            zs, zm = self.partition_latent(z)
            out.update({"z_shared": zs, "z_modality": zm})
        return out

    def sample_conditional(self, num_samples: int, modality_indices: torch.Tensor, device: torch.device) -> torch.Tensor:
        """:456-482 — N(0, 1) latents shifted by (modality - 2) * 0.3, decoded with the modality heads."""
        r = self.encoder_out_res
        z = torch.randn(num_samples, self.latent_dim, r, r, device=device)
        shift = (modality_indices.to(device).float() - 2.0) * 0.3
        z = z + shift.view(-1, 1, 1, 1)
        return self.decode(z, modality_indices)


class _TotalFn(torch.autograd.Function):
    """the guarded weighted total of DisentangledVAELoss (:527-565) on four device scalars."""

    @staticmethod
    def forward(ctx, recon, kl, sep, con, weights):
        terms = [t.float().reshape(()).contiguous() for t in (recon, kl, sep, con)]
        out5, mask4 = ops.dis_total(*terms, weights)
        ctx.save_for_backward(mask4)
        ctx.weights = weights
        return tuple(out5[i].clone() for i in range(5))

    @staticmethod
    def backward(ctx, g_total, *g_terms):
        (mask4,) = ctx.saved_tensors
        grads = []
        for i in range(4):
            g = mask4[i] * (g_total * ctx.weights[i] if g_total is not None else 0.0)
            if g_terms[i] is not None:
                g = g + mask4[i] * g_terms[i]
            grads.append(g)
        return (*grads, None)


class DisentangledVAELoss(nn.Module):
    """src/models/disentangled_conditional_vae.py:485-573."""

    def __init__(self, recon_loss_type: str = "mse", kl_weight: float = 1.0, recon_weight: float = 1.0,
                 separation_weight: float = 0.1, contrastive_weight: float = 0.05):
        super().__init__()
        if recon_loss_type not in ("mse", "l1"):
            raise ValueError(f"Unknown reconstruction loss: {recon_loss_type}")
        self.recon_loss_type = recon_loss_type
        self.kl_weight = kl_weight
        self.recon_weight = recon_weight
        self.separation_weight = separation_weight
        self.contrastive_weight = contrastive_weight

    def forward(self, outputs: Dict[str, torch.Tensor], targets: torch.Tensor) -> Dict[str, torch.Tensor]:
        L.require_device()
        recon, mu, logvar = outputs["reconstruction"], outputs["mu"], outputs["logvar"]
        if recon.shape != targets.shape:
            recon = recon.expand_as(targets)   # nn.MSELoss broadcasting (1-channel reconstruction of a padded batch)
        post = outputs.get("posterior")
        kl_sum = getattr(post, "_mvae_kl_sum", None) if post is not None and getattr(post, "_mvae_logvar", None) is logvar \
            else None
        # recon = mean over elements; kl = -0.5 * sum(1 + lv - mu^2 - e^lv) / targets.numel()  (:521-525)
        _loss, recon_loss, kl_loss = _VAELossFn.apply(recon, targets, mu, logvar, kl_sum, _KIND[self.recon_loss_type],
                                                      1.0, 1.0, float(targets.numel()))
        total, rl, kl, sep, con = _TotalFn.apply(recon_loss, kl_loss, outputs["separation_loss"],
                                                 outputs["contrastive_loss"],
                                                 (float(self.recon_weight), float(self.kl_weight),
                                                  float(self.separation_weight), float(self.contrastive_weight)))
        return {"loss": total, "recon_loss": rl, "kl_loss": kl, "separation_loss": sep, "contrastive_loss": con}
