"""Validation / metric step on the fused metric kernels — drop-in for src/utils/metrics.py:14-102
(`compute_reconstruction_metrics`, `compute_kl_metrics`, `compute_latent_metrics`) and for the metric
half of `VAELightningModule.validation_step` (src/lightning_module.py:220-310).

The reference pays one ``.item()`` host synchronisation per metric (4 + 4 per validation batch) and
runs torchmetrics' PSNR / SSIM as separate multi-kernel pipelines.  Here MSE, MAE, PSNR and SSIM come
out of ONE pass over (original, reconstruction), the KL statistics out of one pass over
(mean, logvar), everything lands in one 11-float device vector, and ``validation_metrics`` reads it
back once.  There is no CPU fallback.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from . import _lib as L
from . import ops

RECON_KEYS = ("mse", "mae", "psnr", "ssim")
KL_KEYS = ("kl_total", "kl_mean", "kl_std", "kl_per_dim_mean")
LATENT_KEYS = ("latent_mean_avg", "latent_std_avg", "latent_sparsity")


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t.detach().float().contiguous()


def reconstruction_metrics_device(original: torch.Tensor, reconstructed: torch.Tensor, data_range: float = 1.0,
                                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """-> fp32 device tensor (mse, mae, psnr, ssim); no host synchronisation."""
    L.require_device()
    if original.shape != reconstructed.shape or original.dim() != 4:
        raise RuntimeError(f"expected two (B, C, H, W) tensors of the same shape, got {tuple(original.shape)} "
                           f"and {tuple(reconstructed.shape)}")
    out = torch.empty(4, dtype=torch.float32, device=original.device) if out is None else out
    ops.recon_metrics(_f32c(original), _f32c(reconstructed), data_range, out)
    return out


def kl_metrics_device(mean: torch.Tensor, logvar: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """-> fp32 device tensor (kl_total, kl_mean, kl_std, kl_per_dim_mean)."""
    L.require_device()
    if mean.shape != logvar.shape or mean.dim() < 2:
        raise RuntimeError("mean / logvar must have the same shape (B, D, ...)")
    out = torch.empty(4, dtype=torch.float32, device=mean.device) if out is None else out
    ops.kl_metrics(_f32c(mean), _f32c(logvar), out)
    return out


def latent_metrics_device(latents: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """-> fp32 device tensor (latent_mean_avg, latent_std_avg, latent_sparsity)."""
    L.require_device()
    z = _f32c(latents)
    if z.dim() > 2:
        z = z.flatten(1)
    out = torch.empty(3, dtype=torch.float32, device=z.device) if out is None else out
    ops.latent_metrics(z, 0.1, out)
    return out


def compute_reconstruction_metrics(original: torch.Tensor, reconstructed: torch.Tensor) -> Dict[str, float]:
    """src/utils/metrics.py:14-46 (one device->host read instead of four)."""
    return dict(zip(RECON_KEYS, reconstruction_metrics_device(original, reconstructed).tolist()))


def compute_kl_metrics(mean: torch.Tensor, logvar: torch.Tensor) -> Dict[str, float]:
    """src/utils/metrics.py:49-75."""
    return dict(zip(KL_KEYS, kl_metrics_device(mean, logvar).tolist()))


def compute_latent_metrics(latents: torch.Tensor) -> Dict[str, float]:
    """src/utils/metrics.py:78-102."""
    return dict(zip(LATENT_KEYS, latent_metrics_device(latents).tolist()))


def validation_metrics_device(x: torch.Tensor, outputs: Dict[str, torch.Tensor], with_latent: bool = False) -> torch.Tensor:
    """the 8 (or 11) numbers validation_step logs (lightning_module.py:248-257), as one device vector
    in the order RECON_KEYS + KL_KEYS (+ LATENT_KEYS)."""
    n = 11 if with_latent else 8
    out = torch.empty(n, dtype=torch.float32, device=x.device)
    recon = outputs["reconstruction"]
    if recon.shape != x.shape:
        recon = recon.expand_as(x)
    reconstruction_metrics_device(x, recon, 1.0, out[0:4])
    kl_metrics_device(outputs["mean"], outputs["logvar"], out[4:8])
    if with_latent:
        latent_metrics_device(outputs["z"], out[8:11])
    return out


def validation_metrics(x: torch.Tensor, outputs: Dict[str, torch.Tensor], with_latent: bool = False) -> Dict[str, float]:
    """-> {"val/mse": ..., ..., "val/kl_per_dim_mean": ...} with ONE device->host read."""
    keys = RECON_KEYS + KL_KEYS + (LATENT_KEYS if with_latent else ())
    vals = validation_metrics_device(x, outputs, with_latent).tolist()
    return {f"val/{k}": v for k, v in zip(keys, vals)}
