"""CPU oracle of the validation / metric step — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this file; the
product path (medvae_disentangled_multimodal_b200.metrics) runs CUDA kernels and never falls back here.

What it restates, each on plain torch CPU ops:

* ``compute_reconstruction_metrics``  — src/utils/metrics.py:14-46: MSE, MAE and the two
  torchmetrics functionals the reference calls with ``data_range=1.0``.
* ``compute_kl_metrics``              — src/utils/metrics.py:49-75 (note ``sum(dim=1)``: with the
  4-D mean/logvar the validation step passes, a "sample" is one (b, y, x) position).
* ``compute_latent_metrics``          — src/utils/metrics.py:78-102.

Third-party arithmetic: ``torchmetrics`` (pyproject.toml: ``torchmetrics>=1.0.0``; not vendored under
/root/reference and not installed in this image, no network).  Its published algorithm is restated:

  peak_signal_noise_ratio(preds, target, data_range=r), base 10, whole-tensor reduction:
      10 * log10(r^2 / mean((preds - target)^2))
  structural_similarity_index_measure(preds, target, data_range=r), defaults gaussian_kernel=True,
  sigma=1.5, kernel_size=11, k1=0.01, k2=0.03, reduction="elementwise_mean":
      g = normalised 11-tap gaussian (sigma 1.5), window = g^T g per channel (depthwise);
      inputs reflect-padded by 5, windowed means of p, t, p^2, t^2, p*t;
      var_p = max(E[p^2] - mu_p^2, 0), var_t likewise, cov = E[pt] - mu_p mu_t;
      ssim map = ((2 mu_p mu_t + c1)(2 cov + c2)) / ((mu_p^2 + mu_t^2 + c1)(var_p + var_t + c2)),
      c1 = (k1 r)^2, c2 = (k2 r)^2; the map is CROPPED by 5 on every side (so the reflect padding
      never reaches the result), averaged per image, then over the batch.

Parity status: the KL / latent functions are pinned against the reference's own code (imported with
torchmetrics / sklearn stubbed, oracle/make_golden_metrics.py -> tests/golden/metrics.npz).  PSNR and
SSIM are pinned only by the two known-answer examples torchmetrics publishes in its docstrings
(tests/test_metrics_oracle.py) and by closed forms (identical / constant images): **parity unpinned**
against a live torchmetrics.
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F


def gaussian_1d(kernel_size: int = 11, sigma: float = 1.5, dtype=torch.float32) -> torch.Tensor:
    dist = torch.arange((1 - kernel_size) / 2, (1 + kernel_size) / 2, step=1, dtype=dtype)
    g = torch.exp(-torch.pow(dist / sigma, 2) / 2)
    return g / g.sum()


def psnr(preds: torch.Tensor, target: torch.Tensor, data_range: float = 1.0) -> torch.Tensor:
    mse = torch.mean((preds - target) ** 2)
    return 10.0 * torch.log10(torch.tensor(float(data_range) ** 2, dtype=mse.dtype) / mse)


def ssim(preds: torch.Tensor, target: torch.Tensor, data_range: float = 1.0, sigma: float = 1.5,
         kernel_size: int = 11, k1: float = 0.01, k2: float = 0.03) -> torch.Tensor:
    b, c, h, w = preds.shape
    pad = (kernel_size - 1) // 2
    g = gaussian_1d(kernel_size, sigma, preds.dtype)
    window = (g[:, None] * g[None, :]).expand(c, 1, kernel_size, kernel_size)
    c1, c2 = (k1 * data_range) ** 2, (k2 * data_range) ** 2
    p = F.pad(preds, (pad, pad, pad, pad), mode="reflect")
    t = F.pad(target, (pad, pad, pad, pad), mode="reflect")
    stack = torch.cat([p, t, p * p, t * t, p * t])
    out = F.conv2d(stack, window, groups=c)
    mu_p, mu_t, e_pp, e_tt, e_pt = out.split(b)
    var_p = torch.clamp(e_pp - mu_p * mu_p, min=0.0)
    var_t = torch.clamp(e_tt - mu_t * mu_t, min=0.0)
    cov = e_pt - mu_p * mu_t
    m = ((2 * mu_p * mu_t + c1) * (2 * cov + c2)) / ((mu_p * mu_p + mu_t * mu_t + c1) * (var_p + var_t + c2))
    m = m[..., pad:-pad, pad:-pad]
    return m.reshape(b, -1).mean(-1).mean()


def compute_reconstruction_metrics(original: torch.Tensor, reconstructed: torch.Tensor) -> Dict[str, float]:
    """src/utils/metrics.py:14-46"""
    return {
        "mse": float(F.mse_loss(reconstructed, original)),
        "mae": float(F.l1_loss(reconstructed, original)),
        "psnr": float(psnr(reconstructed, original, 1.0)),
        "ssim": float(ssim(reconstructed, original, 1.0)),
    }


def compute_kl_metrics(mean: torch.Tensor, logvar: torch.Tensor) -> Dict[str, float]:
    """src/utils/metrics.py:49-75"""
    kl = 0.5 * (mean.pow(2) + logvar.exp() - logvar - 1)
    per_sample = kl.sum(dim=1)
    return {"kl_total": float(kl.sum()), "kl_mean": float(per_sample.mean()), "kl_std": float(per_sample.std()),
            "kl_per_dim_mean": float(kl.mean(dim=0).mean())}


def compute_latent_metrics(latents: torch.Tensor) -> Dict[str, float]:
    """src/utils/metrics.py:78-102"""
    if latents.dim() > 2:
        latents = latents.flatten(1)
    return {"latent_mean_avg": float(latents.mean(dim=0).mean()), "latent_std_avg": float(latents.std(dim=0).mean()),
            "latent_sparsity": float((latents.abs() < 0.1).float().mean())}
