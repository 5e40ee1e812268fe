"""Names the reference exports next to the hot path that this package does NOT implement
("re-export / stub only"): the perceptual / adversarial losses and the discriminator
(src/losses/vae_losses.py:67-330, src/models/discriminator.py:11-70).  They need pretrained LPIPS /
BiomedCLIP weights or a BatchNorm discriminator that the training step of BASELINE.json never builds.
Constructing one raises with that explanation instead of an ImportError, so a Hydra target swap
 fails loudly and specifically."""
from __future__ import annotations

import torch.nn as nn


class _OutOfScope(nn.Module):
    _what = ""

    def __init__(self, *args, **kwargs):
        super().__init__()
        raise NotImplementedError(
            f"{type(self).__name__} ({self._what}) is outside the B200 hot path (VAE / Beta-VAE / conditional / "
            "disentangled training step with VAELoss / DisentangledVAELoss); use the reference's own implementation "
            "for it — it composes with these models because they keep the reference's tensor API")


class LPIPSLoss(_OutOfScope):
    _what = "src/losses/vae_losses.py:67-94, needs the lpips package and its pretrained weights"


class BiomedCLIPLoss(_OutOfScope):
    _what = "src/losses/vae_losses.py:97-186, needs open_clip and the BiomedCLIP checkpoint"


class LPIPSWithDiscriminator(_OutOfScope):
    _what = "src/losses/vae_losses.py:214-330, adversarial + perceptual loss with manual optimisation"


class NLayerDiscriminator(_OutOfScope):
    _what = "src/models/discriminator.py:11-70, PatchGAN discriminator with BatchNorm"
