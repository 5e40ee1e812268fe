"""Loss functions package."""

from .vae_losses import VAELoss, LPIPSLoss, BiomedCLIPLoss, LPIPSWithDiscriminator

__all__ = [
    "VAELoss",
    "LPIPSLoss",
    "BiomedCLIPLoss",
    "LPIPSWithDiscriminator",
]
