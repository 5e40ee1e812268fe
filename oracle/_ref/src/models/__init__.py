"""Medical VAE models package."""

from .base_vae import BaseVAE
from .beta_vae import BetaVAE
from .conditional_vae import ConditionalVAE
from .disentangled_conditional_vae import (
    DisentangledConditionalVAE,
    DisentangledVAELoss,
)
from .encoder_decoder import Encoder, Decoder
from .discriminator import NLayerDiscriminator

__all__ = [
    "BaseVAE",
    "BetaVAE",
    "ConditionalVAE",
    "DisentangledConditionalVAE",
    "DisentangledVAELoss",
    "Encoder",
    "Decoder",
    "NLayerDiscriminator",
]
