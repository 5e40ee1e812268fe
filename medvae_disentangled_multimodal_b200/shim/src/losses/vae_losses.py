"""src/losses/vae_losses.py:17-330"""
from medvae_disentangled_multimodal_b200 import BiomedCLIPLoss, LPIPSLoss, LPIPSWithDiscriminator, VAELoss  # noqa: F401
