"""src/models/beta_vae.py:13-43"""
from medvae_disentangled_multimodal_b200 import BetaVAE  # noqa: F401
