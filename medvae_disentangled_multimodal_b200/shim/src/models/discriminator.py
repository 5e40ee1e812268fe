"""src/models/discriminator.py:11-70 — outside the hot path: a stub that says so when constructed"""
from medvae_disentangled_multimodal_b200 import NLayerDiscriminator  # noqa: F401
