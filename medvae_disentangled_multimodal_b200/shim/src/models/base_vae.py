"""src/models/base_vae.py:14-153"""
from medvae_disentangled_multimodal_b200 import BaseVAE  # noqa: F401
