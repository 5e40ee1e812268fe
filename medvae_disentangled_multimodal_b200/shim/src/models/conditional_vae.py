"""src/models/conditional_vae.py:14-203"""
from medvae_disentangled_multimodal_b200 import ConditionalVAE, FiLMLayer  # noqa: F401
