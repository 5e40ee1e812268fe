"""src/models/disentangled_conditional_vae.py:14-573 (imported by submodule path at src/lightning_module.py:14,
quick_test.py, test_modality_channels.py, basic_test.py)"""
from medvae_disentangled_multimodal_b200 import DisentangledConditionalVAE, DisentangledVAELoss  # noqa: F401
