"""`src.losses` (src/losses/__init__.py:3-10) served by the B200 path; the perceptual / adversarial losses are stubs
that raise on construction (out_of_scope.py)."""
from medvae_disentangled_multimodal_b200 import BiomedCLIPLoss, LPIPSLoss, LPIPSWithDiscriminator, VAELoss

__all__ = ["VAELoss", "LPIPSLoss", "BiomedCLIPLoss", "LPIPSWithDiscriminator"]
