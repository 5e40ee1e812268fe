"""medvae_disentangled_multimodal_b200 — B200-native (sm_100a) conditional-VAE training step.

Drop-in for the `src.models` / `src.losses` class surface of
parsakzr/medvae-disentangled-multimodal: same class names, constructor
kwargs, methods, output-dict keys and state_dict keys; the arithmetic runs in hand-written CUDA
kernels behind the C ABI in include/medvae_b200.h.  There is no CPU fallback.
"""
from .blocks import AttnBlock, Decoder, Downsample, Encoder, LinAttnBlock, LinearAttention, ResnetBlock, Upsample
from .losses import VAELoss
from .out_of_scope import BiomedCLIPLoss, LPIPSLoss, LPIPSWithDiscriminator, NLayerDiscriminator
from .vae import BaseVAE, BetaVAE, ConditionalVAE, FiLMLayer

__all__ = ["BaseVAE", "BetaVAE", "ConditionalVAE", "Encoder", "Decoder", "VAELoss", "ResnetBlock", "AttnBlock",
           "Downsample", "Upsample", "FiLMLayer", "LinearAttention", "LinAttnBlock",
           # exported by the reference next to the hot path; stubs that say so (out_of_scope.py)
           "LPIPSLoss", "BiomedCLIPLoss", "LPIPSWithDiscriminator", "NLayerDiscriminator"]

try:  # the disentangled model lives in its own module
    from .disentangled import DisentangledConditionalVAE, DisentangledVAELoss  # noqa: F401
    __all__ += ["DisentangledConditionalVAE", "DisentangledVAELoss"]
except ImportError:  # pragma: no cover - module not built yet
    pass
