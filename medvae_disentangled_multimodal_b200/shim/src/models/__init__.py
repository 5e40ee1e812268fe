"""`src.models` of parsakzr/medvae-disentangled-multimodal (src/models/__init__.py:3-22), served by the B200 path.

This directory is a PORTION of the reference's namespace package `src` (the reference ships no src/__init__.py): put
`medvae_disentangled_multimodal_b200/shim` on sys.path AHEAD of the reference root and `src.models` / `src.losses`
resolve here while `src.lightning_module`, `src.data` and `src.utils` still resolve to the reference's own files —
no edit to the reference.  `python -m medvae_disentangled_multimodal_b200.launch main.py ...` does exactly that."""
from medvae_disentangled_multimodal_b200 import (BaseVAE, BetaVAE, ConditionalVAE, Decoder,
                                                 DisentangledConditionalVAE, DisentangledVAELoss, Encoder,
                                                 NLayerDiscriminator)

__all__ = ["BaseVAE", "BetaVAE", "ConditionalVAE", "DisentangledConditionalVAE", "DisentangledVAELoss", "Encoder",
           "Decoder", "NLayerDiscriminator"]
