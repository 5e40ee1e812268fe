"""src/models/encoder_decoder.py:13-451"""
from medvae_disentangled_multimodal_b200 import (AttnBlock, Decoder, Downsample, Encoder, LinAttnBlock,  # noqa: F401
                                                 LinearAttention, ResnetBlock, Upsample)
