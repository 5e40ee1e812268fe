"""CPU oracle of the batch feed and of FiLMLayer — TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench cpu_baseline).

Restates, on plain torch CPU ops, what the reference does on the host for every sample and batch:
  * ``getitem_image``  — MedMNISTDataset.__getitem__, src/data/medmnist_data.py:205-225 (ToTensor, channel
                          conversion to the modality's channel count) + Normalize(0.5, 0.5), :372-380
  * ``collate``        — mixed_modality_collate_fn, :16-72 (zero-pad to the widest sample)
  * ``film``           — FiLMLayer.forward, src/models/conditional_vae.py:197-203
Parity: PINNED — tests/golden/feed.npz and film.npz were produced by the reference's own code
(oracle/make_golden_widen.py); tests/test_oracle_widen.py checks these functions against them bit for bit (feed)
/ to 1e-6 (film).
"""
from __future__ import annotations

from typing import Sequence

import torch
import torch.nn.functional as F

GRAY_DATASETS_12 = (0, 3, 9, 10, 11)   # chest, pneumonia, organ{a,c,s}mnist (:157-163)


def getitem_image(img_u8_hwc: torch.Tensor, src_channels: int, target_channels: int) -> torch.Tensor:
    """uint8 (H, W, 3) staged sample (gray stored in channel 0) -> normalised float32 (target_channels, H, W)"""
    image = img_u8_hwc[..., :src_channels].permute(2, 0, 1).to(torch.float32).div(255)     # transforms.ToTensor
    if target_channels == 1:
        if image.shape[0] == 3:
            image = (0.299 * image[0] + 0.587 * image[1] + 0.114 * image[2]).unsqueeze(0)
    elif target_channels == 3:
        if image.shape[0] == 1:
            image = image.repeat(3, 1, 1)
    return (image - 0.5) / 0.5                                                                # Normalize(0.5, 0.5)


def collate(images: Sequence[torch.Tensor]) -> torch.Tensor:
    widest = max(i.shape[0] for i in images)
    return torch.stack([torch.cat([i, torch.zeros(widest - i.shape[0], *i.shape[1:])]) if i.shape[0] < widest else i
                        for i in images])


def feed_batch(staged_u8, src_c, modality_idx, sel, n_modalities: int = 12, gray=GRAY_DATASETS_12):
    """-> (x, onehot, idx) for the samples `sel` of a staged store"""
    imgs = [getitem_image(staged_u8[i], int(src_c[i]), 1 if int(modality_idx[i]) in gray else 3) for i in sel]
    idx = torch.as_tensor([int(modality_idx[i]) for i in sel], dtype=torch.int64)
    return collate(imgs), F.one_hot(idx, n_modalities).float(), idx


def film(features, cond, sd):
    scale = F.linear(cond, sd["scale_transform.weight"], sd["scale_transform.bias"]).unsqueeze(-1).unsqueeze(-1)
    shift = F.linear(cond, sd["shift_transform.weight"], sd["shift_transform.bias"]).unsqueeze(-1).unsqueeze(-1)
    return features * scale + shift
