"""Synthetic MedMNIST-shaped batches in the reference's batch-tuple layout
``(x, labels, modality_onehot, modality_idx)`` (src/data/medmnist_data.py:251): x float32 NCHW
in [-1, 1] (``Normalize(0.5, 0.5)``, :372-380), one-hot float32 (B, n_modalities), and the
``mixed_modality_collate_fn`` convention that 1-channel modalities are zero-padded to 3 channels
(:16-72).  Gray modalities are indices {0, 3, 9, 10, 11} of the 12-modality list (:138-163).
"""
from __future__ import annotations

from typing import Sequence

import torch

GRAY_12 = (0, 3, 9, 10, 11)
GRAY_5 = (0, 3)


def synthetic_batch(batch: int, channels: int, res: int, n_modalities: int = 12, seed: int = 42,
                    gray: Sequence[int] = GRAY_12, first_index: int = 0, pin: bool = False):
    """-> (x, labels, onehot, idx).  ``first_index`` is the global index of sample 0 so that rank
    shards of one global batch get the modality pattern ``i mod n_modalities`` of the whole batch."""
    g = torch.Generator().manual_seed(seed + first_index)
    x = torch.rand(batch, channels, res, res, generator=g) * 2 - 1
    idx = (torch.arange(batch) + first_index) % n_modalities
    onehot = torch.nn.functional.one_hot(idx, n_modalities).float()
    if channels == 3:
        for m in gray:
            x[idx == m, 1:] = 0.0
    labels = torch.zeros(batch, 1, dtype=torch.int64)
    if pin and torch.cuda.is_available():
        x, onehot = x.pin_memory(), onehot.pin_memory()
    return x, labels, onehot, idx
