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


# --------------------------------------------------------------------------------------------------
# on-device batch feed
# --------------------------------------------------------------------------------------------------
# dataset order of MedMNISTDataset._create_modality_map (src/data/medmnist_data.py:138-153) and the channel
# count each modality is converted to (_get_modality_channels, :155-184)
DATASETS_12 = ("chestmnist", "pathmnist", "octmnist", "pneumoniamnist", "dermamnist", "bloodmnist", "tissuemnist",
               "retinamnist", "breastmnist", "organamnist", "organcmnist", "organsmnist")
TARGET_CHANNELS_12 = tuple(1 if m in GRAY_12 else 3 for m in range(12))


class DeviceBatchFeed:
    """Staged uint8 samples -> training batches, on the GPU.

    What the reference does per sample in DataLoader worker processes — ``ToTensor``, RGB<->gray
    conversion to the modality's channel count, ``Normalize(0.5, 0.5)`` (src/data/medmnist_data.py:186-251,
    372-380) — and per batch in ``mixed_modality_collate_fn`` (zero-pad 1-channel samples to the widest
    sample of the batch, :16-72) is one kernel over a uint8 batch that was gathered on the device; the
    result is bit-identical fp32.  Returns the reference's batch tuple
    ``(x, labels, modality_onehot, modality_idx)``.

    images: uint8 (N, H, W, SC), SC in {1, 3}, host or device; src_channels: (N,) valid channels per
    sample (1: grayscale stored in channel 0); modality_idx: (N,) int64; labels: (N, 1) int64 or None.
    """

    def __init__(self, images: torch.Tensor, src_channels, modality_idx, labels=None, n_modalities: int = 12,
                 target_channels: Sequence[int] = TARGET_CHANNELS_12, device="cuda"):
        if images.dtype != torch.uint8 or images.dim() != 4 or images.shape[-1] not in (1, 3):
            raise ValueError("images must be uint8 (N, H, W, 1|3)")
        self.device = torch.device(device)
        self.n_modalities = n_modalities
        self._tgt_host = [int(target_channels[int(m)]) for m in modality_idx]
        self.images = images.to(self.device).contiguous()
        self.src_c = torch.as_tensor(src_channels, dtype=torch.int32).to(self.device)
        self.tgt_c = torch.tensor(self._tgt_host, dtype=torch.int32).to(self.device)
        self.idx = torch.as_tensor(modality_idx, dtype=torch.int64).to(self.device)
        n = images.shape[0]
        self.labels = (torch.zeros(n, 1, dtype=torch.int64) if labels is None else torch.as_tensor(labels)).to(self.device)

    def __len__(self):
        return self.images.shape[0]

    def batch(self, indices):
        """indices: host sequence / tensor of sample numbers -> (x, labels, onehot, idx) on the device.
        The batch's channel count is that of its widest sample (decided on the host, no device read)."""
        from . import _lib as L
        from . import ops
        L.require_device()
        host = indices.tolist() if torch.is_tensor(indices) else list(indices)
        out_c = max(self._tgt_host[i] for i in host)
        sel = torch.as_tensor(host, dtype=torch.int64).to(self.device, non_blocking=True)
        src = self.images.index_select(0, sel)
        idx = self.idx.index_select(0, sel)
        x, onehot = ops.batch_feed_u8(src, self.src_c.index_select(0, sel), self.tgt_c.index_select(0, sel), idx, out_c,
                                      self.n_modalities)
        return x, self.labels.index_select(0, sel), onehot, idx
