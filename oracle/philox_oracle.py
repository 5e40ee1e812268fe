"""TEST INFRASTRUCTURE — numpy restatement of the counter-based RNG of the B200 path (not part of the
reference: the reference draws torch.randn_like at src/models/base_vae.py:85 and nn.Dropout masks at
src/models/encoder_decoder.py:164 from torch's global generator; SURVEY.md 8a row a12 asks for a
Philox4x32-10 counter RNG instead).  The algorithm is the published one (Salmon, Moraes, Dror, Shaw:
"Parallel random numbers: as easy as 1, 2, 3", SC'11; Random123 ``philox4x32-10``); it is pinned by
that library's known-answer vectors (KAT) in tests/test_philox_cpu.py.  Only tests/ may import this."""
from __future__ import annotations

import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)

# Random123 kat_vectors, generator "philox4x32 10": (counter[4], key[2]) -> output[4]
KAT = [
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000),
     (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
    ((0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF, 0xFFFFFFFF),
     (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
    ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
     (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
]


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """vectorised over numpy arrays of uint32 counters; scalar keys.  -> 4 uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & MASK for c in (c0, c1, c2, c3))
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0          # 32 x 32 -> 64 bit products (no overflow in uint64)
        p1 = M1 * c2
        n0 = (p1 >> np.uint64(32)) ^ c1 ^ np.uint64(k0)
        n2 = (p0 >> np.uint64(32)) ^ c3 ^ np.uint64(k1)
        c0, c1, c2, c3 = n0, p1 & MASK, n2, p0 & MASK
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def block(seed: int, ctr_lo, ctr_hi):
    """key = 64-bit seed, counter = {ctr_lo, ctr_hi} as 64-bit halves (csrc/philox.cuh:Philox)."""
    lo = np.asarray(ctr_lo, dtype=np.uint64)
    hi = np.asarray(ctr_hi, dtype=np.uint64)
    return philox4x32_10(lo & MASK, lo >> np.uint64(32), hi & MASK, hi >> np.uint64(32),
                         seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)


def dropout_keep_mask(n: int, p: float, seed: int, offset: int, elem_base: int = 0) -> np.ndarray:
    """keep[e] = u >= p with u = (word >> 8) * 2^-24 (exact in fp32), word (g & 3) of block g // 4 of stream
    (seed, offset), g = elem_base + e the element's index in the global batch: csrc/feed.cu:dropout_kernel.
    Integer work + one exact float compare: bit-exact."""
    g = np.arange(n, dtype=np.uint64) + np.uint64(elem_base)
    q0 = int(elem_base) // 4
    q = np.arange(q0, (int(elem_base) + n + 3) // 4 + 1, dtype=np.uint64)
    words = np.stack(block(seed, q, np.full_like(q, offset)), axis=1).reshape(-1)
    words = words[(g - np.uint64(4 * q0)).astype(np.int64)]
    u = (words >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    return u >= np.float32(p)


def normal_words(n: int, seed: int, offset: int, elem_base: int = 0):
    """the two words element e draws: (x, y) of block e >> 1 for even e, (z, w) for odd e.  elem_base = the first
    element's index in the global batch (first sample x latent elements per sample)."""
    e = np.arange(n, dtype=np.uint64) + np.uint64(elem_base)
    x, y, z, w = block(seed, e >> np.uint64(1), np.full_like(e, offset))
    odd = (e & np.uint64(1)).astype(bool)
    return np.where(odd, z, x), np.where(odd, w, y)


def uniform_open(words) -> np.ndarray:
    """((word >> 8) + 0.5) * 2^-24 evaluated in fp32 as the kernel does (the sum rounds to even above 2^23,
    so the largest value is exactly 1.0 and the smallest 2^-25)."""
    return ((words >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 16777216.0)


def normal_noise(n: int, seed: int, offset: int, elem_base: int = 0) -> np.ndarray:
    """Box-Muller on uniforms ((word >> 8) + 0.5) * 2^-24 (csrc/loss.cu:philox_normal).  The device
    evaluates log / cos with fast intrinsics in fp32, so this agrees to ~1e-4 absolute, not bit for bit;
    the words are exact."""
    a, b = normal_words(n, seed, offset, elem_base)
    u1, u2 = (uniform_open(w).astype(np.float64) for w in (a, b))
    return (np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)).astype(np.float32)
