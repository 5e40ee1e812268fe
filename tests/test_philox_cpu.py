"""Philox4x32-10 (counter-based RNG for the reparameterisation noise and the dropout
masks), checked without a GPU: the shipped header csrc/philox.cuh — compiled for the host by
tests/philox_emu.cu — and the numpy restatement oracle/philox_oracle.py both reproduce the Random123
known-answer vectors, and agree with each other on the stream layouts the kernels use.  Integer work:
bit-exact."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import philox_oracle as P

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def emu():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    out_dir = os.path.join(HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libphilox_emu.so")
    src = os.path.join(HERE, "philox_emu.cu")
    deps = [src, os.path.join(HERE, "..", "medvae_disentangled_multimodal_b200", "csrc", "philox.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.run([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-shared",
                        "-Xcompiler", "-fPIC", "-o", so, src], check=True, capture_output=True)
    lib = C.CDLL(so)
    lib.emu_philox_block.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint32)]
    lib.emu_keep_mask.argtypes = [C.c_longlong, C.c_float, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint8)]
    lib.emu_normal_uniforms.argtypes = [C.c_longlong, C.c_uint64, C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    return lib


@pytest.mark.parametrize("ctr,key,want", P.KAT)
def test_known_answer_vectors(emu, ctr, key, want):
    out = (C.c_uint32 * 4)()
    emu.emu_philox_raw((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
    assert tuple(out) == want                                               # the shipped header
    got = P.philox4x32_10(*[np.array([c]) for c in ctr], *key)
    assert tuple(int(g[0]) for g in got) == want                            # the numpy restatement


def test_seed_and_counter_packing(emu):
    """key = {seed lo, seed hi}; counter = {ctr_lo lo, ctr_lo hi, ctr_hi lo, ctr_hi hi}"""
    rng = np.random.default_rng(0)
    for _ in range(64):
        seed, lo, hi = (int(v) for v in rng.integers(0, 2 ** 63, size=3, dtype=np.uint64))
        out = (C.c_uint32 * 4)()
        emu.emu_philox_block(seed, lo, hi, out)
        want = P.block(seed, np.array([lo], dtype=np.uint64), np.array([hi], dtype=np.uint64))
        assert tuple(out) == tuple(int(w[0]) for w in want)
    # the third KAT written as (seed, ctr_lo, ctr_hi)
    ctr, key, want = P.KAT[2]
    out = (C.c_uint32 * 4)()
    emu.emu_philox_block(key[0] | (key[1] << 32), ctr[0] | (ctr[1] << 32), ctr[2] | (ctr[3] << 32), out)
    assert tuple(out) == want


@pytest.mark.parametrize("n,p,seed,offset", [(1, 0.5, 1, 0), (7, 0.1, 1234, 5), (4096, 0.1, 0xDEADBEEFCAFE, 3 << 20),
                                             (1001, 0.9, 2 ** 63 + 11, 2 ** 40), (64, 0.0, 9, 9)])
def test_dropout_keep_mask_layout(emu, n, p, seed, offset):
    mask = np.zeros(n, dtype=np.uint8)
    emu.emu_keep_mask(n, p, seed, offset, mask.ctypes.data_as(C.POINTER(C.c_uint8)))
    want = P.dropout_keep_mask(n, p, seed, offset)
    assert np.array_equal(mask.astype(bool), want)
    if p == 0.0:
        assert want.all()
    elif n >= 1000:
        assert abs(want.mean() - (1.0 - p)) < 0.05


def test_normal_stream_layout_and_moments(emu):
    n, seed, offset = 1 << 16, 1234, 1
    u1 = np.zeros(n, dtype=np.float32)
    u2 = np.zeros(n, dtype=np.float32)
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    emu.emu_normal_uniforms(n, seed, offset, fp(u1), fp(u2))
    a, b = P.normal_words(n, seed, offset)
    assert np.array_equal(u1, P.uniform_open(a)) and np.array_equal(u2, P.uniform_open(b))
    assert u1.min() > 0.0 and u1.max() <= 1.0                               # log(u1) finite, never positive
    z = P.normal_noise(n, seed, offset).astype(np.float64)
    assert abs(z.mean()) < 0.02 and abs(z.std() - 1.0) < 0.02 and abs((z ** 4).mean() - 3.0) < 0.15
    # element e and e ^ 1 share a block, different blocks are unrelated
    assert abs(np.corrcoef(z[0::2], z[1::2])[0, 1]) < 0.02
    assert not np.array_equal(P.normal_noise(64, seed, offset + 1), z[:64].astype(np.float32))
