"""Host logic of the conv path, checked without a GPU: the tap programs of csrc/conv_lower.h (what both
the tcgen05 and the SIMT launchers execute) are interpreted on the CPU by tests/lowering_emu.cu and
compared with the reference's own ops and their autograd — nn.Conv2d s1 (encoder_decoder.py:128-146),
Downsample = F.pad(0,1,0,1) + conv3x3 s2 (:184-188), Upsample = nearest-2x + conv3x3 (:205-209).
The packed-weight layouts ([slab][N][K], 16 parity slabs for Upsample) are restated here in numpy from
the description in include/medvae_b200.h / csrc/pack.cu."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from medvae_disentangled_multimodal_b200 import _lib as L
from medvae_disentangled_multimodal_b200 import ops

HERE = os.path.dirname(os.path.abspath(__file__))
S1, DOWN, UP = 0, 1, 2


@pytest.fixture(scope="module")
def emu():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    out_dir = os.path.join(HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "liblowering_emu.so")
    src = os.path.join(HERE, "lowering_emu.cu")
    deps = [src] + [os.path.join(HERE, "..", "medvae_disentangled_multimodal_b200", "csrc", f)
                    for f in ("conv_lower.h", "common.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.run([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-shared",
                        "-Xcompiler", "-fPIC", "-o", so, src], check=True, capture_output=True)
    lib = C.CDLL(so)
    for f in (lib.emu_fprop, lib.emu_dgrad, lib.emu_wgrad, lib.emu_program_shape):
        f.restype = C.c_int
    return lib


def up_rset(a, ui):
    """3x3 taps that land on source offset ui of output parity a (csrc/pack.cu:up_rset)."""
    return {(0, 0): [0], (0, 1): [1, 2], (1, 0): [0, 1], (1, 1): [2]}[(a, ui)]


def slab_taps(mode, k):
    """-> per slab the list of (r, s) filter taps it sums."""
    if mode != UP:
        return [[(r, s)] for r in range(k) for s in range(k)]
    out = []
    for a in range(2):
        for b in range(2):
            for ui in range(2):
                for vi in range(2):
                    out.append([(r, s) for r in up_rset(a, ui) for s in up_rset(b, vi)])
    return out


def pack(w, mode):
    """OIHW -> fprop operand [slab][Cout][Cin] and dgrad operand [slab][Cin][Cout]."""
    k = w.shape[-1]
    wf = np.stack([sum(w[:, :, r, s] for r, s in taps) for taps in slab_taps(mode, k)])
    return np.ascontiguousarray(wf, dtype=np.float32), np.ascontiguousarray(wf.transpose(0, 2, 1), dtype=np.float32)


def unpack_dw(dw_slabs, mode, k):
    """slab-major wgrad partials -> OIHW (what wgrad_finalize_kernel does): every filter tap collects the
    slabs it contributed to."""
    cout, cin = dw_slabs.shape[1:]
    dw = np.zeros((cout, cin, k, k), dtype=np.float64)
    for slab, taps in enumerate(slab_taps(mode, k)):
        for r, s in taps:
            dw[:, :, r, s] += dw_slabs[slab]
    return dw


def reference(x, w, mode):
    if mode == DOWN:
        return F.conv2d(F.pad(x, (0, 1, 0, 1), mode="constant", value=0), w, stride=2, padding=0)
    if mode == UP:
        x = F.interpolate(x, scale_factor=2.0, mode="nearest")
    return F.conv2d(x, w, stride=1, padding=(w.shape[-1] - 1) // 2)


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


CASES = [  # (B, H, W, Cin, Cout, k, mode)
    (2, 7, 7, 3, 4, 3, S1), (1, 5, 9, 4, 2, 3, S1), (2, 6, 6, 5, 3, 1, S1),
    (2, 8, 8, 3, 4, 3, DOWN), (1, 7, 7, 2, 3, 3, DOWN), (2, 14, 14, 2, 2, 3, DOWN), (1, 5, 8, 3, 2, 3, DOWN),
    (2, 4, 4, 3, 4, 3, UP), (1, 7, 7, 2, 3, 3, UP), (1, 3, 5, 4, 2, 3, UP),
]


@pytest.mark.parametrize("B,H,W,Cin,Cout,k,mode", CASES)
def test_tap_programs_match_torch_conv_and_autograd(emu, B, H, W, Cin, Cout, k, mode):
    g = torch.Generator().manual_seed(1000 * mode + 10 * H + k)
    x = torch.randn(B, Cin, H, W, generator=g, dtype=torch.float64, requires_grad=True)
    w = torch.randn(Cout, Cin, k, k, generator=g, dtype=torch.float64, requires_grad=True)
    y = reference(x, w, mode)
    dy = torch.randn(y.shape, generator=g, dtype=torch.float64)
    y.backward(dy)

    d = ops.conv_desc(B, H, W, Cin, Cout, k, mode, L.F32)
    P, Q = ops.conv_out_hw(d)
    assert (P, Q) == tuple(y.shape[-2:])
    wf, wd = pack(w.detach().numpy(), mode)
    assert wf.shape[0] == ops.conv_num_slabs(d)
    x_nhwc = np.ascontiguousarray(x.detach().permute(0, 2, 3, 1).numpy(), dtype=np.float32)
    dy_nhwc = np.ascontiguousarray(dy.permute(0, 2, 3, 1).numpy(), dtype=np.float32)

    y_emu = np.full((B, P, Q, Cout), np.nan, dtype=np.float32)          # every output written exactly once
    assert emu.emu_fprop(C.byref(d), fptr(x_nhwc), fptr(wf), fptr(y_emu)) == 0
    np.testing.assert_allclose(y_emu.transpose(0, 3, 1, 2), y.detach().numpy(), rtol=2e-5, atol=2e-5)

    dx_emu = np.full((B, H, W, Cin), np.nan, dtype=np.float32)
    assert emu.emu_dgrad(C.byref(d), fptr(dy_nhwc), fptr(wd), fptr(dx_emu)) == 0
    np.testing.assert_allclose(dx_emu.transpose(0, 3, 1, 2), x.grad.numpy(), rtol=2e-5, atol=2e-5)

    dw_slabs = np.full((wf.shape[0], Cout, Cin), np.nan, dtype=np.float32)
    assert emu.emu_wgrad(C.byref(d), fptr(x_nhwc), fptr(dy_nhwc), fptr(dw_slabs)) == 0
    np.testing.assert_allclose(unpack_dw(dw_slabs.astype(np.float64), mode, k), w.grad.numpy(), rtol=5e-5, atol=5e-5)


@pytest.mark.parametrize("mode,k,expect", [
    (S1, 3, ((1, 9), (1, 9), 9)), (S1, 1, ((1, 1), (1, 1), 1)),
    (DOWN, 3, ((1, 9), (4, 9), 9)),      # dgrad: 4 parity problems of 4 + 2 + 2 + 1 taps
    (UP, 3, ((4, 16), (1, 16), 16)),     # fprop: 4 parity problems of 2x2 taps instead of 9 taps on the 2x grid
])
def test_program_shapes(emu, mode, k, expect):
    d = ops.conv_desc(2, 8, 8, 4, 4, k, mode, L.F32)
    npr, nt = C.c_int(), C.c_int()
    for which, want in zip((0, 1), expect[:2]):
        assert emu.emu_program_shape(C.byref(d), which, C.byref(npr), C.byref(nt)) == 0
        assert (npr.value, nt.value) == want
    assert emu.emu_program_shape(C.byref(d), 2, C.byref(npr), C.byref(nt)) == 0
    assert nt.value == expect[2]


def test_upsample_fold_saves_flops():
    """the four 2x2-tap parity problems run 16 taps per 4 output pixels instead of 36: 2.25x fewer MMAs
    than the reference's conv on the materialised 2x tensor."""
    taps = slab_taps(UP, 3)
    assert len(taps) == 16 and sum(len(t) for t in taps) == 36
    assert all(sorted(sum((taps[p * 4 + i] for i in range(4)), [])) == [(r, s) for r in range(3) for s in range(3)]
               for p in range(4))
