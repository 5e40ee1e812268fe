"""The no-allocation entry point ``mvae_attn_fused_bwd_static`` (include/medvae_b200.h; no counterpart in the reference):
the fused attention backward with its row-dot scratch in a file-scope static
``__device__`` array instead of a caller-provided buffer.  Called through the C ABI (ctypes) and compared with the
existing entry point ``mvae_attn_fused_bwd`` on two calls in a row that share the array.

This file is named to be collected LAST (pytest -x must not lose other tests to a failure here) and
passes both alone in a fresh process and after the rest of the suite."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
MAX_ROWS = 1048576      # MVAE_ATTN_STATIC_MAX_ROWS


def _mods():
    from medvae_disentangled_multimodal_b200 import _lib as L
    return L


def _case(L, B, T, c, seed):
    """-> (dqkv by the existing entry point, dqkv by the static-scratch one), everything through the C ABI"""
    lib = L.lib()
    g = torch.Generator().manual_seed(seed)
    qkv = (torch.randn(B, T, 3 * c, generator=g) * 0.7).to(DEV, torch.bfloat16)
    dout = torch.randn(B, T, c, generator=g).to(DEV, torch.bfloat16)
    out = torch.empty(B, T, c, dtype=torch.bfloat16, device=DEV)
    lse2 = torch.empty(B, T, dtype=torch.float32, device=DEV)
    st = L.stream()
    assert lib.mvae_attn_fused_ok(T, c, L.BF16) == 1
    L.check(lib.mvae_attn_fused_fwd(L.ptr(qkv), L.ptr(out), L.ptr(lse2), B, T, c, L.BF16, st), "attn_fused_fwd")
    rowdot = torch.empty(B * T, dtype=torch.float32, device=DEV)
    ref = torch.full_like(qkv, float("nan"))
    L.check(lib.mvae_attn_fused_bwd(L.ptr(qkv), L.ptr(out), L.ptr(lse2), L.ptr(dout), L.ptr(ref), L.ptr(rowdot), B, T, c,
                                    L.BF16, st), "attn_fused_bwd")
    got = torch.full_like(qkv, float("nan"))
    L.check(lib.mvae_attn_fused_bwd_static(L.ptr(qkv), L.ptr(out), L.ptr(lse2), L.ptr(dout), L.ptr(got), B, T, c, L.BF16,
                                           st), "attn_fused_bwd_static")
    torch.cuda.synchronize()
    return ref, got


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def test_static_scratch_backward_matches_the_existing_entry_point_on_two_calls_in_a_row():
    L = _mods()
    # the second call is smaller than the first and has other data: it runs over the first call's leftovers in the array
    for B, T, c, seed in [(24, 256, 128, 11), (7, 49, 64, 12)]:
        ref, got = _case(L, B, T, c, seed)
        assert torch.isfinite(ref.float()).all() and torch.isfinite(got.float()).all()
        # the same kernels on the same inputs; 2e-2 is the bf16 budget of tests/test_kernels_gpu.py (TOL), bit equality
        # is what is expected
        assert _rel(got, ref) < 2e-2
        assert torch.equal(got, ref)


def test_static_scratch_backward_rejects_sizes_beyond_the_maximum():
    L = _mods()
    lib = L.lib()
    B, T, c = MAX_ROWS // 256 + 1, 256, 64      # batch * tokens = MAX_ROWS + 256
    one = torch.zeros(16, dtype=torch.float32, device=DEV)      # never touched: the call must fail before any launch
    rc = lib.mvae_attn_fused_bwd_static(L.ptr(one), L.ptr(one), L.ptr(one), L.ptr(one), L.ptr(one), B, T, c, L.BF16,
                                        L.stream())
    assert rc != 0
    msg = L.last_error()
    assert "MVAE_ATTN_STATIC_MAX_ROWS" in msg and str(MAX_ROWS) in msg
    torch.cuda.synchronize()
