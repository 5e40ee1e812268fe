"""Tensor-level wrappers over the C ABI (one Python function per entry point family).

All activations are NHWC torch tensors ([B, H, W, C], bf16 or fp32) living on the current CUDA
device; work is enqueued on torch's current stream.  Nothing here computes with torch ops.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L
from . import profile as P
from ._lib import ConvDesc, check, lib, ptr, stream

_scratch = {}


def scratch(name: str, nbytes: int, device) -> torch.Tensor:
    """grow-only byte scratch buffers, one per (name, device, stream-agnostic): kernels that use
    them are ordered on the current stream."""
    key = (name, str(device))
    t = _scratch.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.empty(max(nbytes, 1024), dtype=torch.uint8, device=device)
        _scratch[key] = t
    return t


def conv_desc(batch, h, w, cin, cout, ksize, mode, dtype_code, engine=L.ENGINE_AUTO) -> ConvDesc:
    return ConvDesc(batch, h, w, cin, cout, ksize, mode, dtype_code, engine)


def conv_out_hw(d: ConvDesc):
    oh, ow = C.c_int32(), C.c_int32()
    check(lib().mvae_conv2d_out_hw(C.byref(d), C.byref(oh), C.byref(ow)), "conv2d_out_hw")
    return oh.value, ow.value


def conv_num_slabs(d: ConvDesc) -> int:
    return lib().mvae_conv2d_num_slabs(C.byref(d))


def conv_pack_weights(d: ConvDesc, w_oihw, wf, wd, ctot, coff):
    check(lib().mvae_conv2d_pack_weights(C.byref(d), ptr(w_oihw), ptr(wf), ptr(wd), ctot, coff, stream()),
          "conv2d_pack_weights")


def conv_pack_weights_batched(table, n_jobs: int, total_tiles: int, dtype_code: int):
    """table: device uint8 tensor holding n_jobs mvae_pack_job structs (blocks.pack_entries)"""
    with P.region("pack_weights", 0.0, 0.0):
        check(lib().mvae_conv2d_pack_weights_batched(ptr(table), int(n_jobs), int(total_tiles), int(dtype_code), stream()),
              "conv2d_pack_weights_batched")


def conv_flops(d: ConvDesc) -> float:
    """algorithmic FLOPs (2*MAC) of one direction of this conv as the reference executes it
    (Upsample = 9 taps on the 2x grid, although the kernels run 4 taps per output parity)."""
    oh, ow = conv_out_hw(d)
    # a skinny conv run with zero-padded channels (blocks.ConvOp.pad_in / pad_out) counts its REAL channels:
    # ConvOp.desc() leaves them on the descriptor object
    cin, cout = getattr(d, "real_channels", (d.in_c, d.out_c))
    return 2.0 * d.batch * oh * ow * cout * cin * d.ksize * d.ksize


def conv_fprop(d: ConvDesc, x, wf, bias, residual, out=None):
    oh, ow = conv_out_hw(d)
    y = out if out is not None else torch.empty((d.batch, oh, ow, d.out_c), dtype=x.dtype, device=x.device)
    if P.active() is None:
        check(lib().mvae_conv2d_fprop(C.byref(d), ptr(x), ptr(wf), ptr(bias), ptr(residual), ptr(y), stream()),
              "conv2d_fprop")
        return y
    kind = "conv_fprop_umma" if lib().mvae_conv2d_uses_umma(C.byref(d)) else "conv_fprop_simt"
    with P.region(kind, conv_flops(d), (x.numel() + y.numel()) * x.element_size()):
        check(lib().mvae_conv2d_fprop(C.byref(d), ptr(x), ptr(wf), ptr(bias), ptr(residual), ptr(y), stream()),
              "conv2d_fprop")
    return y


def conv_fuses_gn(d: ConvDesc, groups: int, direction: int) -> bool:
    """can this conv carry the GroupNorm hook in its epilogue (direction 0: fprop / output channels,
    1: dgrad / input channels)?"""
    return bool(lib().mvae_conv2d_fuses_gn(C.byref(d), int(groups), int(direction)))


def conv_fprop_gn(d: ConvDesc, x, wf, bias, residual, groups: int):
    """fprop that also returns the per-(sample, group) (sum, sum of squares) of its output."""
    oh, ow = conv_out_hw(d)
    y = torch.empty((d.batch, oh, ow, d.out_c), dtype=x.dtype, device=x.device)
    sums = torch.empty((d.batch, groups, 2), dtype=torch.float32, device=x.device)
    with P.region("conv_fprop_umma", conv_flops(d) if P.active() else 0.0, (x.numel() + y.numel()) * x.element_size()):
        check(lib().mvae_conv2d_fprop_gn(C.byref(d), ptr(x), ptr(wf), ptr(bias), ptr(residual), ptr(y), ptr(sums),
                                         int(groups), stream()), "conv2d_fprop_gn")
    return y, sums


def conv_dgrad_gn(d: ConvDesc, dy, wd, gn_x, stats, gamma, beta, groups: int, silu: bool):
    """dgrad whose epilogue applies the SiLU / affine part of the GroupNorm backward:
    -> (du, s12) with s12[n][g] = (sum du*gamma, sum du*gamma*xhat)."""
    du = torch.empty((d.batch, d.in_h, d.in_w, d.in_c), dtype=dy.dtype, device=dy.device)
    s12 = torch.empty((d.batch, groups, 2), dtype=torch.float32, device=dy.device)
    with P.region("conv_dgrad_umma", conv_flops(d) if P.active() else 0.0,
                  (dy.numel() + 2 * du.numel()) * dy.element_size()):
        check(lib().mvae_conv2d_dgrad_gn(C.byref(d), ptr(dy), ptr(wd), ptr(du), ptr(gn_x), ptr(stats), ptr(gamma),
                                         ptr(beta), int(groups), int(silu), ptr(s12), stream()), "conv2d_dgrad_gn")
    return du, s12


def gn_apply_fwd(x, gamma, beta, sums, groups, eps, silu):
    """one-pass GroupNorm(+SiLU) from precomputed sums -> (y, stats)"""
    b, h, w, c = x.shape
    y = torch.empty_like(x)
    stats = torch.empty((b, groups, 2), dtype=torch.float32, device=x.device)
    with P.region("groupnorm_fwd", 0.0, 2 * x.numel() * x.element_size()):
        check(lib().mvae_groupnorm_apply_fwd(ptr(x), ptr(gamma), ptr(beta), ptr(y), ptr(sums), ptr(stats), b, h * w, c,
                                             groups, eps, int(silu), L.code_of(x.dtype), stream()),
              "groupnorm_apply_fwd")
    return y, stats


def gn_apply_bwd(du, x, gamma, stats, s12, dx_add, groups):
    """one-pass GroupNorm backward from du and the group sums -> (dx, dgamma, dbeta, colsum(dx))"""
    b, h, w, c = x.shape
    dx = torch.empty_like(x)
    out3 = torch.empty((3, c), dtype=torch.float32, device=x.device)
    sc = scratch("gn1", 4 * lib().mvae_groupnorm_apply_bwd_scratch_floats(b, h * w, c), x.device)
    with P.region("groupnorm_bwd", 0.0, (3 + (dx_add is not None)) * x.numel() * x.element_size()):
        check(lib().mvae_groupnorm_apply_bwd(ptr(du), ptr(x), ptr(gamma), ptr(stats), ptr(s12), ptr(dx_add), ptr(dx),
                                             ptr(out3), ptr(sc), b, h * w, c, groups, L.code_of(x.dtype), stream()),
              "groupnorm_apply_bwd")
    return dx, out3[0], out3[1], out3[2]


def conv_dgrad(d: ConvDesc, dy, wd):
    dx = torch.empty((d.batch, d.in_h, d.in_w, d.in_c), dtype=dy.dtype, device=dy.device)
    kind = "conv_dgrad"
    if P.active() is not None:
        kind = "conv_dgrad_umma" if lib().mvae_conv2d_uses_umma(C.byref(d)) else "conv_dgrad_simt"
    with P.region(kind, conv_flops(d) if P.active() else 0.0, (dy.numel() + dx.numel()) * dy.element_size()):
        check(lib().mvae_conv2d_dgrad(C.byref(d), ptr(dy), ptr(wd), ptr(dx), stream()), "conv2d_dgrad")
    return dx


def conv_wgrad(d: ConvDesc, x, dy, want_bias=True):
    dw = torch.empty((d.out_c, d.in_c, d.ksize, d.ksize), dtype=torch.float32, device=x.device)
    db = torch.empty((d.out_c,), dtype=torch.float32, device=x.device) if want_bias else None
    nbytes = lib().mvae_conv2d_wgrad_workspace_bytes(C.byref(d))
    ws = scratch("wgrad", nbytes, x.device)
    kind = "conv_wgrad"
    if P.active() is not None:
        kind = "conv_wgrad_umma" if lib().mvae_conv2d_uses_umma(C.byref(d)) else "conv_wgrad_simt"
    with P.region(kind, conv_flops(d) if P.active() else 0.0, (x.numel() + dy.numel()) * x.element_size()):
        check(lib().mvae_conv2d_wgrad(C.byref(d), ptr(x), ptr(dy), ptr(dw), ptr(db), ptr(ws), ws.numel(), stream()),
              "conv2d_wgrad")
    return dw, db


def conv_wgrad_acc(d: ConvDesc, x, dy, dw_view, db_view, beta=1.0):
    """wgrad accumulated in place: dw_view = beta*dw_view + dW (contiguous OIHW fp32 view, e.g. a
    slice of the training step's flat gradient buffer); db_view likewise, or None to skip the bias."""
    nbytes = lib().mvae_conv2d_wgrad_workspace_bytes(C.byref(d))
    ws = scratch("wgrad", nbytes, x.device)
    kind = "conv_wgrad"
    if P.active() is not None:
        kind = "conv_wgrad_umma" if lib().mvae_conv2d_uses_umma(C.byref(d)) else "conv_wgrad_simt"
    with P.region(kind, conv_flops(d) if P.active() else 0.0, (x.numel() + dy.numel()) * x.element_size()):
        check(lib().mvae_conv2d_wgrad_acc(C.byref(d), ptr(x), ptr(dy), ptr(dw_view), ptr(db_view), ptr(ws), ws.numel(),
                                          float(beta), stream()), "conv2d_wgrad_acc")


def gn_fwd(x, gamma, beta, groups, eps, silu):
    b, h, w, c = x.shape
    y = torch.empty_like(x)
    stats = torch.empty((b, groups, 2), dtype=torch.float32, device=x.device)
    sc = scratch("gn", 4 * lib().mvae_groupnorm_scratch_floats(b, h * w, c, groups, 0), x.device)
    with P.region("groupnorm_fwd", 0.0, 2 * x.numel() * x.element_size()):
        check(lib().mvae_groupnorm_silu_fwd(ptr(x), ptr(gamma), ptr(beta), ptr(y), ptr(stats), ptr(sc), b, h * w, c,
                                            groups, eps, int(silu), L.code_of(x.dtype), stream()),
              "groupnorm_silu_fwd")
    return y, stats


def gn_bwd(dy, x, gamma, beta, stats, dx_add, groups, silu):
    b, h, w, c = x.shape
    dx = torch.empty_like(x)
    dgb = torch.empty((2, c), dtype=torch.float32, device=x.device)
    sc = scratch("gn", 4 * lib().mvae_groupnorm_scratch_floats(b, h * w, c, groups, 1), x.device)
    with P.region("groupnorm_bwd", 0.0, (3 + (dx_add is not None)) * x.numel() * x.element_size()):
        check(lib().mvae_groupnorm_silu_bwd(ptr(dy), ptr(x), ptr(gamma), ptr(beta), ptr(stats), ptr(dx_add), ptr(dx),
                                            ptr(dgb[0]), ptr(dgb[1]), ptr(sc), b, h * w, c, groups, int(silu),
                                            L.code_of(x.dtype), stream()), "groupnorm_silu_bwd")
    return dx, dgb[0], dgb[1]


def attn_fwd(qkv):
    """-> (out, saved): saved is what attn_bwd needs besides qkv — ("fused", out, lse2 [B, T] fp32) on the fused tcgen05
    kernel (bf16, C % 64 == 0, T <= 256: the probabilities are recomputed in the backward), else ("probs", [B, T, T])"""
    b, h, w, c3 = qkv.shape
    c, t = c3 // 3, h * w
    out = torch.empty((b, h, w, c), dtype=qkv.dtype, device=qkv.device)
    code = L.code_of(qkv.dtype)
    if lib().mvae_attn_fused_ok(t, c, code):
        lse2 = torch.empty((b, t), dtype=torch.float32, device=qkv.device)
        with P.region("attn_fwd", 4.0 * b * t * t * c, (qkv.numel() + out.numel()) * 2):
            check(lib().mvae_attn_fused_fwd(ptr(qkv), ptr(out), ptr(lse2), b, t, c, code, stream()), "attn_fused_fwd")
        return out, ("fused", out, lse2)
    # the tcgen05 batched-GEMM path keeps P in bf16 (pad8(T) columns), the SIMT / fp32 path in fp32
    probs = torch.empty((b, t, t), dtype=torch.float32, device=qkv.device)
    with P.region("attn_fwd", 4.0 * b * t * t * c, 0.0):
        check(lib().mvae_attn_core_fwd(ptr(qkv), ptr(out), ptr(probs), b, t, c, code, stream()), "attn_core_fwd")
    return out, ("probs", probs)


def attn_bwd(qkv, saved, dout):
    b, h, w, c3 = qkv.shape
    c, t = c3 // 3, h * w
    dqkv = torch.empty_like(qkv)
    if saved[0] == "fused":
        _tag, out, lse2 = saved
        rowdot = scratch("attn_rowdot", 4 * b * t, qkv.device)
        with P.region("attn_bwd", 10.0 * b * t * t * c, (2 * qkv.numel() + 2 * dout.numel()) * 2):
            check(lib().mvae_attn_fused_bwd(ptr(qkv), ptr(out), ptr(lse2), ptr(dout), ptr(dqkv), ptr(rowdot), b, t, c,
                                            L.code_of(qkv.dtype), stream()), "attn_fused_bwd")
        return dqkv
    ds = scratch("attn_ds", 4 * b * t * t, qkv.device)
    with P.region("attn_bwd", 8.0 * b * t * t * c, 0.0):
        check(lib().mvae_attn_core_bwd(ptr(qkv), ptr(saved[1]), ptr(dout), ptr(dqkv), ptr(ds), b, t, c,
                                       L.code_of(qkv.dtype), stream()), "attn_core_bwd")
    return dqkv


def input_pack_fwd(x_nchw, cond, w, b, out_c, dtype):
    bsz, c, h, wd_ = x_nchw.shape
    out = torch.empty((bsz, h, wd_, out_c), dtype=dtype, device=x_nchw.device)
    cond_dim = cond.shape[1] if cond is not None else 0
    check(lib().mvae_input_pack_fwd(ptr(x_nchw), ptr(cond), ptr(w), ptr(b), ptr(out), bsz, c, h, wd_, cond_dim,
                                    out_c, L.code_of(dtype), stream()), "input_pack_fwd")
    return out


def input_pack_bwd(d_in, cond, w, b, c):
    bsz, h, wd_, out_c = d_in.shape
    dw = torch.empty_like(w)
    db = torch.empty_like(b)
    check(lib().mvae_input_pack_bwd(ptr(d_in), ptr(cond), ptr(w), ptr(b), ptr(dw), ptr(db), bsz, c, h, wd_,
                                    cond.shape[1], out_c, L.code_of(d_in.dtype), stream()), "input_pack_bwd")
    return dw, db


def nhwc_to_nchw_f32(src, c_off=0, c_dst=None):
    b, h, w, c = src.shape
    c_dst = c - c_off if c_dst is None else c_dst
    dst = torch.empty((b, c_dst, h, w), dtype=torch.float32, device=src.device)
    check(lib().mvae_nhwc_to_nchw_f32(ptr(src), ptr(dst), b, h, w, c, c_off, c_dst, L.code_of(src.dtype), stream()),
          "nhwc_to_nchw_f32")
    return dst


def nchw_f32_to_nhwc(src, dtype):
    b, c, h, w = src.shape
    dst = torch.empty((b, h, w, c), dtype=dtype, device=src.device)
    check(lib().mvae_nchw_f32_to_nhwc(ptr(src), ptr(dst), b, h, w, c, L.code_of(dtype), stream()), "nchw_f32_to_nhwc")
    return dst


def reparam_kl_fwd(moments, eps, zc, clamp, seed, offset, want_kl=True, want_std=True, step_state=None, sample_base=0):
    """sample_base: index of this call's first sample in the global batch (rank and micro-batch start); the Philox
    element counter starts at sample_base * zc * r * r"""
    b, r, _, c2 = moments.shape
    dev = moments.device
    mean = torch.empty((b, zc, r, r), dtype=torch.float32, device=dev)
    logvar = torch.empty_like(mean)
    z = torch.empty_like(mean)
    std = torch.empty_like(mean) if want_std else None
    eps_out = None
    if eps is None:
        eps_out = torch.empty_like(mean)
    kl = torch.zeros((), dtype=torch.float64, device=dev) if want_kl else None
    # algorithmic bytes: read the moments, write z (the API also returns fp32 mean / logvar / std / eps)
    with P.region("loss_kernels", 0.0, moments.numel() * moments.element_size() + 4 * mean.numel()):
        check(lib().mvae_reparam_kl_fwd(ptr(moments), ptr(eps), ptr(eps_out), ptr(mean), ptr(logvar), ptr(std), ptr(z),
                                        None, ptr(kl), b, zc, r, float(clamp), int(seed), int(offset),
                                        int(sample_base) * zc * r * r, ptr(step_state),
                                        L.code_of(moments.dtype), stream()), "reparam_kl_fwd")
    return mean, logvar, z, std, (eps if eps is not None else eps_out), kl


def reparam_kl_bwd(d_mean, d_logvar, d_z, logvar, eps, moments, clamp):
    b, r, _, c2 = moments.shape
    d_mom = torch.empty_like(moments)
    with P.region("loss_kernels", 0.0, 2 * moments.numel() * moments.element_size() + 4 * (moments.numel() // 2)):
        check(lib().mvae_reparam_kl_bwd(ptr(d_mean), ptr(d_logvar), ptr(d_z), None, ptr(logvar), ptr(eps), ptr(moments),
                                        ptr(d_mom), b, c2 // 2, r, float(clamp), L.code_of(moments.dtype), stream()),
              "reparam_kl_bwd")
    return d_mom


def recon_loss(recon, x, kind, want_grad=True):
    n = recon.numel()
    s = torch.zeros((), dtype=torch.float64, device=recon.device)
    d_unit = torch.empty_like(recon) if want_grad else None
    with P.region("loss_kernels", 0.0, (3 if want_grad else 2) * 4 * n):
        check(lib().mvae_recon_loss_fwd_bwd(ptr(recon), ptr(x), ptr(s), ptr(d_unit), n, kind, stream()), "recon_loss")
    return s, d_unit


def kl_sum(mean, logvar):
    s = torch.zeros((), dtype=torch.float64, device=mean.device)
    check(lib().mvae_kl_fwd(ptr(mean), ptr(logvar), ptr(s), mean.numel(), stream()), "kl_fwd")
    return s


def kl_bwd(mean, logvar, coef):
    dm, dl = torch.empty_like(mean), torch.empty_like(logvar)
    with P.region("loss_kernels", 0.0, 16 * mean.numel()):
        check(lib().mvae_kl_bwd(ptr(mean), ptr(logvar), ptr(coef), ptr(dm), ptr(dl), mean.numel(), stream()), "kl_bwd")
    return dm, dl


def scale_by(t, scale_dev):
    out = torch.empty_like(t)
    with P.region("loss_kernels", 0.0, 8 * t.numel()):
        check(lib().mvae_scale_by_device_scalar(ptr(t), ptr(scale_dev), ptr(out), t.numel(), stream()), "scale")
    return out


# ------------------------------------------------------------------ disentangled model extras
def dis_input_fwd(x_nchw, idx, pw, pb, is_gray, out_c, dtype):
    b, xc, h, w = x_nchw.shape
    out = torch.empty((b, h, w, out_c), dtype=dtype, device=x_nchw.device)
    check(lib().mvae_dis_input_fwd(ptr(x_nchw), ptr(idx), ptr(pw), ptr(pb), ptr(is_gray), ptr(out), b, xc, h, w,
                                   pw.shape[0], out_c, L.code_of(dtype), stream()), "dis_input_fwd")
    return out


def dis_input_bwd(d_in, x_nchw, idx, is_gray, n_mod):
    b, xc, h, w = x_nchw.shape
    dpw = torch.empty((n_mod, 3), dtype=torch.float32, device=x_nchw.device)
    dpb = torch.empty((n_mod, 3), dtype=torch.float32, device=x_nchw.device)
    check(lib().mvae_dis_input_bwd(ptr(d_in), ptr(x_nchw), ptr(idx), ptr(is_gray), ptr(dpw), ptr(dpb), b, xc, h, w, n_mod,
                                   d_in.shape[-1], L.code_of(d_in.dtype), stream()), "dis_input_bwd")
    return dpw, dpb


def dis_head_fwd(r_nhwc, idx, w1, b1, w2, b2, pw, pb, is_gray, out_c):
    b, h, w, rc = r_nhwc.shape
    dev = r_nhwc.device
    hid = torch.empty((b, h, w, 3), dtype=torch.float32, device=dev)
    y = torch.empty((b, h, w, 3), dtype=torch.float32, device=dev)
    out = torch.empty((b, out_c, h, w), dtype=torch.float32, device=dev)
    check(lib().mvae_dis_head_fwd(ptr(r_nhwc), rc, ptr(idx), ptr(w1), ptr(b1), ptr(w2), ptr(b2), ptr(pw), ptr(pb),
                                  ptr(is_gray), ptr(hid), ptr(y), ptr(out), b, h, w, w1.shape[0], out_c,
                                  L.code_of(r_nhwc.dtype), stream()), "dis_head_fwd")
    return out, hid, y


def dis_head_bwd(d_out, r_nhwc, hid, y, idx, w1, w2, pw, is_gray):
    b, h, w, rc = r_nhwc.shape
    n_mod = w1.shape[0]
    d_r = torch.zeros_like(r_nhwc)
    grads = torch.empty((n_mod, 172), dtype=torch.float32, device=r_nhwc.device)
    sc = scratch("dis_head", 4 * 6 * b * h * w, r_nhwc.device)
    check(lib().mvae_dis_head_bwd(ptr(d_out), ptr(r_nhwc), rc, ptr(hid), ptr(y), ptr(idx), ptr(w1), ptr(w2), ptr(pw),
                                  ptr(is_gray), ptr(d_r), ptr(grads), ptr(sc), b, h, w, n_mod, d_out.shape[1],
                                  L.code_of(r_nhwc.dtype), stream()), "dis_head_bwd")
    return d_r, grads


def dis_aux_fwd(z_flat, idx, shared, modal, n_mod, tau):
    b, d = z_flat.shape
    aux = torch.empty((lib().mvae_dis_aux_floats(b, modal, n_mod),), dtype=torch.float32, device=z_flat.device)
    check(lib().mvae_dis_aux_loss_fwd(ptr(z_flat), ptr(idx), ptr(aux), b, d, shared, modal, n_mod, float(tau), stream()),
          "dis_aux_loss_fwd")
    return aux


def dis_aux_bwd(z_flat, idx, aux, g_sep, g_con, shared, modal, n_mod, tau):
    b, d = z_flat.shape
    dz = torch.zeros_like(z_flat)
    check(lib().mvae_dis_aux_loss_bwd(ptr(z_flat), ptr(idx), ptr(aux), ptr(g_sep), ptr(g_con), ptr(dz), b, d, shared,
                                      modal, n_mod, float(tau), stream()), "dis_aux_loss_bwd")
    return dz


def dis_total(recon, kl, sep, con, weights):
    out5 = torch.empty((5,), dtype=torch.float32, device=recon.device)
    mask4 = torch.empty((4,), dtype=torch.float32, device=recon.device)
    check(lib().mvae_dis_loss_total(ptr(recon), ptr(kl), ptr(sep), ptr(con), *[float(w) for w in weights], ptr(out5),
                                    ptr(mask4), stream()), "dis_loss_total")
    return out5, mask4


# ------------------------------------------------------------------------------------------------
# callers either side of the training step: linear attention, dropout, FiLM, batch feed,
# validation metrics
# ------------------------------------------------------------------------------------------------
def linattn_fwd(qkv):
    """qkv NHWC [B,H,W,3C] -> (out [B,H,W,C], ks fp32 [B,T,C], ctx fp32 [B,C,C])"""
    b, h, w, c3 = qkv.shape
    c, t = c3 // 3, h * w
    out = torch.empty((b, h, w, c), dtype=qkv.dtype, device=qkv.device)
    ks = torch.empty((b, t, c), dtype=torch.float32, device=qkv.device)
    ctx = torch.empty((b, c, c), dtype=torch.float32, device=qkv.device)
    check(lib().mvae_linattn_core_fwd(ptr(qkv), ptr(out), ptr(ks), ptr(ctx), b, t, c, L.code_of(qkv.dtype), stream()),
          "linattn_core_fwd")
    return out, ks, ctx


def linattn_bwd(qkv, ks, ctx, dout):
    b, h, w, c3 = qkv.shape
    c, t = c3 // 3, h * w
    dqkv = torch.empty_like(qkv)
    dctx = scratch("linattn_dctx", 4 * b * c * c, qkv.device)
    dks = scratch("linattn_dks", 4 * b * t * c, qkv.device)
    check(lib().mvae_linattn_core_bwd(ptr(qkv), ptr(ks), ptr(ctx), ptr(dout), ptr(dqkv), ptr(dctx), ptr(dks), b, t, c,
                                      L.code_of(qkv.dtype), stream()), "linattn_core_bwd")
    return dqkv


def dropout_apply(x, p, seed, offset, mask_in=None, want_mask=False, step_state=None, elem_base=0):
    """-> (y, mask or None); x any contiguous bf16/fp32 tensor, mask uint8 of the same shape.  elem_base: index of x's
    first element in the global batch (first sample * elements per sample)"""
    y = torch.empty_like(x)
    mask = torch.empty(x.shape, dtype=torch.uint8, device=x.device) if want_mask else None
    check(lib().mvae_dropout_apply(ptr(x), ptr(y), x.numel(), float(p), int(seed) & 0xFFFFFFFFFFFFFFFF, int(offset),
                                   int(elem_base), ptr(step_state), ptr(mask_in), ptr(mask), L.code_of(x.dtype), stream()),
          "dropout_apply")
    return y, mask


def film_fwd(features, cond, ws, bs, wt, bt):
    b, c, h, w = features.shape
    out = torch.empty_like(features)
    st = torch.empty((2, b, c), dtype=torch.float32, device=features.device)
    check(lib().mvae_film_fwd(ptr(features), ptr(cond), ptr(ws), ptr(bs), ptr(wt), ptr(bt), ptr(out), ptr(st), b, c, h * w,
                              cond.shape[1], stream()), "film_fwd")
    return out, st


def film_bwd(d_out, features, cond, ws, wt, st, need_dcond):
    b, c, h, w = features.shape
    k = cond.shape[1]
    dev = features.device
    df = torch.empty_like(features)
    dst = torch.empty((2, b, c), dtype=torch.float32, device=dev)
    dws, dwt = torch.empty((c, k), dtype=torch.float32, device=dev), torch.empty((c, k), dtype=torch.float32, device=dev)
    dbs, dbt = torch.empty((c,), dtype=torch.float32, device=dev), torch.empty((c,), dtype=torch.float32, device=dev)
    dcond = torch.empty((b, k), dtype=torch.float32, device=dev) if need_dcond else None
    check(lib().mvae_film_bwd(ptr(d_out), ptr(features), ptr(cond), ptr(ws), ptr(wt), ptr(st), ptr(df), ptr(dst), ptr(dws),
                              ptr(dbs), ptr(dwt), ptr(dbt), ptr(dcond), b, c, h * w, k, stream()), "film_bwd")
    return df, dws, dbs, dwt, dbt, dcond


def batch_feed_u8(src, src_c, tgt_c, idx, out_c, n_mod):
    """src uint8 [B,H,W,SC] (device) -> (x fp32 NCHW [B,out_c,H,W], onehot fp32 [B,n_mod] or None)"""
    b, h, w, sc = src.shape
    x = torch.empty((b, out_c, h, w), dtype=torch.float32, device=src.device)
    onehot = torch.empty((b, n_mod), dtype=torch.float32, device=src.device) if n_mod > 0 else None
    check(lib().mvae_batch_feed_u8(ptr(src), ptr(src_c), ptr(tgt_c), ptr(idx), ptr(x), ptr(onehot), b, h, w, sc, out_c,
                                   max(n_mod, 0), stream()), "batch_feed_u8")
    return x, onehot


def recon_metrics(original, recon, data_range, out):
    """original/recon NCHW fp32 contiguous; out: fp32 device view of 4 elements (mse, mae, psnr, ssim)"""
    b, c, h, w = original.shape
    ws = scratch("metrics_ws", 8 * lib().mvae_recon_metrics_ws_doubles(b * c, h, w), original.device)
    check(lib().mvae_recon_metrics(ptr(original), ptr(recon), b * c, h, w, float(data_range), ptr(ws), ptr(out), stream()),
          "recon_metrics")


def kl_metrics(mean, logvar, out):
    """mean/logvar fp32 contiguous, >= 2-D; out: 4 floats (kl_total, kl_mean, kl_std, kl_per_dim_mean)"""
    b, c = mean.shape[0], mean.shape[1]
    s = mean.numel() // (b * c)
    ws = scratch("metrics_ws_kl", 8 * lib().mvae_stat_metrics_ws_doubles(), mean.device)
    check(lib().mvae_kl_metrics(ptr(mean), ptr(logvar), b, c, s, ptr(ws), ptr(out), stream()), "kl_metrics")


def latent_metrics(z, thr, out):
    """z fp32 [B, D] contiguous; out: 3 floats (latent_mean_avg, latent_std_avg, latent_sparsity)"""
    ws = scratch("metrics_ws_lat", 8 * lib().mvae_stat_metrics_ws_doubles(), z.device)
    check(lib().mvae_latent_metrics(ptr(z), z.shape[0], z.shape[1], float(thr), ptr(ws), ptr(out), stream()),
          "latent_metrics")


def step_state_advance(step_state, rng_increment: float):
    """step_state: float64 device tensor {optimizer step, learning rate, Philox position}"""
    check(lib().mvae_step_state_advance(ptr(step_state), float(rng_increment), stream()), "step_state_advance")
