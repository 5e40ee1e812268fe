"""ctypes binding of libmedvae_b200.so (the C ABI in include/medvae_b200.h).

There is no CPU or eager-PyTorch fallback: if the library is missing, or a compute entry point
is called without a Blackwell device, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmedvae_b200.so")

ABI_VERSION = 5          # include/medvae_b200.h MVAE_ABI_VERSION
F32, BF16 = 0, 1
CONV_S1, CONV_DOWN, CONV_UP = 0, 1, 2
ENGINE_AUTO, ENGINE_SIMT, ENGINE_UMMA = 0, 1, 2


class MvaeError(RuntimeError):
    pass


class ConvDesc(C.Structure):
    _fields_ = [("batch", C.c_int32), ("in_h", C.c_int32), ("in_w", C.c_int32), ("in_c", C.c_int32),
                ("out_c", C.c_int32), ("ksize", C.c_int32), ("mode", C.c_int32), ("dtype", C.c_int32),
                ("engine", C.c_int32)]


_lib = None
_vp, _i32, _i64, _f32, _u64, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_uint64, C.c_size_t
_dp = C.POINTER(ConvDesc)

# name -> (restype, argtypes).  Every symbol declared in include/medvae_b200.h is listed here;
# tests/test_abi.py cross-checks the two.
SIGNATURES = {
    "mvae_last_error": (C.c_char_p, []),
    "mvae_abi_version": (C.c_int, []),
    "mvae_device_ok": (C.c_int, []),
    "mvae_launch_count": (C.c_uint64, []),
    "mvae_set_debug": (C.c_int, [C.c_int, C.c_int]),
    "mvae_conv2d_out_hw": (C.c_int, [_dp, C.POINTER(_i32), C.POINTER(_i32)]),
    "mvae_conv2d_num_slabs": (C.c_int, [_dp]),
    "mvae_conv2d_uses_umma": (C.c_int, [_dp]),
    "mvae_conv2d_pack_weights": (C.c_int, [_dp, _vp, _vp, _vp, _i32, _i32, _vp]),
    "mvae_conv2d_pack_weights_batched": (C.c_int, [_vp, _i32, _i32, _i32, _vp]),
    "mvae_conv2d_fprop": (C.c_int, [_dp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "mvae_conv2d_dgrad": (C.c_int, [_dp, _vp, _vp, _vp, _vp]),
    "mvae_conv2d_wgrad_workspace_bytes": (_sz, [_dp]),
    "mvae_conv2d_wgrad": (C.c_int, [_dp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "mvae_conv2d_wgrad_acc": (C.c_int, [_dp, _vp, _vp, _vp, _vp, _vp, _sz, _f32, _vp]),
    "mvae_groupnorm_splits": (C.c_int, [_i32, _i32, _i32]),
    "mvae_groupnorm_scratch_floats": (C.c_int64, [_i32, _i32, _i32, _i32, _i32]),
    "mvae_groupnorm_silu_fwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _f32, _i32, _i32, _vp]),
    "mvae_groupnorm_silu_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32,
                                          _i32, _i32, _vp]),
    "mvae_conv2d_fuses_gn": (C.c_int, [_dp, _i32, _i32]),
    "mvae_conv2d_fprop_gn": (C.c_int, [_dp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp]),
    "mvae_conv2d_dgrad_gn": (C.c_int, [_dp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp]),
    "mvae_groupnorm_fused_ok": (C.c_int, [_i32, _i32]),
    "mvae_groupnorm_apply_bwd_scratch_floats": (C.c_int64, [_i32, _i32, _i32]),
    "mvae_groupnorm_apply_fwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _f32, _i32, _i32, _vp]),
    "mvae_groupnorm_apply_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32,
                                           _vp]),
    "mvae_attn_core_fwd": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "mvae_attn_core_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "mvae_attn_fused_ok": (C.c_int, [_i32, _i32, _i32]),
    "mvae_attn_fused_fwd": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "mvae_attn_fused_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "mvae_attn_fused_bwd_static": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "mvae_input_pack_fwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "mvae_input_pack_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "mvae_nhwc_to_nchw_f32": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "mvae_nchw_f32_to_nhwc": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp]),
    "mvae_reparam_kl_fwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _f32, _u64, _u64,
                                      _u64, _vp, _i32, _vp]),
    "mvae_reparam_kl_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _f32, _i32, _vp]),
    "mvae_recon_loss_fwd_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i32, _vp]),
    "mvae_kl_fwd": (C.c_int, [_vp, _vp, _vp, _i64, _vp]),
    "mvae_kl_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _vp]),
    "mvae_scale_by_device_scalar": (C.c_int, [_vp, _vp, _vp, _i64, _vp]),
    "mvae_grad_sqnorm": (C.c_int, [_vp, _vp, _i32, _vp, _vp]),
    "mvae_adamw_clip_step": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _vp, _f32, _f32, _f32, _f32, _f32, _f32, _i32,
                                       _i32, _vp]),
    "mvae_grad_sqnorm_chunked": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp]),
    "mvae_adamw_clip_step_chunked": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i32, _vp,
                                               _f32, _f32, _f32, _f32, _f32, _f32, _i32, _i32, _vp, _vp]),
    "mvae_step_state_advance": (C.c_int, [_vp, C.c_double, _vp]),
    "mvae_dis_input_fwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "mvae_dis_input_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "mvae_dis_head_fwd": (C.c_int, [_vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32,
                                    _i32, _i32, _vp]),
    "mvae_dis_head_bwd": (C.c_int, [_vp, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32,
                                    _i32, _i32, _vp]),
    "mvae_dis_aux_floats": (C.c_int64, [_i32, _i32, _i32]),
    "mvae_dis_aux_loss_fwd": (C.c_int, [_vp, _vp, _vp, _i32, _i64, _i32, _i32, _i32, _f32, _vp]),
    "mvae_dis_aux_loss_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i64, _i32, _i32, _i32, _f32, _vp]),
    "mvae_dis_loss_total": (C.c_int, [_vp, _vp, _vp, _vp, _f32, _f32, _f32, _f32, _vp, _vp, _vp]),
    "mvae_linattn_core_fwd": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "mvae_linattn_core_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "mvae_dropout_apply": (C.c_int, [_vp, _vp, _i64, _f32, _u64, _u64, _u64, _vp, _vp, _vp, _i32, _vp]),
    "mvae_film_fwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "mvae_film_bwd": (C.c_int, [_vp] * 13 + [_i32, _i32, _i32, _i32, _vp]),
    "mvae_batch_feed_u8": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "mvae_recon_metrics_ws_doubles": (C.c_int64, [_i32, _i32, _i32]),
    "mvae_stat_metrics_ws_doubles": (C.c_int64, []),
    "mvae_recon_metrics": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _f32, _vp, _vp, _vp]),
    "mvae_kl_metrics": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp]),
    "mvae_latent_metrics": (C.c_int, [_vp, _i32, _i64, _f32, _vp, _vp, _vp]),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MvaeError(
                f"{LIB_PATH} is missing: build it with `python -m medvae_disentangled_multimodal_b200.build` "
                "(there is no fallback path)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def last_error() -> str:
    return lib().mvae_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = ""):
    if rc != 0:
        raise MvaeError(f"{what}: {last_error()}")


def require_device():
    if not torch.cuda.is_available():
        raise MvaeError("medvae_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    if not lib().mvae_device_ok():
        raise MvaeError("medvae_b200 kernels are built for sm_100a (B200) only")


def ptr(t):
    """device pointer of a tensor (None -> NULL)"""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def code_of(dtype: torch.dtype) -> int:
    if dtype == torch.float32:
        return F32
    if dtype == torch.bfloat16:
        return BF16
    raise MvaeError(f"unsupported activation dtype {dtype}")
