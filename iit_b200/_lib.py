"""ctypes binding of libiit_b200.so (the C-ABI declared in include/iit_b200.h).

There is no fallback: if the shared library is missing or a symbol is absent, importing this module
raises. Every entry point returns an int status; non-zero raises RuntimeError(iit_last_error()).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# IIT_B200_LIB: another build of the same library (e.g. a kernel variant for a same-machine A/B measurement)
LIB_PATH = os.environ.get("IIT_B200_LIB") or os.path.join(_HERE, "libiit_b200.so")

IIT_MAX_RANK = 6
IIT_F32, IIT_BF16 = 0, 1
IOI_NODE_IDS = {None: 0, "all_nodes_hook": 1, "hook_duplicate": 2, "hook_s_inhibition": 3, "hook_name_mover": 4}


class Box(C.Structure):
    _fields_ = [
        ("rank", C.c_int32),
        ("elem_bytes", C.c_int32),
        ("start", C.c_int64 * IIT_MAX_RANK),
        ("extent", C.c_int64 * IIT_MAX_RANK),
        ("dst_stride", C.c_int64 * IIT_MAX_RANK),
        ("src_stride", C.c_int64 * IIT_MAX_RANK),
    ]


class Conv(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n", "h", "w", "c", "k", "r", "s", "stride", "pad", "p", "q")]


class Relayout(C.Structure):
    """iit_relayout_t"""
    _fields_ = [("w", C.c_void_p), ("out", C.c_void_p), ("stride", C.c_int64 * 4),
                ("k", C.c_int32), ("c", C.c_int32), ("r", C.c_int32), ("s", C.c_int32),
                ("to_crsk", C.c_int32), ("dtype", C.c_int32), ("out_pitch", C.c_int64)]


class WgradMap(C.Structure):
    """iit_wgrad_map_t"""
    _fields_ = [("dst", C.c_void_p), ("size", C.c_int32 * 3), ("reserved", C.c_int32), ("stride", C.c_int64 * 3), ("offset", C.c_int64)]


class HlSpec(C.Structure):
    """iit_hl_spec_t"""
    _fields_ = [("mode", C.c_int32), ("node", C.c_int32), ("base", C.c_void_p), ("src", C.c_void_p), ("names", C.c_void_p),
                ("n_names", C.c_int32), ("seq", C.c_int32), ("d_vocab", C.c_int64)]


_vp, _i, _i64, _f, _d, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double, C.c_size_t
_BoxP, _ConvP = C.POINTER(Box), C.POINTER(Conv)

# name -> argtypes (restype is int unless listed in _RESTYPES). Mirrors include/iit_b200.h one to one.
SIGNATURES = {
    "iit_patch_fwd": [_vp, _vp, _BoxP, _vp],
    "iit_patch_bwd": [_vp, _BoxP, _vp],
    "iit_capture_box": [_vp, _vp, _BoxP, _vp],
    "iit_hl_pvr": [_vp, _vp, _i, _i64, _vp, _vp, _vp],
    "iit_hl_ioi_last_argmax": [_vp, _vp, _i, _vp, _i, _i64, _i, _i64, _vp, _vp],
    "iit_hl_ioi_dense": [_vp, _vp, _i, _vp, _i, _i64, _i, _i64, _vp, _vp],
    "iit_hl_ce_fwd": [_vp, _i, _i64, _i64, _i64, C.POINTER(HlSpec), _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "iit_ce_index_fwd": [_vp, _i, _i64, _i64, _i64, _vp, _vp, _vp, _vp],
    "iit_ce_index_bwd": [_vp, _i, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _i64, _vp],
    "iit_ce_prob_fwd": [_vp, _i, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp],
    "iit_ce_prob_bwd": [_vp, _i, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _i64, _vp],
    "iit_iia_count": [_vp, _i, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp],
    "iit_conv2d_fprop": [_vp, _vp, _vp, _ConvP, _i, _vp, _vp, _vp],
    "iit_conv2d_dgrad": [_vp, _vp, _vp, _ConvP, _i, _vp],
    "iit_conv2d_tc_supported": [_ConvP, _i],
    "iit_conv2d_fprop_tc": [_vp, _vp, _vp, _ConvP, _vp, _vp, _vp],
    "iit_conv2d_dgrad_tc": [_vp, _vp, _vp, _ConvP, _vp],
    "iit_conv2d_dgrad_tc_accumulate": [_vp, _vp, _vp, _ConvP, _vp],
    "iit_conv2d_patch_fusable": [_ConvP],
    "iit_conv2d_fprop_tc_patch": [_vp, _vp, _vp, _ConvP, _vp, _vp, _vp, C.POINTER(C.c_int32), _vp],
    "iit_bn_act_bwd_apply_box": [_vp, _vp, _vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, C.POINTER(C.c_int32), _vp],
    "iit_gemm_bias_act_tc": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp],
    "iit_gemm_dgrad_gelu_tc": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _vp],
    "iit_conv2d_wgrad_tc_scratch_bytes": [_ConvP],
    "iit_conv2d_wgrad_tc": [_vp, _vp, _vp, _ConvP, _vp, _sz, _vp],
    "iit_conv2d_wgrad_tc_accumulate": [_vp, _vp, _vp, _ConvP, _vp, _sz, _vp],
    "iit_conv2d_fprop_tc_affine": [_vp, _vp, _vp, _ConvP, _vp, _vp, _vp, _i, _vp],
    "iit_bn_eval_affine": [_vp, _vp, _vp, _vp, _i, _f, _vp, _vp, _vp],
    "iit_conv2d_wgrad_tc_scatter": [_vp, _vp, _ConvP, _vp, _sz, _vp, _i, _i64, _vp],
    "iit_bn_param_grads": [_vp, _i, _vp, _vp, _i, _i, _vp],
    "iit_conv2d_wgrad_scratch_bytes": [_ConvP],
    "iit_conv2d_wgrad": [_vp, _vp, _vp, _ConvP, _i, _vp, _sz, _vp],
    "iit_stem_im2col": [_vp, _i, C.POINTER(C.c_int64), _vp, _i, _i, _i, _vp],
    "iit_conv2d_wgrad_noalloc": [_vp, _vp, _vp, _ConvP, _i, _vp],
    "iit_filter_relayout": [_vp, C.POINTER(C.c_int64), _vp, _i, _i, _i, _i, _i, _i, _vp],
    "iit_bn_stats": [_vp, _i, _i64, _i, _vp, _vp, _vp],
    "iit_bn_finalize": [_vp, _vp, _i64, _i, _f, _f, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "iit_bn_eval_invstd": [_vp, _i, _f, _vp, _vp],
    "iit_bn_act_fwd": [_vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp],
    "iit_bn_act_bwd_reduce": [_vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp],
    "iit_bn_act_bwd_apply": [_vp, _vp, _vp, _vp, _vp, _i, _i64, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp],
    "iit_div_scalar": [_vp, _vp, _i, _i64, _f, _vp],
    "iit_relu_fwd": [_vp, _vp, _i, _i64, _vp],
    "iit_relu_bwd": [_vp, _vp, _vp, _i, _i64, _vp],
    "iit_add_relu_fwd": [_vp, _vp, _vp, _i, _i64, _i, _vp],
    "iit_maxpool3x3s2_fwd": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "iit_maxpool3x3s2_bwd": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "iit_avgpool_fwd": [_vp, _vp, _i, _i, _i, _i, _vp],
    "iit_avgpool_bwd": [_vp, _vp, _i, _i, _i, _i, _vp],
    "iit_linear_fwd": [_vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "iit_linear_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "iit_embed_fwd": [_vp, _vp, _vp, _vp, _i, _i64, _i, _i, _i64, _vp],
    "iit_embed_bwd": [_vp, _vp, _i, _vp, _vp, _i64, _i, _i, _i64, _vp],
    "iit_lnpre_fwd": [_vp, _vp, _vp, _i, _i64, _i, _f, _vp],
    "iit_lnpre_bwd": [_vp, _vp, _vp, _vp, _i, _i64, _i, _vp],
    "iit_lnpre_bwd_accumulate": [_vp, _vp, _vp, _vp, _i, _i64, _i, _vp],
    "iit_bias_act_fwd": [_vp, _vp, _vp, _vp, _vp, _i, _i64, _i, _i, _vp],
    "iit_gelu_bwd": [_vp, _vp, _vp, _i, _i64, _vp],
    "iit_gelu_bwd_bias_colsum": [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp],
    "iit_colsum": [_vp, _i, _vp, _i64, _i64, _vp],
    "iit_attn_fwd": [_vp, _vp, _i, _i64, _i, _i, _i, _vp],
    "iit_attn_bwd": [_vp, _vp, _vp, _i, _i64, _i, _i, _i, _vp],
    "iit_bn_relu_maxpool_fwd": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "iit_bn_relu_maxpool_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "iit_conv_win_enable": [_i],
    "iit_conv_pair_enable": [_i],
    "iit_conv_win_tma_store": [_i],
    "iit_stem_s2d": [_vp, _i, _vp, _vp, _i, _i, _i, _vp],
    "iit_stem_filter_s2d": [_vp, _vp, _vp, _vp],
    "iit_stem_conv_s2d": [_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp],
    "iit_stem_wgrad_s2d_scratch_bytes": [_i, _i, _i],
    "iit_stem_wgrad_s2d": [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, C.c_size_t, _vp],
    "iit_gather_rows": [_vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp],
    "iit_gather_box": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i64, _vp, _vp],
    "iit_pick_other_class": [_vp, _vp, _i, _i64, C.c_uint64, _i64, _vp, _vp, _vp, _vp],
    "iit_filter_relayout_batched": [_vp, _i, _i64, _vp],
    "iit_kl_div_rows": [_vp, _vp, _i64, _i64, _i64, _i64, _i, _i, _vp, _vp],
    "iit_sqnorm": [_vp, _i64, _vp, _vp],
    "iit_adam_step": [_vp, _vp, _vp, _vp, _i64, _d, _d, _d, _d, _i64, _vp, _d, _vp],
    "iit_adam_step_dev": [_vp, _vp, _vp, _vp, _i64, _vp, _d, _d, _d, _vp, _vp, _d, _vp],
    "iit_last_error": [],
    "iit_version": [],
    "iit_launch_count": [],
    "iit_kernel_variant_slots": [],
    "iit_kernel_variant_name": [_i],
    "iit_kernel_variant_count": [_i],
}
_RESTYPES = {"iit_last_error": C.c_char_p, "iit_launch_count": C.c_ulonglong,
             "iit_kernel_variant_name": C.c_char_p, "iit_kernel_variant_count": C.c_ulonglong,
             "iit_conv2d_wgrad_scratch_bytes": C.c_size_t, "iit_conv2d_wgrad_tc_scratch_bytes": C.c_size_t,
             "iit_stem_wgrad_s2d_scratch_bytes": C.c_size_t}
_NO_STATUS = set(_RESTYPES) | {"iit_version", "iit_kernel_variant_slots", "iit_conv2d_tc_supported", "iit_conv_win_enable", "iit_conv_pair_enable", "iit_conv_win_tma_store",
                               "iit_conv2d_patch_fusable"}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA extension was not built. Run `python -m iit_b200.build` "
            "(or __graft_entry__.build()). There is no CPU or PyTorch fallback for these kernels.")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, C.c_int)
    return lib


lib = _load()


def last_error() -> str:
    return lib.iit_last_error().decode()


def call(name: str, *args):
    """Invoke a status-returning entry point; raise on failure."""
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed (status {rc}): {last_error()}")


def launch_count() -> int:
    return int(lib.iit_launch_count())


def kernel_variant_counts() -> dict:
    """{kernel instantiation name: launches since load} for the tensor-core convolution dispatchers."""
    return {lib.iit_kernel_variant_name(i).decode(): int(lib.iit_kernel_variant_count(i))
            for i in range(int(lib.iit_kernel_variant_slots()))}
