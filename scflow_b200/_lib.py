"""ctypes binding of libscflow_b200.so (the C ABI declared in include/scflow_b200.h).

The product path has no CPU or library fallback: if the shared object is missing or a launcher
returns an error, this module raises.  Loading the library needs no GPU (cudart is linked
statically and initialises lazily), so symbol checks run in the CPU-only test tier.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libscflow_b200.so")

_P = c_void_p
ABI_VERSION = 7
# name -> (restype, argtypes); mirrors include/scflow_b200.h one to one
SIGNATURES = {
    "scflow_abi_version": (c_int, []),
    "scflow_strerror": (c_char_p, [c_int]),
    "scflow_last_cuda_error": (c_int, []),
    "scflow_corr_pyramid_workspace_bytes": (c_size_t, [c_int] * 5),
    "scflow_corr_pyramid_fwd": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int,
                                        _P, c_size_t, _P]),
    "scflow_corr_lookup_fwd": (c_int, [_P, c_int, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, _P]),
    "scflow_pack_nhwc": (c_int, [_P, c_int, _P, c_int, _P, c_int, _P, c_int, c_int, ctypes.c_longlong, c_int,
                                 c_int, _P]),
    "scflow_nhwc_to_nchw": (c_int, [_P, _P, c_int, c_int, c_int, c_int, _P]),
    "scflow_lift_points": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, _P]),
    "scflow_pose_flow_fwd": (c_int, [_P] * 14 + [c_int, c_int, c_int, c_int, c_float, c_float, _P]),
    "scflow_upsample_fwd": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_float, _P]),
    "scflow_downsample_fwd": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, c_float, _P]),
    "scflow_group_norm4_act_nhwc": (c_int, [_P, _P, _P, _P, _P, ctypes.c_longlong, c_int, c_int, c_int, c_int, c_float,
                                            c_int, _P]),
    "scflow_render_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "scflow_render_fwd": (c_int, [_P] * 6 + [c_int, c_int] + [_P] * 4 + [c_int, c_int, c_int, c_float, c_float, c_float]
                          + [_P] * 5),
    "scflow_add_error_fwd": (c_int, [_P, _P, c_int] + [_P] * 9 + [c_int, _P]),
    "scflow_expand_taps_fwd": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "scflow_pose_tail_fwd": (c_int, [_P] * 9 + [c_int, c_int, c_int, c_int, _P]),
    "scflow_pose_fc_workspace_floats": (c_size_t, [c_int, c_int, c_int, c_int]),
    "scflow_pose_fc_fwd": (c_int, [_P] * 13 + [c_int] * 6 + [_P]),
    "scflow_pose_fc_fwd_static": (c_int, [_P] * 12 + [c_int] * 6 + [_P]),
    "scflow_stem_s2d_split_fwd": (c_int, [_P, _P, c_int, c_int, c_int, c_int, _P]),
    "scflow_conv_small_cin_fwd": (c_int, [_P, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_longlong, _P, _P, _P,
                                          _P, ctypes.c_longlong, c_int,
                                          c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P]),
    "scflow_conv_small_cout_fwd": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, _P]),
    "scflow_conv_tc_fwd": (c_int, [_P, _P]),
    "scflow_conv_tc_packed_bytes": (c_size_t, [c_int] * 5),
    "scflow_conv_tc_pack_weights": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "scflow_split_planar": (c_int, [_P, _P, ctypes.c_longlong, c_int, c_int, c_int, ctypes.c_longlong, _P]),
    "scflow_norm_apply_split": (c_int, [_P, _P, _P, ctypes.c_longlong, c_int, _P, _P, _P, ctypes.c_longlong, c_int,
                                        _P, c_int, c_int, c_int, c_float, c_int, _P]),
    "scflow_split_bf16": (c_int, [_P, ctypes.c_longlong, _P, ctypes.c_longlong, c_int, c_int, ctypes.c_longlong, _P]),
    "scflow_corr_lookup_bwd": (c_int, [_P, c_int, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, _P]),
    "scflow_tap_sum_fwd": (c_int, [_P, c_int, c_int, _P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "scflow_corr_fold_bwd": (c_int, [_P, c_int, _P, c_int, c_int, c_int, c_int, _P]),
    "scflow_corr_pyramid_bwd": (c_int, [_P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, _P]),
    "scflow_upsample_bwd": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, c_float, _P]),
    "scflow_gru_reset_fwd": (c_int, [_P, _P, _P, ctypes.c_longlong, _P]),
    "scflow_gru_reset_bwd": (c_int, [_P, _P, _P, _P, _P, ctypes.c_longlong, _P]),
    "scflow_gru_blend_fwd": (c_int, [_P, _P, _P, _P, ctypes.c_longlong, _P]),
    "scflow_gru_blend_bwd": (c_int, [_P, _P, _P, _P, _P, _P, _P, ctypes.c_longlong, _P]),
    "scflow_pose_compose_fwd": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_float, _P]),
    "scflow_pose_compose_bwd": (c_int, [_P] * 10 + [c_int, c_float, c_int, _P]),
    "scflow_pose_flow_bwd": (c_int, [_P] * 8 + [c_int, c_int, c_int, _P]),
}

_lib = None


class ScflowError(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    """Load the in-tree shared object; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ScflowError(
            f"{LIB_PATH} not found: build it with `python -m scflow_b200.build` "
            "(scflow_b200 has no CPU or PyTorch fallback by design)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    if lib.scflow_abi_version() != ABI_VERSION:
        raise ScflowError("libscflow_b200.so ABI version mismatch; rebuild")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        lib = load()
        msg = lib.scflow_strerror(rc).decode()
        raise ScflowError(f"{what}: {msg} (rc={rc}, cudaError={lib.scflow_last_cuda_error()})")
