"""ctypes binding of libncnet_b200.so (the C ABI declared in include/ncnet_b200.h).

There is no CPU or PyTorch fallback: if the shared library cannot be loaded (and cannot be
built because nvcc is absent) importing this module's ``lib()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

_lock = threading.Lock()
_lib = None

F32, F16, CB16, CB16PM = 0, 1, 2, 3
W_OIKKKK, W_KOIKKK = 0, 1
CONV_FP32, CONV_TC16 = 0, 1

_p, _i, _ll, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_size_t

# name -> (restype, argtypes); mirrors include/ncnet_b200.h one to one
SIGNATURES = {
    "ncnet_version": (_i, []),
    "ncnet_last_error": (C.c_char_p, []),
    "ncnet_feature_l2norm": (_i, [_p, _p, _i, _i, _i, _i, _p]),
    "ncnet_corr4d_fwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "ncnet_corr4d_tc_scratch_bytes": (_sz, [_i] * 4),
    "ncnet_corr4d_fwd_tc": (_i, [_p, _p, _p, _p] + [_i] * 7 + [_p]),
    "ncnet_corr4d_pool_fwd_tc": (_i, [_p] * 8 + [_i] * 10 + [_p]),
    "ncnet_corr4d_bwd": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ncnet_mutual_matching_keys_bytes": (_sz, [_i, _i, _i]),
    "ncnet_mutual_matching_sums_bytes": (_sz, [_i, _i, _i]),
    "ncnet_mutual_matching_fwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ncnet_mutual_matching_fwd_infer": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ncnet_mutual_matching_max": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "ncnet_mutual_matching_apply": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ncnet_mutual_matching_bwd": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "ncnet_mutual_matching_fwd_noalloc": (_i, [_p, _p, _i, _i, _i, _i, _i, _p]),
    "ncnet_match_score_fwd": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p]),
    "ncnet_match_score_bwd": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p]),
    "ncnet_conv4d_scratch_bytes": (_sz, [_i] * 10),
    "ncnet_conv4d_fwd": (_i, [_p, _p, _p, _p, _p] + [_i] * 15 + [_p]),
    "ncnet_conv4d_dgrad": (_i, [_p, _p, _p, _p, _p] + [_i] * 15 + [_p]),
    "ncnet_conv4d_wgrad_scratch_bytes": (_sz, [_i] * 10),
    "ncnet_conv4d_wgrad": (_i, [_p, _p, _p, _p, _p] + [_i] * 14 + [_p, _p]),
    "ncnet_conv4d_tc_supported": (_i, [_i] * 5),
    "ncnet_conv4d_tc_fwd_supported": (_i, [_i] * 5),
    "ncnet_cb16_bytes": (_sz, [_i] * 6),
    "ncnet_cb16_pack": (_i, [_p, _p] + [_i] * 8 + [_p]),
    "ncnet_cb16_unpack": (_i, [_p, _p] + [_i] * 8 + [_p]),
    "ncnet_relu_mask_bcast": (_i, [_p, _p, _p, _i, _i, _ll, _i, _p, _p]),
    "ncnet_grad_scale": (_i, [_p, _ll, _i, _p, _p]),
    "ncnet_maxpool4d_fwd": (_i, [_p] * 6 + [_i] * 7 + [_p]),
    "ncnet_matches_unique_scratch_bytes": (_sz, [_i]),
    "ncnet_matches_sort_unique": (_i, [_p] * 5 + [_i] + [_p] * 5 + [_i] * 5 + [_p] * 4),
    "ncnet_points_axis_transform": (_i, [_p, _p, _i, _p, _i, _i, _i, _p]),
    "ncnet_bilinear_point_tnf": (_i, [_p] * 5 + [_i, _p, _p, _i, _i, _p]),
    "ncnet_pck": (_i, [_p, _p, _p, C.c_float, _p, _i, _i, _p]),
    "ncnet_corr_to_matches_scratch_bytes": (_sz, [_i] * 7),
    "ncnet_corr_to_matches": (_i, [_p] + [_i] * 8 + [_p] * 4 + [_i] + [_p] * 4 + [_p] * 9 + [_p, _i, _p]),
}


def library_path() -> str:
    return _build.LIB


def lib() -> C.CDLL:
    """Load (building first if the in-tree .so is stale or absent) and type the library."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = _build.LIB
        if not _build.up_to_date():
            nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
            if os.path.exists(nvcc):
                _build.build()
            elif not os.path.exists(path):
                raise RuntimeError(
                    f"ncnet_b200: {path} is missing and nvcc is not available to build it; "
                    "there is no CPU fallback (run `python -m ncnet_b200.build`)")
        handle = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError if the library lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = handle
        return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().ncnet_last_error()
        raise RuntimeError(f"{what}: {msg.decode() if msg else 'error'}")
