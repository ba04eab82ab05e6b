"""ctypes binding of libmelissa_b200.so (the C ABI in include/melissa_b200.h).

There is no CPU fallback: if the CUDA library is missing or a call fails the
caller gets an exception, never a silently different code path.
"""
from __future__ import annotations

import ctypes as C
import os

from .build import LIB_PATH

c_i32, c_i64, c_dbl = C.c_int32, C.c_int64, C.c_double
p_i32, p_i64, p_dbl, p_u8 = C.POINTER(c_i32), C.POINTER(c_i64), C.POINTER(c_dbl), C.POINTER(C.c_uint8)
handle_t = C.c_void_p

PROGRESS_FN = C.CFUNCTYPE(C.c_int, c_i32, c_dbl, c_dbl, C.c_void_p)

# name -> (restype, argtypes); must list every symbol the header declares.
SIGNATURES = {
    "melissa_last_error": (C.c_char_p, []),
    "melissa_abi_version": (C.c_int, []),
    "melissa_set_progress": (C.c_int, [PROGRESS_FN, C.c_void_p]),
    "melissa_pack": (C.c_int, [c_i32, c_i32, c_i64, p_i64, p_i32, p_i32, p_dbl, p_u8, c_i32, c_i32, p_dbl, c_dbl,
                               c_i32, c_i32, C.POINTER(handle_t)]),
    "melissa_pack_device": (C.c_int, [c_i32, c_i32, c_i64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      c_i32, c_i32, p_dbl, c_dbl, c_i32, c_i32, C.POINTER(handle_t)]),
    "melissa_set_option": (C.c_int, [C.c_char_p, c_i64]),
    "melissa_free": (C.c_int, [handle_t]),
    "melissa_profile_enable": (C.c_int, [handle_t, c_i32]),
    "melissa_profile_read": (C.c_int, [handle_t, c_i32, p_dbl, p_i64]),
    "melissa_vb_run": (C.c_int, [handle_t, c_i32, p_dbl, p_dbl, c_dbl, c_dbl, c_i32, c_dbl, c_i32,
                                 p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, c_i32,
                                 p_i32, p_i32, p_i32]),
    "melissa_pack_multi": (C.c_int, [c_i32, c_i32, c_i64, p_i64, p_i32, p_i32, p_dbl, p_u8, c_i32, c_i32, p_dbl, c_dbl,
                                     c_i32, p_i32, C.POINTER(handle_t)]),
    "melissa_multi_create": (C.c_int, [c_i32, C.POINTER(handle_t), C.POINTER(handle_t)]),
    "melissa_vb_run_multi": (C.c_int, [handle_t, c_i32, p_dbl, p_dbl, c_dbl, c_dbl, c_i32, c_dbl, c_i32,
                                       p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, c_i32,
                                       p_i32, p_i32, p_i32]),
    "melissa_multi_free": (C.c_int, [handle_t]),
    "melissa_multi_gpus": (c_i32, [handle_t]),
    "melissa_multi_last_loop_ms": (c_dbl, [handle_t]),
    "melissa_multi_launch_count": (c_i64, [handle_t]),
    "melissa_vb_begin": (C.c_int, [handle_t, c_i32, p_dbl, p_dbl, c_dbl, c_dbl, c_i64, C.c_void_p]),
    "melissa_vb_phase_a": (C.c_int, [handle_t]),
    "melissa_vb_phase_b": (C.c_int, [handle_t, c_i32]),
    "melissa_vb_lower_bound": (C.c_int, [handle_t, p_dbl]),
    "melissa_vb_reduce_buffer": (C.c_int, [handle_t, C.POINTER(C.c_void_p), p_i64]),
    "melissa_vb_lb_buffer": (C.c_int, [handle_t, C.POINTER(C.c_void_p), p_i64]),
    "melissa_vb_fetch": (C.c_int, [handle_t, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl, p_dbl]),
    "melissa_vb_fetch_HtEz": (C.c_int, [handle_t, p_dbl]),
    "melissa_profiles_mle": (C.c_int, [handle_t, c_dbl, c_i32, p_dbl]),
    "melissa_profiles_mle_device": (C.c_int, [handle_t, c_dbl, c_i32, C.c_void_p]),
    "melissa_sync": (C.c_int, [handle_t]),
    "melissa_launch_count": (c_i64, [handle_t]),
    "melissa_layout_info": (C.c_int, [handle_t, p_i64, c_i32]),
    "melissa_impute": (C.c_int, [c_i32, c_i32, c_i32, c_i32, c_i32, p_dbl, c_dbl, p_dbl, p_dbl, c_i64, p_i64, p_i32,
                                 p_i32, p_dbl, p_dbl, c_i32]),
    "melissa_impute_noalloc": (C.c_int, [c_i32, c_i32, c_i32, c_i32, c_i32, p_dbl, c_dbl, p_dbl, p_dbl, c_i64, p_i64,
                                         p_i32, p_i32, p_dbl, p_dbl, c_i32]),
}

_lib = None


class MelissaError(RuntimeError):
    """A non-zero status from the C ABI (message from melissa_last_error)."""

    def __init__(self, status: int, message: str):
        super().__init__(message)
        self.status = status


def load():
    """Load the shared library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(needs nvcc). melissa_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int):
    if status != 0:
        raise MelissaError(status, load().melissa_last_error().decode("utf-8", "replace"))


def ptr(arr, ctype):
    """ctypes pointer to a C-contiguous numpy array (None -> NULL)."""
    if arr is None:
        return None
    return arr.ctypes.data_as(C.POINTER(ctype))
