"""ctypes binding of libppm_sm100.so (the C ABI declared in include/ppm_b200.h).

The library is built in-tree by ``ppmstereo_b200/build.py``.  There is no fallback: if the
library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p, POINTER

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libppm_sm100.so")

PPM_F32, PPM_BF16 = 0, 1

# name -> (restype, argtypes); must list every symbol of include/ppm_b200.h
SIGNATURES = {
    "ppm_abi_version": (c_int, []),
    "ppm_last_error": (c_char_p, []),
    "ppm_corr_build": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                               POINTER(c_void_p), c_int, c_int, c_void_p]),
    "ppm_corr_lookup": (c_int, [POINTER(c_void_p), c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                c_void_p, c_void_p]),
    "ppm_corr_lookup_l0": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ppm_qk_similarity": (c_int, [c_void_p, c_void_p, c_int, c_int64, c_int64, c_int64, c_int, c_int, c_int, c_int,
                                  c_void_p, c_void_p, c_void_p]),
    "ppm_qam_pick": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p,
                             c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ppm_qam_pick_static": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p]),
    "ppm_pack_query": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p, c_int, c_int, c_int,
                               c_void_p, c_void_p]),
    "ppm_pack_key": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_int,
                             c_int, c_int, c_int, c_void_p, c_void_p]),
    "ppm_pack_value": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int64, c_int, c_int, c_int, c_void_p,
                               c_void_p]),
    "ppm_memory_attend": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64,
                                  c_int64, c_void_p, c_float, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ppm_memory_attend_band": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int64,
                                       c_int64, c_void_p, c_float, c_int, c_int, c_int, c_int, c_int, c_int,
                                       c_void_p, c_void_p]),
    "ppm_convex_upsample": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ppm_sepconv_gru3d_workspace": (c_int, [c_int, c_int, c_int, c_int, POINTER(c_size_t)]),
    "ppm_sepconv_gru3d": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p, c_int, c_int64, c_int64, c_int64,
                                  POINTER(c_void_p), POINTER(c_void_p), c_int, c_int, c_int, c_int, c_void_p, c_size_t,
                                  c_void_p, c_int, c_void_p]),
    "ppm_sepconv_gru3d_state_offset": (c_int, [c_int, c_int, c_int, c_int, POINTER(c_size_t)]),
    "ppm_flow_head3d_workspace": (c_int, [c_int, c_int, c_int, c_int, POINTER(c_size_t)]),
    "ppm_flow_head3d": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p, c_int, c_void_p]),
    "ppm_motion_corr_front": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                      c_void_p, c_void_p, c_void_p]),
    "ppm_motion_encoder_workspace": (c_int, [c_int, c_int, c_int, POINTER(c_size_t)]),
    "ppm_motion_encoder": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                   c_void_p, c_size_t, c_void_p, c_void_p, c_void_p]),
    "ppm_value_uncertainty_workspace": (c_int, [c_int, c_int, c_int, POINTER(c_size_t)]),
    "ppm_value_uncertainty": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p]),
    "ppm_update_block3d_workspace": (c_int, [c_int, c_int, c_int, c_int, POINTER(c_size_t)]),
    "ppm_update_block3d": (c_int, [c_void_p, c_int, POINTER(c_void_p), POINTER(c_int), POINTER(c_void_p), POINTER(c_void_p),
                                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                   c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ppm_pack_bytes": (c_int, [c_int, c_int, c_int, c_int, POINTER(c_size_t), POINTER(c_size_t),
                               POINTER(c_size_t)]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load the library (once).  Raises RuntimeError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m ppmstereo_b200.build` "
            "(nvcc, sm_100a).  ppmstereo_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype, fn.argtypes = res, args
    if lib.ppm_abi_version() != 1:
        raise RuntimeError("libppm_sm100.so ABI version mismatch")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().ppm_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def dtype_code(t) -> int:
    import torch
    if t.dtype == torch.float32:
        return PPM_F32
    if t.dtype == torch.bfloat16:
        return PPM_BF16
    raise RuntimeError(f"unsupported dtype {t.dtype}: expected float32 or bfloat16")


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


def ptr_array(tensors):
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr
