"""ctypes binding of ``libhive_b200.so`` (include/hive_b200.h).

There is no CPU fallback: if the CUDA library is missing or a GPU is absent the product path raises.
"""
from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhive_b200.so")

HV_OK = 0
HV_E_INVALID = -1
HV_E_CUDA = -2
HV_E_NOSURFACE = -3
HV_E_LEVEL_RANGE = -4
HV_E_CAPACITY = -5

_lib = None
_lock = threading.Lock()

c_void_p, c_int, c_int64, c_double, c_float = (ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double,
                                                ctypes.c_float)

# name -> argtypes; every function returns int.  Kept in step with include/hive_b200.h (tests/test_abi.py checks it).
SIGNATURES = {
    "hv_version": [],
    "hv_build_id": [ctypes.c_char_p, c_int],
    "hv_last_error": [ctypes.c_char_p, c_int],
    "hv_sm_count": [],
    "hv_pack_frame": [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p],
    "hv_upload_frame": [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p],
    "hv_integrate": [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                     c_void_p, c_int, c_int, c_double, c_double, c_int, c_void_p, c_void_p],
    "hv_mc_scratch_bytes": [c_void_p, c_int, c_int, c_void_p],
    "hv_mc_scratch_init": [c_void_p, c_int, c_int, c_void_p, c_int64, c_void_p],
    "hv_mc_work_bytes": [c_int64, c_int64, c_void_p],
    "hv_mc_count": [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_double, c_void_p, c_int64, c_void_p, c_int64,
                    c_int64, c_void_p, c_void_p],
    "hv_mc_emit": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_double, c_void_p,
                   c_int64, c_void_p, c_int64, c_int64, c_int64, c_int64, c_float, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "hv_mc_extract": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_double, c_void_p,
                      c_int64, c_void_p, c_int64, c_int64, c_int64, c_int64, c_float, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p,
                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "hv_mc_column_counts": [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_double, c_void_p, c_int64, c_void_p, c_int64,
                            c_int64, c_void_p, c_void_p, c_void_p],
    "hv_mc_shift_scratch_bytes": [c_int64, c_void_p],
    "hv_mc_column_shifts": [c_void_p, c_int64, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p],
    "hv_mc_emit_placed": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_double, c_void_p,
                          c_int64, c_void_p, c_int64, c_int64, c_int64, c_int64, c_float, c_void_p, c_int64, c_int64, c_void_p,
                          c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p],
    "hv_scatter_rows": [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p],
    "hv_mc_emit_direct": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_double, c_void_p,
                          c_int64, c_void_p, c_int64, c_int64, c_int64, c_int64, c_float, c_void_p, c_int64, c_int64, c_void_p,
                          c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "hv_peer_alloc": [c_int64, c_void_p, c_void_p],
    "hv_peer_free": [c_void_p],
    "hv_peer_open": [c_void_p, c_void_p],
    "hv_peer_close": [c_void_p],
    "hv_copy_async": [c_void_p, c_void_p, c_int64, c_void_p],
    "hv_peer_signal": [c_void_p, c_int, ctypes.c_uint, c_void_p],
    "hv_peer_wait": [c_void_p, c_int, ctypes.c_uint, c_void_p, c_void_p],
    "hv_copy_planes": [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p],
    "hv_mc_column_shifts_peers": [c_void_p, c_int, c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p],
    "hv_d2m_scratch_bytes": [c_int, c_int, c_int, c_void_p],
    "hv_d2m_count": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p, c_int64, c_void_p,
                     c_void_p],
    "hv_d2m_emit": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p, c_void_p, c_void_p,
                    c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "hv_d2m_emit_capped": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p, c_void_p, c_void_p,
                           c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    "hv_image2world": [c_void_p, c_void_p, c_int64, c_void_p, c_double, c_void_p, c_void_p],
    "hv_world2image": [c_void_p, c_int64, c_void_p, c_double, c_int, c_void_p, c_void_p, c_void_p],
    "hv_filter_faces_scratch_bytes": [c_int64, c_void_p],
    "hv_dilate_scratch_bytes": [c_int, c_int, c_int, c_void_p],
    "hv_dilate_mask": [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p],
    "hv_dilate_static_max_bytes": [c_void_p],
    "hv_dilate_mask_static": [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p],
    "hv_pack_frame_masked": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p],
    "hv_upload_frame_masked": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_int64, c_void_p, c_void_p],
    "hv_keyframe_overlap": [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p],
    "hv_depth_max": [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_void_p],
    "hv_byte_lut": [c_void_p, c_int64, c_void_p, c_void_p],
    "hv_atlas_pack": [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int64,
                      c_void_p, c_void_p],
    "hv_cc_scratch_bytes": [c_int64, c_void_p],
    "hv_cc_mask": [c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p],
    "hv_cc_cleanup_scratch_bytes": [c_int64, c_int64, c_void_p],
    "hv_cc_cleanup": [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p],
    "hv_count_nonzero": [c_void_p, c_int, c_int64, c_void_p, c_void_p],
    "hv_uv_bbox": [c_void_p, c_int64, c_void_p, c_void_p],
    "hv_cc_compact_scratch_bytes": [c_int64, c_void_p],
    "hv_cc_compact_vertices": [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                               c_int64, c_void_p],
    "hv_meshwrite_ply": [ctypes.c_char_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int],
    "hv_pcwrite_ply": [ctypes.c_char_p, c_void_p, c_void_p, c_int64, c_int],
    "hv_filter_faces": [c_void_p, c_void_p, c_int, c_void_p, c_int64, c_double, c_double, c_void_p, c_int64, c_int,
                        c_void_p, c_void_p, c_void_p],
}


class HiveB200Error(RuntimeError):
    pass


def load():
    """Load the shared library (building is ``python -m hive_b200.build`` / ``__graft_entry__.build()``)."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise HiveB200Error(
                    f"{LIB_PATH} is missing: build it with `python -m hive_b200.build` -- hive_b200 has no CPU fallback")
            lib = ctypes.CDLL(LIB_PATH)
            for name, argtypes in SIGNATURES.items():
                fn = getattr(lib, name)
                fn.argtypes = argtypes
                fn.restype = c_int
            _lib = lib
    return _lib


def build_id() -> str:
    """Hash of the sources the loaded library was compiled from (compare with ``build.source_hash()``)."""
    buf = ctypes.create_string_buffer(64)
    load().hv_build_id(buf, 64)
    return buf.value.decode()


def last_error() -> str:
    buf = ctypes.create_string_buffer(512)
    load().hv_last_error(buf, 512)
    return buf.value.decode("utf-8", "replace")


def check(code: int) -> int:
    """Map a C-ABI return code to the exception type the reference raises for the same condition."""
    if code >= 0:
        return code
    msg = last_error()
    if code == HV_E_INVALID:
        raise AssertionError(msg)
    if code == HV_E_LEVEL_RANGE:
        raise ValueError("Surface level must be within volume data range.")
    if code == HV_E_NOSURFACE:
        raise RuntimeError("No surface found at the given iso value.")
    raise HiveB200Error(f"hive_b200 error {code}: {msg}")


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise HiveB200Error("hive_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch
