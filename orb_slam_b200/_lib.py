"""ctypes binding of the C ABI in include/orbx.h / include/orbm.h.

The shared library is built in-tree (orb_slam_b200/libfrontend_b200.so) by
`__graft_entry__.build()` / `make -C orb_slam_b200/csrc`.  There is no CPU
fallback: if the library is missing or no sm_100 GPU is present, loading or
`orbx_create` fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ORB_B200_LIB") or os.path.join(_HERE, "libfrontend_b200.so")     # (override: instrumented debug builds)

KEYPOINT_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
     ("octave", "<i4"), ("class_id", "<i4")]
)
assert KEYPOINT_DTYPE.itemsize == 28

ORBX_OK = 0
ORBX_E_CAPACITY = -3

_lib = None


class FrontendError(RuntimeError):
    pass


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FrontendError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C orb_slam_b200/csrc` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    L.orbx_create.argtypes = [C.POINTER(vp), i, f, i, i, i, i]; L.orbx_create.restype = i
    L.orbx_destroy.argtypes = [vp]; L.orbx_destroy.restype = None
    L.orbx_levels.argtypes = [vp]; L.orbx_levels.restype = i
    L.orbx_scale_tables.argtypes = [vp, vp, vp, vp, vp, vp]; L.orbx_scale_tables.restype = i
    L.orbx_max_keypoints.argtypes = [vp, i, i]; L.orbx_max_keypoints.restype = i
    L.orbx_extract.argtypes = [vp, vp, i, i, sz, vp, vp, i, C.POINTER(i)]; L.orbx_extract.restype = i
    L.orbx_extract_batch.argtypes = [vp, C.POINTER(vp), i, i, i, sz, vp, vp, i, vp]; L.orbx_extract_batch.restype = i
    L.orbx_extract_batch_device.argtypes = [vp, vp, i, i, i, sz, sz, vp, vp, i, vp]
    L.orbx_extract_batch_device.restype = i
    L.orbx_stream.argtypes = [vp]; L.orbx_stream.restype = vp
    L.orbx_synchronize.argtypes = [vp]; L.orbx_synchronize.restype = i
    L.orbx_level_size.argtypes = [vp, i, C.POINTER(i), C.POINTER(i)]; L.orbx_level_size.restype = i
    L.orbx_get_level.argtypes = [vp, i, i, i, i, vp, sz]; L.orbx_get_level.restype = i
    L.orbx_pyramid_host.argtypes = [vp, i, C.POINTER(vp), vp, vp]; L.orbx_pyramid_host.restype = i
    L.orbx_get_candidates.argtypes = [vp, i, i, vp, i, C.POINTER(i)]; L.orbx_get_candidates.restype = i
    L.orbx_stereo_match.argtypes = [vp, vp, i, vp, vp, i, vp, vp, i, f, f, vp, vp, C.POINTER(i)]; L.orbx_stereo_match.restype = i
    L.orbx_stereo_match_batch_device.argtypes = [vp, vp, i, vp, vp, vp, vp, vp, vp, i, f, f, vp, vp, vp]
    L.orbx_stereo_match_batch_device.restype = i
    L.orbx_launches_per_call.argtypes = [vp]; L.orbx_launches_per_call.restype = i
    L.orbx_set_profiling.argtypes = [vp, i]; L.orbx_set_profiling.restype = i
    L.orbx_stage_times.argtypes = [vp, vp, i]; L.orbx_stage_times.restype = i
    L.orbx_error_string.argtypes = [i]; L.orbx_error_string.restype = C.c_char_p
    _lib = L
    return L


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().orbx_error_string(rc).decode()
        raise FrontendError(f"{what} failed: {msg} (code {rc})")
