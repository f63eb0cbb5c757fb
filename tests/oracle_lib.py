"""Loader for the CPU oracle (oracle/liborb_oracle.so) and, when present, the
reference build (oracle/_ref/liborb_ref.so).  Test infrastructure only."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liborb_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "liborb_ref.so")

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
               ("octave", "<i4"), ("class_id", "<i4")])

_oracle = None
_ref = None


def oracle() -> C.CDLL:
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
        L = C.CDLL(ORACLE_SO)
        vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
        L.orbo_create.restype = vp; L.orbo_create.argtypes = [i, f, i, i, i]
        L.orbo_destroy.argtypes = [vp]
        L.orbo_levels.argtypes = [vp]
        L.orbo_scale_tables.argtypes = [vp, vp, vp, vp, vp, vp]
        L.orbo_extract.argtypes = [vp, vp, i, i, sz, vp, vp, i]
        L.orbo_level_size.argtypes = [vp, i, vp, vp]
        L.orbo_get_level.argtypes = [vp, i, i, i, vp, sz]
        L.orbo_get_candidates.argtypes = [vp, i, vp, i]
        L.orbo_distribute_octtree.argtypes = [vp, i, i, i, i, i, i, vp, i]
        L.orbo_ic_angle.argtypes = [vp, sz, i, i]; L.orbo_ic_angle.restype = f
        L.orbo_descriptor.argtypes = [vp, sz, i, i, f, vp]
        _oracle = L
    return _oracle


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def ref() -> C.CDLL:
    global _ref
    if _ref is None:
        L = C.CDLL(REF_SO)
        vp, i, f = C.c_void_p, C.c_int, C.c_float
        L.orbref_create.restype = vp; L.orbref_create.argtypes = [i, f, i, i, i]
        L.orbref_destroy.argtypes = [vp]
        L.orbref_extract.argtypes = [vp, vp, i, i, i, i, vp, vp, i]
        L.orbref_level_size.argtypes = [vp, i, vp, vp]
        L.orbref_get_level.argtypes = [vp, i, i, vp, i]
        _ref = L
    return _ref


class OracleExtractor:
    """CPU oracle with the ORBextractor call shape."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniTh, minTh):
        self.L = oracle()
        self.h = self.L.orbo_create(nfeatures, scaleFactor, nlevels, iniTh, minTh)
        self.nlevels = nlevels
        self.nfeatures = nfeatures

    def __del__(self):
        try:
            self.L.orbo_destroy(self.h)
        except Exception:
            pass

    def tables(self):
        n = self.nlevels
        a = [np.zeros(n, np.float32) for _ in range(4)] + [np.zeros(n, np.int32)]
        self.L.orbo_scale_tables(self.h, *[x.ctypes.data for x in a])
        return a

    def __call__(self, image):
        image = np.ascontiguousarray(image)
        h, w = image.shape
        cap = self.nfeatures * 2 + 64 * self.nlevels
        kps = np.zeros(cap, KP); desc = np.zeros((cap, 32), np.uint8)
        n = self.L.orbo_extract(self.h, image.ctypes.data, w, h, image.strides[0], kps.ctypes.data, desc.ctypes.data, cap)
        if n < 0:
            raise RuntimeError(f"oracle extract returned {n}")
        assert n <= cap
        return kps[:n].copy(), desc[:n].copy()

    def level_size(self, l):
        w, h = C.c_int(), C.c_int()
        self.L.orbo_level_size(self.h, l, C.byref(w), C.byref(h))
        return w.value, h.value

    def get_level(self, l, border=0, blurred=False):
        w, h = self.level_size(l)
        out = np.zeros((h + 2 * border, w + 2 * border), np.uint8)
        rc = self.L.orbo_get_level(self.h, l, border, int(blurred), out.ctypes.data, out.strides[0])
        return out if rc == 0 else None

    def get_candidates(self, l):
        n = self.L.orbo_get_candidates(self.h, l, None, 0)
        out = np.zeros((max(n, 1), 3), np.int32)
        self.L.orbo_get_candidates(self.h, l, out.ctypes.data, n)
        return out[:n]


class RefExtractor:
    """The reference's own ORBextractor (oracle/_ref), canonical tie-break by default."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniTh, minTh):
        self.L = ref()
        self.h = self.L.orbref_create(nfeatures, scaleFactor, nlevels, iniTh, minTh)
        self.nfeatures = nfeatures; self.nlevels = nlevels

    def __call__(self, image, canonical=True):
        image = np.ascontiguousarray(image)
        h, w = image.shape
        cap = self.nfeatures * 2 + 64 * self.nlevels
        kps = np.zeros(cap, KP); desc = np.zeros((cap, 32), np.uint8)
        n = self.L.orbref_extract(self.h, image.ctypes.data, w, h, image.strides[0], int(canonical), kps.ctypes.data,
                                  desc.ctypes.data, cap)
        if n < 0:
            raise RuntimeError(f"reference extract returned {n}")
        return kps[:n].copy(), desc[:n].copy()

    def get_level(self, l, border=0):
        w, h = C.c_int(), C.c_int()
        self.L.orbref_level_size(self.h, l, C.byref(w), C.byref(h))
        out = np.zeros((h.value + 2 * border, w.value + 2 * border), np.uint8)
        self.L.orbref_get_level(self.h, l, border, out.ctypes.data, out.strides[0])
        return out
