"""Python mirror of ORBVocabulary::transform (Frame::ComputeBoW, reference src/Frame.cc:521-528 ->
Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1194) and of the batched
MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:241-306) over the C ABI of include/orbm.h."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import check, lib
from .matcher import FeatureVector, MatcherContext

TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3
L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = 0, 1, 2, 3, 4, 5


def _bind(L):
    if getattr(L, "_vocab_bound", False):
        return L
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    L.orbm_vocab_create.argtypes = [vp, i, i, i, i, i, vp, vp, vp, vp, C.POINTER(vp)]; L.orbm_vocab_create.restype = i
    L.orbm_vocab_load_text.argtypes = [vp, C.c_char_p, C.POINTER(vp)]; L.orbm_vocab_load_text.restype = i
    L.orbm_vocab_destroy.argtypes = [vp]; L.orbm_vocab_destroy.restype = None
    L.orbm_vocab_info.argtypes = [vp] + [C.POINTER(i)] * 6; L.orbm_vocab_info.restype = i
    L.orbm_vocab_descend.argtypes = [vp, vp, vp, i, i, vp, vp, vp]; L.orbm_vocab_descend.restype = i
    L.orbm_vocab_descend_device.argtypes = [vp, vp, vp, sz, i, vp]; L.orbm_vocab_descend_device.restype = i
    L.orbm_vocab_transform.argtypes = [vp, vp, vp, i, i, vp, vp, C.POINTER(i), vp, vp, vp, C.POINTER(i)]
    L.orbm_vocab_transform.restype = i
    L.orbm_distinctive_descriptors.argtypes = [vp, vp, vp, i, vp]; L.orbm_distinctive_descriptors.restype = i
    L._vocab_bound = True
    return L


class ORBVocabulary:
    """A vocabulary tree resident on the GPU.  Build with `from_arrays` (the node table loadFromTextFile would
    produce) or `loadFromTextFile` (the ORBvoc.txt format)."""

    def __init__(self, ctx: MatcherContext, handle):
        self._ctx, self._h = ctx, handle
        L = _bind(lib())
        v = [C.c_int() for _ in range(6)]
        check(L.orbm_vocab_info(self._h, *[C.byref(x) for x in v]), "orbm_vocab_info")
        self.k, self.L, self.nnodes, self.nwords, self.scoring, self.weighting = (x.value for x in v)

    @classmethod
    def from_arrays(cls, ctx, k, L, parent, is_leaf, desc, weight, scoring=L1_NORM, weighting=TF_IDF):
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); weight = np.ascontiguousarray(weight, np.float64)
        n = len(parent)
        assert len(is_leaf) == n and len(desc) == n and len(weight) == n
        h = C.c_void_p()
        check(_bind(lib()).orbm_vocab_create(ctx.h, k, L, scoring, weighting, n, parent.ctypes.data, is_leaf.ctypes.data,
                                             desc.ctypes.data, weight.ctypes.data, C.byref(h)), "orbm_vocab_create")
        return cls(ctx, h)

    @classmethod
    def loadFromTextFile(cls, ctx, path):
        h = C.c_void_p()
        check(_bind(lib()).orbm_vocab_load_text(ctx.h, str(path).encode(), C.byref(h)), "orbm_vocab_load_text")
        return cls(ctx, h)

    def close(self):
        if self._h:
            lib().orbm_vocab_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def size(self):
        return self.nwords

    def descend(self, descriptors, levelsup=4):
        """transform(feature, word_id, weight, &nid, levelsup) per row -> (word ids, node ids, weights)."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        n = len(d)
        w = np.zeros(n, np.uint32); nid = np.zeros(n, np.uint32); wt = np.zeros(n, np.float64)
        check(lib().orbm_vocab_descend(self._ctx.h, self._h, d.ctypes.data, n, levelsup, w.ctypes.data, nid.ctypes.data,
                                       wt.ctypes.data), "orbm_vocab_descend")
        return w, nid, wt

    def transform(self, descriptors, levelsup=4):
        """transform(features, BowVector, FeatureVector, levelsup) -> (word ids, values, FeatureVector)."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        n = len(d)
        wid = np.zeros(max(n, 1), np.uint32); wv = np.zeros(max(n, 1), np.float64)
        nid = np.zeros(max(n, 1), np.uint32); st = np.zeros(n + 1, np.int32); it = np.zeros(max(n, 1), np.uint32)
        nw, nn = C.c_int(), C.c_int()
        check(lib().orbm_vocab_transform(self._ctx.h, self._h, d.ctypes.data, n, levelsup, wid.ctypes.data, wv.ctypes.data,
                                         C.byref(nw), nid.ctypes.data, st.ctypes.data, it.ctypes.data, C.byref(nn)),
              "orbm_vocab_transform")
        fv = FeatureVector.from_flat(nid[:nn.value], st[:nn.value + 1], it[:st[nn.value]])
        return wid[:nw.value].copy(), wv[:nw.value].copy(), fv


def compute_distinctive_descriptors(ctx: MatcherContext, descriptors, start):
    """Batched MapPoint::ComputeDistinctiveDescriptors: best[p] = index inside point p's descriptor list."""
    d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
    start = np.ascontiguousarray(start, np.int32)
    npts = len(start) - 1
    best = np.zeros(max(npts, 1), np.int32)
    check(_bind(lib()).orbm_distinctive_descriptors(ctx.h, d.ctypes.data, start.ctypes.data, npts, best.ctypes.data),
          "orbm_distinctive_descriptors")
    return best[:npts]
