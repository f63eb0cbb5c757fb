"""Python mirror of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:45-111)
over the C ABI of include/orbx.h.  Same constructor arguments, same getters;
`__call__` plays the role of operator()(image, mask, keypoints, descriptors)
and returns (keypoints, descriptors) as numpy arrays (structured cv::KeyPoint
layout, N x 32 uint8)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import KEYPOINT_DTYPE, check, lib


class ORBextractor:
    HARRIS_SCORE = 0
    FAST_SCORE = 1

    def __init__(self, nfeatures: int, scaleFactor: float, nlevels: int, iniThFAST: int, minThFAST: int,
                 device: int = 0):
        self._L = lib()
        h = C.c_void_p()
        check(self._L.orbx_create(C.byref(h), nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device),
              "orbx_create")
        self._h = h
        self.nlevels = nlevels
        n = nlevels
        sf, isf, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        nf = np.zeros(n, np.int32)
        check(self._L.orbx_scale_tables(h, sf.ctypes.data, isf.ctypes.data, s2.ctypes.data, is2.ctypes.data,
                                        nf.ctypes.data), "orbx_scale_tables")
        self._sf, self._isf, self._s2, self._is2, self.mnFeaturesPerLevel = sf, isf, s2, is2, nf
        self._scaleFactor = float(np.float32(scaleFactor))

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # getters of include/ORBextractor.h:63-83
    def GetLevels(self): return self.nlevels
    def GetScaleFactor(self): return self._scaleFactor
    def GetScaleFactors(self): return self._sf.copy()
    def GetInverseScaleFactors(self): return self._isf.copy()
    def GetScaleSigmaSquares(self): return self._s2.copy()
    def GetInverseScaleSigmaSquares(self): return self._is2.copy()

    def max_keypoints(self, width: int, height: int) -> int:
        n = self._L.orbx_max_keypoints(self._h, width, height)
        if n < 0:
            check(n, "orbx_max_keypoints")
        return n

    def __call__(self, image: np.ndarray, mask=None):
        """operator(): mask is ignored, as in the reference (include/ORBextractor.h:58)."""
        assert image.dtype == np.uint8 and image.ndim == 2     # assert(image.type() == CV_8UC1), :1256
        if image.size == 0:
            return None, None                                   # outputs untouched, :1252
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        cap = self.max_keypoints(w, h)
        kps = np.empty(cap, KEYPOINT_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n = C.c_int()
        check(self._L.orbx_extract(self._h, image.ctypes.data, w, h, image.strides[0], kps.ctypes.data,
                                   desc.ctypes.data, cap, C.byref(n)), "orbx_extract")
        return kps[: n.value].copy(), desc[: n.value].copy()

    def extract_batch(self, images: np.ndarray):
        """images: (n, h, w) uint8, C-contiguous.  Returns lists of per-frame (keypoints, descriptors)."""
        assert images.dtype == np.uint8 and images.ndim == 3 and images.flags.c_contiguous
        n, h, w = images.shape
        cap = self.max_keypoints(w, h)
        kps = np.empty((n, cap), KEYPOINT_DTYPE)
        desc = np.empty((n, cap, 32), np.uint8)
        cnt = np.zeros(n, np.int32)
        ptrs = (C.c_void_p * n)(*[images.ctypes.data + i * images.strides[0] for i in range(n)])
        check(self._L.orbx_extract_batch(self._h, ptrs, n, w, h, images.strides[1], kps.ctypes.data,
                                         desc.ctypes.data, cap, cnt.ctypes.data), "orbx_extract_batch")
        return [(kps[i, : cnt[i]].copy(), desc[i, : cnt[i]].copy()) for i in range(n)]

    # mvImagePyramid (include/ORBextractor.h:85)
    def level_size(self, level: int):
        w, h = C.c_int(), C.c_int()
        check(self._L.orbx_level_size(self._h, level, C.byref(w), C.byref(h)), "orbx_level_size")
        return w.value, h.value

    def get_level(self, level: int, frame: int = 0, border: int = 0, blurred: bool = False) -> np.ndarray:
        w, h = self.level_size(level)
        out = np.empty((h + 2 * border, w + 2 * border), np.uint8)
        check(self._L.orbx_get_level(self._h, frame, level, border, int(blurred), out.ctypes.data, out.strides[0]),
              "orbx_get_level")
        return out

    def get_candidates(self, level: int, frame: int = 0) -> np.ndarray:
        n = C.c_int()
        check(self._L.orbx_get_candidates(self._h, frame, level, None, 0, C.byref(n)), "orbx_get_candidates")
        out = np.zeros((max(n.value, 1), 3), np.int32)
        check(self._L.orbx_get_candidates(self._h, frame, level, out.ctypes.data, n.value, C.byref(n)),
              "orbx_get_candidates")
        return out[: n.value]

    def stereo_match(self, right: "ORBextractor", keysL, descL, keysR, descR, mbf: float, mb: float, frame: int = 0):
        """Frame::ComputeStereoMatches (src/Frame.cc:630-872) with self = mpORBextractorLeft, right = mpORBextractorRight,
        on the pyramids of their last extraction.  Returns (mvuRight, mvDepth, number of matches)."""
        kL = np.ascontiguousarray(keysL, KEYPOINT_DTYPE); kR = np.ascontiguousarray(keysR, KEYPOINT_DTYPE)
        dL = np.ascontiguousarray(descL, np.uint8); dR = np.ascontiguousarray(descR, np.uint8)
        u = np.empty(len(kL), np.float32); z = np.empty(len(kL), np.float32); n = C.c_int()
        check(self._L.orbx_stereo_match(self._h, right._h, frame, kL.ctypes.data, dL.ctypes.data, len(kL), kR.ctypes.data,
                                        dR.ctypes.data, len(kR), mbf, mb, u.ctypes.data, z.ctypes.data, C.byref(n)),
              "orbx_stereo_match")
        return u, z, n.value

    def stereo_match_batch_device(self, right: "ORBextractor", n, d_keysL, d_descL, d_nL, d_keysR, d_descR, d_nR, cap, mbf, mb,
                                  d_uRight, d_depth, d_nMatched):
        """Frame::ComputeStereoMatches for the n stereo pairs both extractors have just processed with
        orbx_extract_batch_device; every d_* argument is a device pointer (int).  Enqueued on this extractor's stream."""
        check(self._L.orbx_stereo_match_batch_device(self._h, right._h, n, d_keysL, d_descL, d_nL, d_keysR, d_descR, d_nR, cap,
                                                     mbf, mb, d_uRight, d_depth, d_nMatched), "orbx_stereo_match_batch_device")

    def synchronize(self):
        check(self._L.orbx_synchronize(self._h), "orbx_synchronize")

    @property
    def handle(self):
        return self._h
