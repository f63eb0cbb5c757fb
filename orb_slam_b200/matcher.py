"""Python mirror of the ORBmatcher Hamming hot path and the Frame grid over the C ABI of include/orbm.h
(reference include/ORBmatcher.h:37-102, include/Frame.h:87-89,161)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import KEYPOINT_DTYPE, check, lib

TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30
FRAME_GRID_ROWS, FRAME_GRID_COLS = 48, 64

TOP2_DTYPE = np.dtype([("i1", "<u4"), ("d1", "<u2"), ("d2", "<u2")])
MAPPOINT_DTYPE = np.dtype([
    ("world", "<f4", 3), ("observations", "<i4"), ("descriptor", "u1", 32), ("valid", "u1"), ("bad", "u1"),
    ("outlier", "u1"), ("trackInView", "u1"), ("trackProjX", "<f4"), ("trackProjY", "<f4"), ("trackProjXR", "<f4"),
    ("trackViewCos", "<f4"), ("trackScaleLevel", "<i4")])
assert MAPPOINT_DTYPE.itemsize == 72


class FrameStruct(C.Structure):
    _fields_ = [("N", C.c_int32), ("keys", C.c_void_p), ("keysUn", C.c_void_p), ("descriptors", C.c_void_p),
                ("uRight", C.c_void_p), ("mnMinX", C.c_float), ("mnMaxX", C.c_float), ("mnMinY", C.c_float),
                ("mnMaxY", C.c_float), ("gridElementWidthInv", C.c_float), ("gridElementHeightInv", C.c_float),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("bf", C.c_float),
                ("b", C.c_float), ("Tcw", C.c_float * 16), ("scaleFactors", C.c_void_p), ("nScaleLevels", C.c_int32)]


class Frame:
    """The members of ORB_SLAM2::Frame the matcher reads (include/Frame.h), as numpy arrays."""

    def __init__(self, keys, descriptors, width, height, scale_factors, *, keys_un=None, u_right=None, K=None,
                 bf=0.0, Tcw=None):
        self.mvKeys = np.ascontiguousarray(keys, KEYPOINT_DTYPE)
        self.mvKeysUn = self.mvKeys if keys_un is None else np.ascontiguousarray(keys_un, KEYPOINT_DTYPE)
        self.mDescriptors = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        self.N = len(self.mvKeys)
        self.mvuRight = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        # ComputeImageBounds without distortion (src/Frame.cc:606-612) and the grid scale (:126-127)
        self.mnMinX, self.mnMaxX, self.mnMinY, self.mnMaxY = np.float32(0), np.float32(width), np.float32(0), np.float32(height)
        self.mfGridElementWidthInv = np.float32(FRAME_GRID_COLS) / (self.mnMaxX - self.mnMinX)
        self.mfGridElementHeightInv = np.float32(FRAME_GRID_ROWS) / (self.mnMaxY - self.mnMinY)
        K = np.eye(3, dtype=np.float32) if K is None else np.asarray(K, np.float32)
        self.fx, self.fy, self.cx, self.cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
        self.mbf = np.float32(bf)
        self.mb = np.float32(self.mbf / self.fx)
        self.mTcw = np.eye(4, dtype=np.float32) if Tcw is None else np.ascontiguousarray(Tcw, np.float32)
        self.mvScaleFactors = np.ascontiguousarray(scale_factors, np.float32)

    def struct(self) -> FrameStruct:
        s = FrameStruct()
        s.N = self.N
        s.keys = self.mvKeys.ctypes.data; s.keysUn = self.mvKeysUn.ctypes.data
        s.descriptors = self.mDescriptors.ctypes.data
        s.uRight = None if self.mvuRight is None else self.mvuRight.ctypes.data
        s.mnMinX, s.mnMaxX, s.mnMinY, s.mnMaxY = float(self.mnMinX), float(self.mnMaxX), float(self.mnMinY), float(self.mnMaxY)
        s.gridElementWidthInv = float(self.mfGridElementWidthInv); s.gridElementHeightInv = float(self.mfGridElementHeightInv)
        s.fx, s.fy, s.cx, s.cy, s.bf, s.b = (float(v) for v in (self.fx, self.fy, self.cx, self.cy, self.mbf, self.mb))
        s.Tcw = (C.c_float * 16)(*self.mTcw.ravel().tolist())
        s.scaleFactors = self.mvScaleFactors.ctypes.data; s.nScaleLevels = len(self.mvScaleFactors)
        return s


class FeatVecStruct(C.Structure):
    """orbm_featvec of include/orbm.h."""
    _fields_ = [("nnodes", C.c_int32), ("node_ids", C.c_void_p), ("start", C.c_void_p), ("items", C.c_void_p)]


class FeatureVector:
    """DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) flattened in map order for the C ABI.
    Built from a dict {node_id: [feature indices]} or an iterable of (node_id, indices)."""

    def __init__(self, nodes):
        pairs = sorted(nodes.items() if hasattr(nodes, "items") else nodes, key=lambda kv: kv[0])
        self.node_ids = np.array([k for k, _ in pairs], np.uint32)
        lens = [len(v) for _, v in pairs]
        self.start = np.zeros(len(pairs) + 1, np.int32)
        if lens:
            self.start[1:] = np.cumsum(lens)
        self.items = (np.concatenate([np.asarray(v, np.uint32) for _, v in pairs]) if lens and sum(lens)
                      else np.zeros(0, np.uint32))
        self.items = np.ascontiguousarray(self.items, np.uint32)

    @classmethod
    def from_flat(cls, node_ids, start, items):
        fv = cls({})
        fv.node_ids = np.ascontiguousarray(node_ids, np.uint32).copy()
        fv.start = np.ascontiguousarray(start, np.int32).copy()
        fv.items = np.ascontiguousarray(items, np.uint32).copy()
        return fv

    def as_dict(self):
        return {int(k): self.items[self.start[j]:self.start[j + 1]].tolist() for j, k in enumerate(self.node_ids)}

    def struct(self) -> FeatVecStruct:
        s = FeatVecStruct()
        s.nnodes = len(self.node_ids)
        s.node_ids = self.node_ids.ctypes.data; s.start = self.start.ctypes.data; s.items = self.items.ctypes.data
        return s


class KeyFrameStruct(C.Structure):
    """orbm_keyframe of include/orbm.h."""
    _fields_ = [("N", C.c_int32), ("keysUn", C.c_void_p), ("descriptors", C.c_void_p), ("uRight", C.c_void_p),
                ("has_point", C.c_void_p), ("featvec", C.c_void_p)]


class Sim3SideStruct(C.Structure):
    """orbm_sim3_side of include/orbm.h."""
    _fields_ = [("kf", C.c_void_p), ("points", C.c_void_p), ("skip", C.c_void_p), ("min_distance_inv", C.c_void_p),
                ("max_distance_inv", C.c_void_p), ("max_distance", C.c_void_p), ("log_scale_factor", C.c_float),
                ("predict", C.c_void_p), ("user", C.c_void_p)]


class KeyFrameView:
    """The members of ORB_SLAM2::KeyFrame that SearchForTriangulation reads."""

    def __init__(self, keys_un, descriptors, u_right, has_point, featvec: FeatureVector):
        self.keysUn = np.ascontiguousarray(keys_un, KEYPOINT_DTYPE)
        self.descriptors = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        self.uRight = np.ascontiguousarray(u_right, np.float32)
        self.has_point = np.ascontiguousarray(has_point, np.uint8)
        self.featvec = featvec
        self.N = len(self.keysUn)

    def struct(self) -> KeyFrameStruct:
        self._fv = self.featvec.struct()
        s = KeyFrameStruct()
        s.N = self.N; s.keysUn = self.keysUn.ctypes.data; s.descriptors = self.descriptors.ctypes.data
        s.uRight = self.uRight.ctypes.data; s.has_point = self.has_point.ctypes.data
        s.featvec = C.cast(C.pointer(self._fv), C.c_void_p)
        return s


def _bind(L):
    if getattr(L, "_orbm_bound", False):
        return
    vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    L.orbm_create.argtypes = [C.POINTER(vp), i]; L.orbm_create.restype = i
    L.orbm_destroy.argtypes = [vp]; L.orbm_destroy.restype = None
    L.orbm_stream.argtypes = [vp]; L.orbm_stream.restype = vp
    L.orbm_synchronize.argtypes = [vp]; L.orbm_synchronize.restype = i
    L.orbm_descriptor_distance.argtypes = [vp, vp]; L.orbm_descriptor_distance.restype = i
    L.orbm_three_maxima.argtypes = [vp, i, C.POINTER(i), C.POINTER(i), C.POINTER(i)]; L.orbm_three_maxima.restype = None
    L.orbm_bf2nn_device.argtypes = [vp, vp, i, vp, sz, C.c_uint32, vp]; L.orbm_bf2nn_device.restype = i
    L.orbm_bf2nn_device_noalloc.argtypes = [vp, vp, i, vp, sz, C.c_uint32, vp]; L.orbm_bf2nn_device_noalloc.restype = i
    L.orbm_merge_top2_device.argtypes = [vp, vp, i, i, vp]; L.orbm_merge_top2_device.restype = i
    L.orbm_bf2nn.argtypes = [vp, vp, i, vp, sz, C.c_uint32, vp]; L.orbm_bf2nn.restype = i
    L.orbm_popc_peak.argtypes = [vp, C.POINTER(C.c_double)]; L.orbm_popc_peak.restype = i
    L.orbm_lop3_peak.argtypes = [vp, C.POINTER(C.c_double)]; L.orbm_lop3_peak.restype = i
    L.orbm_pair_mix_peak.argtypes = [vp, C.POINTER(C.c_double)]; L.orbm_pair_mix_peak.restype = i
    L.orbm_nccl_unique_id.argtypes = [vp]; L.orbm_nccl_unique_id.restype = i
    L.orbm_nccl_version.argtypes = [C.POINTER(i)]; L.orbm_nccl_version.restype = i
    L.orbm_shard_create.argtypes = [C.POINTER(vp), vp, vp, i, i]; L.orbm_shard_create.restype = i
    L.orbm_shard_create_with_comm.argtypes = [C.POINTER(vp), vp, vp, i, i]; L.orbm_shard_create_with_comm.restype = i
    L.orbm_shard_destroy.argtypes = [vp]; L.orbm_shard_destroy.restype = None
    L.orbm_shard_set_rows.argtypes = [vp, vp, sz, C.c_uint32]; L.orbm_shard_set_rows.restype = i
    L.orbm_shard_upload_rows.argtypes = [vp, vp, sz, C.c_uint32]; L.orbm_shard_upload_rows.restype = i
    L.orbm_shard_bf2nn_device.argtypes = [vp, vp, i, vp]; L.orbm_shard_bf2nn_device.restype = i
    L.orbm_shard_bf2nn.argtypes = [vp, vp, i, vp]; L.orbm_shard_bf2nn.restype = i
    L.orbm_grid_build.argtypes = [vp, vp, vp, vp]; L.orbm_grid_build.restype = i
    L.orbm_features_in_area.argtypes = [vp, vp, f, f, f, i, i, vp, i, C.POINTER(i)]; L.orbm_features_in_area.restype = i
    L.orbm_search_for_initialization.argtypes = [vp, vp, vp, vp, vp, i, f, i, C.POINTER(i)]
    L.orbm_search_for_initialization.restype = i
    L.orbm_search_by_projection_frame.argtypes = [vp, vp, vp, vp, vp, f, i, i, vp, C.POINTER(i)]
    L.orbm_search_by_projection_frame.restype = i
    L.orbm_search_by_projection_points.argtypes = [vp, vp, vp, i, vp, f, f, vp, C.POINTER(i)]
    L.orbm_search_by_projection_points.restype = i
    L.orbm_search_by_projection_keyframe.argtypes = [vp, vp, vp, i, vp, vp, vp, f, vp, vp, vp, vp, vp, f, i, i, vp, C.POINTER(i)]
    L.orbm_search_by_projection_keyframe.restype = i
    L.orbm_fuse_select.argtypes = [vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, f, vp, vp, f, vp]
    L.orbm_fuse_select.restype = i
    L.orbm_search_by_projection_sim3.argtypes = [vp, vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, f, vp, vp, vp, i, vp, C.POINTER(i)]
    L.orbm_search_by_projection_sim3.restype = i
    L.orbm_fuse_select_sim3.argtypes = [vp, vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, f, vp, vp, f, vp]
    L.orbm_fuse_select_sim3.restype = i
    L.orbm_search_by_sim3.argtypes = [vp, vp, vp, vp, vp, vp, vp, f, vp, C.POINTER(i)]
    L.orbm_search_by_sim3.restype = i
    L.orbm_search_for_triangulation.argtypes = [vp, vp, vp, vp, vp, vp, f, f, f, f, vp, vp, i, vp, i, i, vp, C.POINTER(i)]
    L.orbm_search_for_triangulation.restype = i
    L.orbm_search_by_bow.argtypes = [vp, i, vp, vp, vp, i, vp, vp, vp, vp, i, vp, f, i, vp, C.POINTER(i)]
    L.orbm_search_by_bow.restype = i
    L._orbm_bound = True


class MatcherContext:
    """Device context shared by ORBmatcher objects of one host thread (orbm_create)."""

    def __init__(self, device: int = 0):
        self._L = lib(); _bind(self._L)
        h = C.c_void_p()
        check(self._L.orbm_create(C.byref(h), device), "orbm_create")
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self._L.orbm_destroy(self.h); self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def search_for_initialization_batch_device(self, d_keys, d_desc, d_n, cap, nframes, width, height, *, d_pairs=None, npairs=None,
                                               d_prev=None, windowSize=100, nnratio=0.9, checkOri=True, d_matches12=None,
                                               d_nmatches=None):
        """orbm_search_for_initialization_batch_device: every argument named d_* is a device pointer (int).  The grid is
        that of an undistorted width x height image (src/Frame.cc:606-612, :126-127)."""
        L = self._L
        if not hasattr(L, "_init_batch_bound"):
            vp, i, f = C.c_void_p, C.c_int, C.c_float
            L.orbm_search_for_initialization_batch_device.argtypes = [vp, vp, vp, vp, i, i, vp, i, f, f, f, f, vp, i, f, i, vp, vp]
            L.orbm_search_for_initialization_batch_device.restype = i
            L._init_batch_bound = True
        if npairs is None:
            npairs = nframes - 1
        invW = float(np.float32(FRAME_GRID_COLS) / (np.float32(width) - np.float32(0)))
        invH = float(np.float32(FRAME_GRID_ROWS) / (np.float32(height) - np.float32(0)))
        check(L.orbm_search_for_initialization_batch_device(self.h, d_keys, d_desc, d_n, cap, nframes, d_pairs, npairs, 0.0, 0.0, invW, invH,
                                                            d_prev, int(windowSize), float(np.float32(nnratio)), int(checkOri), d_matches12,
                                                            d_nmatches), "orbm_search_for_initialization_batch_device")

    def synchronize(self):
        check(self._L.orbm_synchronize(self.h), "orbm_synchronize")

    def popc_peak(self) -> float:
        v = C.c_double()
        check(self._L.orbm_popc_peak(self.h, C.byref(v)), "orbm_popc_peak")
        return v.value

    def pair_mix_peak(self) -> float:
        v = C.c_double()
        check(self._L.orbm_pair_mix_peak(self.h, C.byref(v)), "orbm_pair_mix_peak")
        return v.value

    def lop3_peak(self) -> float:
        v = C.c_double()
        check(self._L.orbm_lop3_peak(self.h, C.byref(v)), "orbm_lop3_peak")
        return v.value

    def grid_build(self, F: Frame):
        cs = np.zeros(FRAME_GRID_COLS * FRAME_GRID_ROWS + 1, np.int32); items = np.zeros(max(F.N, 1), np.int32)
        s = F.struct()
        check(self._L.orbm_grid_build(self.h, C.byref(s), cs.ctypes.data, items.ctypes.data), "orbm_grid_build")
        return cs, items[: cs[-1]]

    def features_in_area(self, F: Frame, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.zeros(max(F.N, 1), np.int32); n = C.c_int()
        s = F.struct()
        check(self._L.orbm_features_in_area(self.h, C.byref(s), x, y, r, minLevel, maxLevel, out.ctypes.data, len(out),
                                            C.byref(n)), "orbm_features_in_area")
        return out[: n.value].copy()

    def bf2nn(self, queries: np.ndarray, d_db_ptr: int, ndb: int, idx_base: int = 0) -> np.ndarray:
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
        out = np.zeros(len(q), TOP2_DTYPE)
        check(self._L.orbm_bf2nn(self.h, q.ctypes.data, len(q), d_db_ptr, ndb, idx_base, out.ctypes.data), "orbm_bf2nn")
        return out


class ORBmatcher:
    """ORBmatcher(nnratio=0.6, checkOri=True) — construction is free, as in the reference."""
    TH_HIGH, TH_LOW, HISTO_LENGTH = TH_HIGH, TH_LOW, HISTO_LENGTH

    def __init__(self, nnratio: float = 0.6, checkOri: bool = True, context: MatcherContext | None = None):
        self.mfNNratio = float(np.float32(nnratio)); self.mbCheckOrientation = bool(checkOri)
        self.ctx = context

    def _c(self):
        if self.ctx is None:
            self.ctx = MatcherContext()
        return self.ctx

    @staticmethod
    def DescriptorDistance(a: np.ndarray, b: np.ndarray) -> int:
        L = lib(); _bind(L)
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return L.orbm_descriptor_distance(a.ctypes.data, b.ctypes.data)

    @staticmethod
    def ComputeThreeMaxima(counts):
        L = lib(); _bind(L)
        c = np.ascontiguousarray(counts, np.int32)
        i1, i2, i3 = C.c_int(-1), C.c_int(-1), C.c_int(-1)
        L.orbm_three_maxima(c.ctypes.data, len(c), C.byref(i1), C.byref(i2), C.byref(i3))
        return i1.value, i2.value, i3.value

    def SearchForInitialization(self, F1: Frame, F2: Frame, vbPrevMatched: np.ndarray, windowSize: int = 10):
        """returns (nmatches, vnMatches12); vbPrevMatched (N1 x 2 float32) is updated in place"""
        c = self._c()
        assert vbPrevMatched.dtype == np.float32 and vbPrevMatched.flags.c_contiguous
        m12 = np.zeros(max(F1.N, 1), np.int32); n = C.c_int()
        s1, s2 = F1.struct(), F2.struct()
        check(c._L.orbm_search_for_initialization(c.h, C.byref(s1), C.byref(s2), vbPrevMatched.ctypes.data, m12.ctypes.data,
                                                  windowSize, self.mfNNratio, int(self.mbCheckOrientation), C.byref(n)),
              "orbm_search_for_initialization")
        return n.value, m12[: F1.N]

    def SearchByProjectionFrame(self, CurrentFrame: Frame, LastFrame: Frame, last_points: np.ndarray,
                                cur_observations: np.ndarray, th: float, bMono: bool):
        c = self._c()
        lp = np.ascontiguousarray(last_points, MAPPOINT_DTYPE); co = np.ascontiguousarray(cur_observations, np.int32)
        out = np.zeros(max(CurrentFrame.N, 1), np.int32); n = C.c_int()
        sc, sl = CurrentFrame.struct(), LastFrame.struct()
        check(c._L.orbm_search_by_projection_frame(c.h, C.byref(sc), C.byref(sl), lp.ctypes.data, co.ctypes.data, th,
                                                   int(bMono), int(self.mbCheckOrientation), out.ctypes.data, C.byref(n)),
              "orbm_search_by_projection_frame")
        return n.value, out[: CurrentFrame.N]

    def SearchByProjectionPoints(self, F: Frame, points: np.ndarray, frame_observations: np.ndarray, th: float = 3.0):
        c = self._c()
        p = np.ascontiguousarray(points, MAPPOINT_DTYPE); fo = np.ascontiguousarray(frame_observations, np.int32)
        out = np.zeros(max(F.N, 1), np.int32); n = C.c_int()
        s = F.struct()
        check(c._L.orbm_search_by_projection_points(c.h, C.byref(s), p.ctypes.data, len(p), fo.ctypes.data, th,
                                                    self.mfNNratio, out.ctypes.data, C.byref(n)),
              "orbm_search_by_projection_points")
        return n.value, out[: F.N]

    def SearchByProjectionKeyFrame(self, CurrentFrame: Frame, kf_points: np.ndarray, min_max_distance: np.ndarray,
                                   already_found: np.ndarray, kf_angle: np.ndarray, cur_has_point: np.ndarray,
                                   th: float, ORBdist: int, log_scale_factor: float | None = None):
        """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1540-1667).
        min_max_distance[i] = (mfMinDistance, mfMaxDistance) of key-frame map point i."""
        c = self._c()
        p = np.ascontiguousarray(kf_points, MAPPOINT_DTYPE)
        mm = np.ascontiguousarray(min_max_distance, np.float32).reshape(-1, 2)
        mn_inv = (np.float32(0.8) * mm[:, 0]).astype(np.float32); mx_inv = (np.float32(1.2) * mm[:, 1]).astype(np.float32)
        mx = np.ascontiguousarray(mm[:, 1])
        af = np.ascontiguousarray(already_found, np.uint8); ang = np.ascontiguousarray(kf_angle, np.float32)
        has = np.ascontiguousarray(cur_has_point, np.uint8)
        if log_scale_factor is None:
            log_scale_factor = float(np.log(np.float32(CurrentFrame.mvScaleFactors[1])))    # Frame::mfLogScaleFactor = log(mfScaleFactor)
        out = np.zeros(max(CurrentFrame.N, 1), np.int32); n = C.c_int()
        sc = CurrentFrame.struct()
        check(c._L.orbm_search_by_projection_keyframe(c.h, C.byref(sc), p.ctypes.data, len(p), mn_inv.ctypes.data, mx_inv.ctypes.data,
                                                      mx.ctypes.data, log_scale_factor, None, None, af.ctypes.data, ang.ctypes.data,
                                                      has.ctypes.data, th, int(ORBdist), int(self.mbCheckOrientation),
                                                      out.ctypes.data, C.byref(n)), "orbm_search_by_projection_keyframe")
        return n.value, out[: CurrentFrame.N]

    def FuseSelect(self, KF: Frame, inv_level_sigma2, points: np.ndarray, in_keyframe, normals, min_max_distance, th: float = 3.0,
                   log_scale_factor: float | None = None):
        """Selection part of Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:893-1040): best_idx[i] = key-frame keypoint map
        point i fuses with, or -1.  KF is the key frame seen through the Frame view (pose in mTcw, mvuRight set)."""
        c = self._c()
        p = np.ascontiguousarray(points, MAPPOINT_DTYPE)
        mm = np.ascontiguousarray(min_max_distance, np.float32).reshape(-1, 2)
        mn_inv = (np.float32(0.8) * mm[:, 0]).astype(np.float32); mx_inv = (np.float32(1.2) * mm[:, 1]).astype(np.float32)
        mx = np.ascontiguousarray(mm[:, 1])
        ik = np.ascontiguousarray(in_keyframe, np.uint8); nr = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        sg = np.ascontiguousarray(inv_level_sigma2, np.float32)
        if log_scale_factor is None:
            log_scale_factor = float(np.log(np.float32(KF.mvScaleFactors[1])))
        out = np.zeros(max(len(p), 1), np.int32)
        sk = KF.struct()
        check(c._L.orbm_fuse_select(c.h, C.byref(sk), None, sg.ctypes.data, p.ctypes.data, len(p), ik.ctypes.data, nr.ctypes.data,
                                    mn_inv.ctypes.data, mx_inv.ctypes.data, mx.ctypes.data, log_scale_factor, None, None, th,
                                    out.ctypes.data), "orbm_fuse_select")
        return out[: len(p)]

    def _sim3_args(self, KF, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, log_scale_factor):
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        p = np.ascontiguousarray(points, MAPPOINT_DTYPE)
        mm = f32(min_max_distance).reshape(-1, 2)
        keep = dict(R=f32(Rcw).ravel(), t=f32(tcw).ravel(), O=f32(Ow).ravel(), p=p, ex=np.ascontiguousarray(excluded, np.uint8),
                    nr=f32(normals).reshape(-1, 3), mn=(np.float32(0.8) * mm[:, 0]).astype(np.float32),
                    mx_inv=(np.float32(1.2) * mm[:, 1]).astype(np.float32), mx=np.ascontiguousarray(mm[:, 1]), sk=KF.struct())
        if log_scale_factor is None:
            log_scale_factor = float(np.log(np.float32(KF.mvScaleFactors[1])))
        k = keep
        args = [C.byref(k["sk"]), k["R"].ctypes.data, k["t"].ctypes.data, k["O"].ctypes.data, p.ctypes.data, len(p), k["ex"].ctypes.data,
                k["nr"].ctypes.data, k["mn"].ctypes.data, k["mx_inv"].ctypes.data, k["mx"].ctypes.data, log_scale_factor, None, None]
        return keep, args

    def SearchByProjectionSim3(self, KF: Frame, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, matched_in, th: int,
                               log_scale_factor: float | None = None):
        """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:290-403) with Scw decomposed by the caller.
        Returns (nmatches, match) with match[idx] = iMP newly stored in vpMatched[idx], -2 untouched."""
        c = self._c()
        keep, args = self._sim3_args(KF, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, log_scale_factor)
        mi = np.ascontiguousarray(matched_in, np.uint8)
        out = np.zeros(max(KF.N, 1), np.int32); n = C.c_int()
        check(c._L.orbm_search_by_projection_sim3(c.h, *args, mi.ctypes.data, int(th), out.ctypes.data, C.byref(n)),
              "orbm_search_by_projection_sim3")
        return n.value, out[: KF.N]

    def FuseSelectSim3(self, KF: Frame, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, th: float,
                       log_scale_factor: float | None = None):
        """Selection part of Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1044-1168)."""
        c = self._c()
        keep, args = self._sim3_args(KF, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, log_scale_factor)
        out = np.zeros(max(len(keep["p"]), 1), np.int32)
        check(c._L.orbm_fuse_select_sim3(c.h, *args, th, out.ctypes.data), "orbm_fuse_select_sim3")
        return out[: len(keep["p"])]

    def SearchBySim3(self, KF1: Frame, KF2: Frame, points1, points2, skip1, skip2, min_max_distance1, min_max_distance2,
                     sR12, t12, sR21, t21, th: float):
        """SearchBySim3 (src/ORBmatcher.cc:1170-1394).  points1/points2: the key frames' map points by keypoint (valid = 0:
        none), skip = already matched or bad.  Returns (nFound, match12) with match12[i1] = keypoint of KF2 or -1."""
        c = self._c()
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        keep = []

        def side(KF, other, pts, skip, mm):
            p = np.ascontiguousarray(pts, MAPPOINT_DTYPE); sk = np.ascontiguousarray(skip, np.uint8)
            mm = f32(mm).reshape(-1, 2)
            mn = (np.float32(0.8) * mm[:, 0]).astype(np.float32); mx_inv = (np.float32(1.2) * mm[:, 1]).astype(np.float32)
            mx = np.ascontiguousarray(mm[:, 1]); st = KF.struct()
            assert len(p) == KF.N and len(sk) == KF.N and len(mm) == KF.N
            s = Sim3SideStruct()
            s.kf = C.cast(C.pointer(st), C.c_void_p); s.points = p.ctypes.data; s.skip = sk.ctypes.data
            s.min_distance_inv = mn.ctypes.data; s.max_distance_inv = mx_inv.ctypes.data; s.max_distance = mx.ctypes.data
            s.log_scale_factor = float(np.log(np.float32(other.mvScaleFactors[1])))
            keep.extend([p, sk, mn, mx_inv, mx, st])
            return s
        s1 = side(KF1, KF2, points1, skip1, min_max_distance1); s2 = side(KF2, KF1, points2, skip2, min_max_distance2)
        a = [f32(sR12).ravel(), f32(t12).ravel(), f32(sR21).ravel(), f32(t21).ravel()]
        out = np.zeros(max(KF1.N, 1), np.int32); n = C.c_int()
        check(c._L.orbm_search_by_sim3(c.h, C.byref(s1), C.byref(s2), a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, a[3].ctypes.data,
                                       th, out.ctypes.data, C.byref(n)), "orbm_search_by_sim3")
        return n.value, out[: KF1.N]

    def _bow(self, variant, desc1, angle1, valid1, fv1: FeatureVector, desc2, angle2, valid2, fv2: FeatureVector):
        c = self._c()
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32); d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        a1 = np.ascontiguousarray(angle1, np.float32); a2 = np.ascontiguousarray(angle2, np.float32)
        v1 = np.ascontiguousarray(valid1, np.uint8)
        v2 = np.ascontiguousarray(valid2, np.uint8) if valid2 is not None else None
        out = np.zeros(max(len(d2) if variant == 0 else len(d1), 1), np.int32); n = C.c_int()
        s1, s2 = fv1.struct(), fv2.struct()
        check(c._L.orbm_search_by_bow(c.h, variant, d1.ctypes.data, a1.ctypes.data, v1.ctypes.data, len(d1), C.byref(s1),
                                      d2.ctypes.data, a2.ctypes.data, v2.ctypes.data if v2 is not None else None, len(d2),
                                      C.byref(s2), self.mfNNratio, int(self.mbCheckOrientation), out.ctypes.data, C.byref(n)),
              "orbm_search_by_bow")
        return n.value, out[: (len(d2) if variant == 0 else len(d1))]

    def SearchByBoWFrame(self, descKF, angleKFun, validKF, fvKF: FeatureVector, descF, angleF, fvF: FeatureVector):
        """SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches) (src/ORBmatcher.cc:159-288).  Returns (nmatches,
        match) with match[iF] = key-frame feature index whose map point lands in vpMapPointMatches[iF], -1 for NULL."""
        return self._bow(0, descKF, angleKFun, validKF, fvKF, descF, angleF, None, fvF)

    def SearchByBoWKeyFrames(self, desc1, angle1, valid1, fv1: FeatureVector, desc2, angle2, valid2, fv2: FeatureVector):
        """SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12) (:590-723).  match[idx1] = idx2 or -1."""
        return self._bow(1, desc1, angle1, valid1, fv1, desc2, angle2, valid2, fv2)

    def SearchForTriangulation(self, kf1: KeyFrameView, kf2: KeyFrameView, Cw1, R2w, t2w, K2, scale_factors2, level_sigma2_2, F12,
                               bOnlyStereo: bool = False):
        """SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:725-891).
        Returns (nmatches, vMatches12); vMatchedPairs = [(i, m) for i, m in enumerate(vMatches12) if m >= 0]."""
        c = self._c()
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        Cw1, R2w, t2w, F12 = f32(Cw1).ravel(), f32(R2w).ravel(), f32(t2w).ravel(), f32(F12).ravel()
        sf, sg = f32(scale_factors2), f32(level_sigma2_2)
        out = np.zeros(max(kf1.N, 1), np.int32); n = C.c_int()
        s1, s2 = kf1.struct(), kf2.struct()
        check(c._L.orbm_search_for_triangulation(c.h, C.byref(s1), C.byref(s2), Cw1.ctypes.data, R2w.ctypes.data, t2w.ctypes.data,
                                                 float(K2[0, 0]), float(K2[1, 1]), float(K2[0, 2]), float(K2[1, 2]), sf.ctypes.data,
                                                 sg.ctypes.data, len(sf), F12.ctypes.data, int(bOnlyStereo), int(self.mbCheckOrientation),
                                                 out.ctypes.data, C.byref(n)), "orbm_search_for_triangulation")
        return n.value, out[: kf1.N]
