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


def _bind_ref(path) -> C.CDLL:
    L = C.CDLL(path)
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    L.orbref_create.restype = vp; L.orbref_create.argtypes = [i, f, i, i, i]
    L.orbref_destroy.argtypes = [vp]
    L.orbref_extract.argtypes = [vp, vp, i, i, i, i, vp, vp, i]
    L.orbref_level_size.argtypes = [vp, i, vp, vp]
    L.orbref_get_level.argtypes = [vp, i, i, vp, i]
    return L


def ref() -> C.CDLL:
    global _ref
    if _ref is None:
        _ref = _bind_ref(REF_SO)
    return _ref


# ---- the reference built a second time for TIMING with its own flags (-O3 -march=..., oracle/Makefile ref_timed) ----
_X86_V3 = {"avx", "avx2", "bmi1", "bmi2", "f16c", "fma", "abm", "movbe", "xsave"}
_X86_V4 = _X86_V3 | {"avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"}
_ISA_PREFIXES = ("avx", "amx", "bmi", "fma", "f16c", "gfni", "vaes", "vpclmul", "sha", "sse", "ssse", "popcnt", "movbe", "abm",
                 "adx", "aes", "pclmul", "rdrand", "rdseed", "clwb", "clflushopt", "movdir", "cldemote", "serialize", "xsave",
                 "rdpid", "fsgsbase", "pku", "waitpkg", "enqcmd", "uintr", "hreset", "ptwrite", "wbnoinvd", "lahf", "cx16")


def _cpu_flags():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("flags"):
                return set(ln.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def timed_variant():
    """(directory, description) of the fastest reference build this host CPU can execute: this container's
    -march=native build when the host has every ISA flag the build machine had, else x86-64-v4 / -v3, else the
    portable parity build (-O3, no -march).  Chosen by cpuid flags only; nothing is executed to probe."""
    base = os.path.join(ROOT, "oracle", "_ref")
    have = _cpu_flags()
    cands = []
    nat = os.path.join(base, "timed_native")
    try:
        built = set(open(os.path.join(nat, "build_cpu_flags.txt")).read().split(":", 1)[1].split())
        need = {f for f in built if f.startswith(_ISA_PREFIXES)}
        cands.append((nat, need))
    except (OSError, IndexError):
        pass
    cands.append((os.path.join(base, "timed_v4"), _X86_V4))
    cands.append((os.path.join(base, "timed_v3"), _X86_V3))
    for d, need in cands:
        if os.path.exists(os.path.join(d, "liborb_ref.so")) and need <= have:
            flags = "-O3 -march=?"
            try:
                flags = open(os.path.join(d, "cxxflags.txt")).read().strip()
                if d == nat:
                    flags += " (" + open(os.path.join(d, "march.txt")).read().split()[-1] + ")"
            except (OSError, IndexError):
                pass
            return d, flags
    return base, "-O3 -ffp-contract=off (portable parity build, no -march)"


_timed = {}


def ref_timed() -> tuple:
    """(bound liborb_ref.so of the timed build, flags string)"""
    d, flags = timed_variant()
    if d not in _timed:
        _timed[d] = _bind_ref(os.path.join(d, "liborb_ref.so"))
    return _timed[d], flags


def ref_timed_bf2nn():
    """(callable(q, db) -> orbm_top2 array running orbref_bf2nn of the timed matcher build, flags) or (None, why)"""
    d, flags = timed_variant()
    so = os.path.join(d, "liborbmatcher_ref.so")
    if not os.path.exists(so):
        return None, "no reference matcher build"
    L = C.CDLL(so)
    L.orbref_bf2nn.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_uint32, C.c_void_p]
    from orb_slam_b200.matcher import TOP2_DTYPE

    def run(q, db):
        q = np.ascontiguousarray(q, np.uint8); db = np.ascontiguousarray(db, np.uint8)
        out = np.zeros(len(q), TOP2_DTYPE)
        L.orbref_bf2nn(q.ctypes.data, len(q), db.ctypes.data, len(db), 0, out.ctypes.data)
        return out
    return run, flags


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

    def __init__(self, nfeatures, scaleFactor, nlevels, iniTh, minTh, lib=None):
        self.L = lib if lib is not None else ref()          # lib: a timed build (ref_timed()), bench.py only
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


def oracle_stereo_matches(exL, exR, kL, dL, kR, dR, mbf, mb):
    """Frame::ComputeStereoMatches restated (oracle/orb_oracle_stereo.cc) on the pyramids of two OracleExtractor
    instances that have just extracted the left / right image."""
    L = oracle()
    n = exL.nlevels
    levL = [exL.get_level(l, border=19) for l in range(n)]
    levR = [exR.get_level(l, border=19) for l in range(n)]
    ptr = lambda a: a.ctypes.data + 19 * a.strides[0] + 19
    pL = (C.c_void_p * n)(*[ptr(a) for a in levL]); pR = (C.c_void_p * n)(*[ptr(a) for a in levR])
    sL = (C.c_size_t * n)(*[a.strides[0] for a in levL]); sR = (C.c_size_t * n)(*[a.strides[0] for a in levR])
    cols = (C.c_int * n)(*[a.shape[1] - 38 for a in levL])
    sf, isf = exL.tables()[0], exL.tables()[1]
    kL = np.ascontiguousarray(kL, KP); kR = np.ascontiguousarray(kR, KP)
    dL = np.ascontiguousarray(dL, np.uint8); dR = np.ascontiguousarray(dR, np.uint8)
    u = np.zeros(len(kL), np.float32); z = np.zeros(len(kL), np.float32)
    L.orbo_compute_stereo_matches.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    kept = L.orbo_compute_stereo_matches(kL.ctypes.data, len(kL), dL.ctypes.data, kR.ctypes.data, len(kR), dR.ctypes.data,
                                         sf.ctypes.data, isf.ctypes.data, n, pL, pR, sL, sR, cols, levL[0].shape[0] - 38,
                                         mbf, mb, u.ctypes.data, z.ctypes.data)
    return u, z, kept


# ---------------- matcher / grid oracle (oracle/orb_oracle_matcher.cc) ----------------
def _bind_matcher(L):
    if getattr(L, "_m_bound", False):
        return L
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    L.orbo_descriptor_distance.argtypes = [vp, vp]
    L.orbo_three_maxima.argtypes = [vp, i, vp, vp, vp]
    L.orbo_grid_build.argtypes = [vp, vp, vp]
    L.orbo_features_in_area.argtypes = [vp, f, f, f, i, i, vp, i]
    L.orbo_search_for_initialization.argtypes = [vp, vp, vp, vp, i, f, i]
    L.orbo_search_by_projection_frame.argtypes = [vp, vp, vp, vp, f, i, i, vp]
    L.orbo_search_by_projection_points.argtypes = [vp, vp, i, vp, f, f, vp]
    L.orbo_bf2nn.argtypes = [vp, i, vp, C.c_size_t, C.c_uint32, vp]
    L.orbo_search_by_projection_keyframe.argtypes = [vp, vp, i, vp, f, vp, vp, vp, f, i, i, vp]
    L.orbo_search_by_projection_sim3.argtypes = [vp, vp, vp, vp, vp, i, vp, vp, vp, f, vp, i, vp]
    L.orbo_fuse_select_sim3.argtypes = [vp, vp, vp, vp, vp, i, vp, vp, vp, f, f, vp]
    L.orbo_search_by_sim3.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, f, f, vp, vp, vp, vp, f, vp]
    L.orbo_fuse_select.argtypes = [vp, vp, vp, vp, i, vp, vp, vp, f, f, vp]
    L.orbo_search_for_triangulation.argtypes = [vp, vp, vp, vp, vp, f, f, f, f, vp, vp, vp, i, i, vp]
    L.orbo_search_by_bow_kf_frame.argtypes = [vp, vp, vp, i, vp, vp, vp, i, vp, f, i, vp]
    L.orbo_search_by_bow_kf_kf.argtypes = [vp, vp, vp, i, vp, vp, vp, vp, i, vp, f, i, vp]
    L._m_bound = True
    return L


class OracleMatcher:
    """CPU oracle with the ORBmatcher call shape; frames are orb_slam_b200.matcher.Frame objects
    (plain data carriers: the struct layout of include/orbm.h)."""

    def __init__(self, nnratio=0.6, checkOri=True):
        self.L = _bind_matcher(oracle())
        self.nnratio = float(np.float32(nnratio)); self.checkOri = bool(checkOri)

    def DescriptorDistance(self, a, b):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return self.L.orbo_descriptor_distance(a.ctypes.data, b.ctypes.data)

    def ComputeThreeMaxima(self, counts):
        c = np.ascontiguousarray(counts, np.int32)
        a, b, d = C.c_int(-1), C.c_int(-1), C.c_int(-1)
        self.L.orbo_three_maxima(c.ctypes.data, len(c), C.byref(a), C.byref(b), C.byref(d))
        return a.value, b.value, d.value

    def grid_build(self, F):
        cs = np.zeros(64 * 48 + 1, np.int32); items = np.zeros(max(F.N, 1), np.int32)
        s = F.struct()
        self.L.orbo_grid_build(C.byref(s), cs.ctypes.data, items.ctypes.data)
        return cs, items[: cs[-1]]

    def features_in_area(self, F, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.zeros(max(F.N, 1), np.int32)
        s = F.struct()
        n = self.L.orbo_features_in_area(C.byref(s), x, y, r, minLevel, maxLevel, out.ctypes.data, len(out))
        return out[:n].copy()

    def SearchForInitialization(self, F1, F2, prev, windowSize=10):
        m12 = np.zeros(max(F1.N, 1), np.int32)
        s1, s2 = F1.struct(), F2.struct()
        n = self.L.orbo_search_for_initialization(C.byref(s1), C.byref(s2), prev.ctypes.data, m12.ctypes.data, windowSize,
                                                  self.nnratio, int(self.checkOri))
        return n, m12[: F1.N]

    def SearchByProjectionFrame(self, cur, last, last_points, cur_obs, th, bMono):
        out = np.zeros(max(cur.N, 1), np.int32)
        sc, sl = cur.struct(), last.struct()
        lp = np.ascontiguousarray(last_points); co = np.ascontiguousarray(cur_obs, np.int32)
        n = self.L.orbo_search_by_projection_frame(C.byref(sc), C.byref(sl), lp.ctypes.data, co.ctypes.data, th, int(bMono),
                                                   int(self.checkOri), out.ctypes.data)
        return n, out[: cur.N]

    def SearchByProjectionPoints(self, F, points, frame_obs, th=3.0):
        out = np.zeros(max(F.N, 1), np.int32)
        s = F.struct()
        p = np.ascontiguousarray(points); fo = np.ascontiguousarray(frame_obs, np.int32)
        n = self.L.orbo_search_by_projection_points(C.byref(s), p.ctypes.data, len(p), fo.ctypes.data, th, self.nnratio,
                                                    out.ctypes.data)
        return n, out[: F.N]

    def bf2nn(self, q, db, idx_base=0):
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32); db = np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
        out = np.zeros(len(q), np.dtype([("i1", "<u4"), ("d1", "<u2"), ("d2", "<u2")]))
        self.L.orbo_bf2nn(q.ctypes.data, len(q), db.ctypes.data, len(db), idx_base, out.ctypes.data)
        return out

    def SearchByProjectionKeyFrame(self, cur, kf_points, min_max_distance, already_found, kf_angle, cur_has_point, th, ORBdist,
                                   log_scale_factor=None):
        from orb_slam_b200.matcher import MAPPOINT_DTYPE
        p = np.ascontiguousarray(kf_points, MAPPOINT_DTYPE); mm = np.ascontiguousarray(min_max_distance, np.float32).reshape(-1, 2)
        af = np.ascontiguousarray(already_found, np.uint8); ang = np.ascontiguousarray(kf_angle, np.float32)
        has = np.ascontiguousarray(cur_has_point, np.uint8)
        if log_scale_factor is None:
            log_scale_factor = float(np.log(np.float32(cur.mvScaleFactors[1])))
        out = np.zeros(max(cur.N, 1), np.int32)
        s = cur.struct()
        n = self.L.orbo_search_by_projection_keyframe(C.byref(s), p.ctypes.data, len(p), mm.ctypes.data, log_scale_factor, af.ctypes.data,
                                                      ang.ctypes.data, has.ctypes.data, th, int(ORBdist), int(self.checkOri), out.ctypes.data)
        return n, out[: cur.N]

    def FuseSelect(self, KF, inv_level_sigma2, points, in_keyframe, normals, min_max_distance, th=3.0, log_scale_factor=None,
                   camera_center=None):
        from orb_slam_b200.matcher import MAPPOINT_DTYPE
        p = np.ascontiguousarray(points, MAPPOINT_DTYPE); mm = np.ascontiguousarray(min_max_distance, np.float32).reshape(-1, 2)
        ik = np.ascontiguousarray(in_keyframe, np.uint8); nr = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        sg = np.ascontiguousarray(inv_level_sigma2, np.float32)
        if log_scale_factor is None:
            log_scale_factor = float(np.log(np.float32(KF.mvScaleFactors[1])))
        out = np.zeros(max(len(p), 1), np.int32)
        s = KF.struct()
        cc = None if camera_center is None else np.ascontiguousarray(camera_center, np.float32)
        self.L.orbo_fuse_select(C.byref(s), None if cc is None else cc.ctypes.data, sg.ctypes.data, p.ctypes.data, len(p), ik.ctypes.data, nr.ctypes.data, mm.ctypes.data,
                                log_scale_factor, th, out.ctypes.data)
        return out[: len(p)]

    def _sim3(self, KF, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, log_scale_factor):
        from orb_slam_b200.matcher import MAPPOINT_DTYPE
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        k = dict(R=f32(Rcw).ravel(), t=f32(tcw).ravel(), O=f32(Ow).ravel(), p=np.ascontiguousarray(points, MAPPOINT_DTYPE),
                 ex=np.ascontiguousarray(excluded, np.uint8), nr=f32(normals).reshape(-1, 3), mm=f32(min_max_distance).reshape(-1, 2), s=KF.struct())
        if log_scale_factor is None:
            log_scale_factor = float(np.log(np.float32(KF.mvScaleFactors[1])))
        args = [C.byref(k["s"]), k["R"].ctypes.data, k["t"].ctypes.data, k["O"].ctypes.data, k["p"].ctypes.data, len(k["p"]), k["ex"].ctypes.data,
                k["nr"].ctypes.data, k["mm"].ctypes.data, log_scale_factor]
        return k, args

    def SearchByProjectionSim3(self, KF, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, matched_in, th, log_scale_factor=None):
        k, args = self._sim3(KF, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, log_scale_factor)
        mi = np.ascontiguousarray(matched_in, np.uint8)
        out = np.zeros(max(KF.N, 1), np.int32)
        n = self.L.orbo_search_by_projection_sim3(*args, mi.ctypes.data, int(th), out.ctypes.data)
        return n, out[: KF.N]

    def FuseSelectSim3(self, KF, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, th, log_scale_factor=None):
        k, args = self._sim3(KF, Rcw, tcw, Ow, points, excluded, normals, min_max_distance, log_scale_factor)
        out = np.zeros(max(len(k["p"]), 1), np.int32)
        self.L.orbo_fuse_select_sim3(*args, th, out.ctypes.data)
        return out[: len(k["p"])]

    def SearchBySim3(self, KF1, KF2, points1, points2, skip1, skip2, mm1, mm2, sR12, t12, sR21, t21, th):
        from orb_slam_b200.matcher import MAPPOINT_DTYPE
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        p1 = np.ascontiguousarray(points1, MAPPOINT_DTYPE); p2 = np.ascontiguousarray(points2, MAPPOINT_DTYPE)
        k1 = np.ascontiguousarray(skip1, np.uint8); k2 = np.ascontiguousarray(skip2, np.uint8)
        m1 = f32(mm1).reshape(-1, 2); m2 = f32(mm2).reshape(-1, 2)
        a = [f32(sR12).ravel(), f32(t12).ravel(), f32(sR21).ravel(), f32(t21).ravel()]
        s1, s2 = KF1.struct(), KF2.struct()
        out = np.zeros(max(KF1.N, 1), np.int32)
        n = self.L.orbo_search_by_sim3(C.byref(s1), C.byref(s2), p1.ctypes.data, p2.ctypes.data, k1.ctypes.data, k2.ctypes.data,
                                       m1.ctypes.data, m2.ctypes.data, float(np.log(np.float32(KF1.mvScaleFactors[1]))),
                                       float(np.log(np.float32(KF2.mvScaleFactors[1]))), a[0].ctypes.data, a[1].ctypes.data,
                                       a[2].ctypes.data, a[3].ctypes.data, th, out.ctypes.data)
        return n, out[: KF1.N]

    def SearchByBoWFrame(self, descKF, angleKFun, validKF, fvKF, descF, angleF, fvF):
        d1 = np.ascontiguousarray(descKF, np.uint8).reshape(-1, 32); d2 = np.ascontiguousarray(descF, np.uint8).reshape(-1, 32)
        a1 = np.ascontiguousarray(angleKFun, np.float32); a2 = np.ascontiguousarray(angleF, np.float32)
        v1 = np.ascontiguousarray(validKF, np.uint8)
        out = np.zeros(max(len(d2), 1), np.int32)
        s1, s2 = fvKF.struct(), fvF.struct()
        n = self.L.orbo_search_by_bow_kf_frame(d1.ctypes.data, a1.ctypes.data, v1.ctypes.data, len(d1), C.byref(s1), d2.ctypes.data,
                                               a2.ctypes.data, len(d2), C.byref(s2), self.nnratio, int(self.checkOri), out.ctypes.data)
        return n, out[: len(d2)]

    def SearchByBoWKeyFrames(self, desc1, angle1, valid1, fv1, desc2, angle2, valid2, fv2):
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32); d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        a1 = np.ascontiguousarray(angle1, np.float32); a2 = np.ascontiguousarray(angle2, np.float32)
        v1 = np.ascontiguousarray(valid1, np.uint8); v2 = np.ascontiguousarray(valid2, np.uint8)
        out = np.zeros(max(len(d1), 1), np.int32)
        s1, s2 = fv1.struct(), fv2.struct()
        n = self.L.orbo_search_by_bow_kf_kf(d1.ctypes.data, a1.ctypes.data, v1.ctypes.data, len(d1), C.byref(s1), d2.ctypes.data,
                                            a2.ctypes.data, v2.ctypes.data, len(d2), C.byref(s2), self.nnratio, int(self.checkOri),
                                            out.ctypes.data)
        return n, out[: len(d1)]

    def SearchForTriangulation(self, kf1, kf2, Cw1, R2w, t2w, K2, scale_factors2, level_sigma2_2, F12, bOnlyStereo=False):
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        Cw1, R2w, t2w, F12 = f32(Cw1).ravel(), f32(R2w).ravel(), f32(t2w).ravel(), f32(F12).ravel()
        sf, sg = f32(scale_factors2), f32(level_sigma2_2)
        out = np.zeros(max(kf1.N, 1), np.int32)
        s1, s2 = kf1.struct(), kf2.struct()
        n = self.L.orbo_search_for_triangulation(C.byref(s1), C.byref(s2), Cw1.ctypes.data, R2w.ctypes.data, t2w.ctypes.data,
                                                 float(K2[0, 0]), float(K2[1, 1]), float(K2[0, 2]), float(K2[1, 2]), sf.ctypes.data,
                                                 sg.ctypes.data, F12.ctypes.data, int(bOnlyStereo), int(self.checkOri), out.ctypes.data)
        return n, out[: kf1.N]


# ---------------- vocabulary (oracle/orb_oracle_vocab.cc) and the reference's own DBoW2 (oracle/_ref/libdbow2_ref.so)
DBOW2_REF_SO = os.path.join(ROOT, "oracle", "_ref", "libdbow2_ref.so")
_dbow2 = None


def have_dbow2_ref() -> bool:
    return os.path.exists(DBOW2_REF_SO)


def _flat_transform(fn, h, desc, levelsup):
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(d)
    wid = np.zeros(max(n, 1), np.uint32); wv = np.zeros(max(n, 1), np.float64)
    nid = np.zeros(max(n, 1), np.uint32); st = np.zeros(n + 1, np.int32); it = np.zeros(max(n, 1), np.uint32)
    nw, nn = C.c_int(), C.c_int()
    fn(h, d.ctypes.data, n, levelsup, wid.ctypes.data, wv.ctypes.data, C.byref(nw), nid.ctypes.data, st.ctypes.data,
       it.ctypes.data, C.byref(nn))
    fv = {int(nid[j]): it[st[j]:st[j + 1]].tolist() for j in range(nn.value)}
    return wid[:nw.value].copy(), wv[:nw.value].copy(), fv


class OracleVocabulary:
    """CPU restatement of ORBVocabulary (oracle/orb_oracle_vocab.cc)."""

    def __init__(self, handle):
        self._h = handle

    @staticmethod
    def _lib():
        L = oracle()
        if not getattr(L, "_vocab_bound", False):
            vp, i = C.c_void_p, C.c_int
            L.orbo_vocab_create.restype = vp; L.orbo_vocab_create.argtypes = [i, i, i, i, i, vp, vp, vp, vp]
            L.orbo_vocab_load_text.restype = vp; L.orbo_vocab_load_text.argtypes = [C.c_char_p]
            L.orbo_vocab_destroy.argtypes = [vp]; L.orbo_vocab_destroy.restype = None
            L.orbo_vocab_words.argtypes = [vp]; L.orbo_vocab_nodes.argtypes = [vp]
            L.orbo_vocab_descend.argtypes = [vp, vp, i, i, vp, vp, vp]; L.orbo_vocab_descend.restype = None
            L.orbo_vocab_transform.argtypes = [vp, vp, i, i, vp, vp, C.POINTER(i), vp, vp, vp, C.POINTER(i)]
            L.orbo_distinctive_descriptors.argtypes = [vp, vp, i, vp]; L.orbo_distinctive_descriptors.restype = None
            L._vocab_bound = True
        return L

    @classmethod
    def from_arrays(cls, k, L, parent, is_leaf, desc, weight, scoring=0, weighting=0):
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); weight = np.ascontiguousarray(weight, np.float64)
        h = cls._lib().orbo_vocab_create(k, L, scoring, weighting, len(parent), parent.ctypes.data, is_leaf.ctypes.data,
                                         desc.ctypes.data, weight.ctypes.data)
        return cls(C.c_void_p(h))

    @classmethod
    def loadFromTextFile(cls, path):
        h = cls._lib().orbo_vocab_load_text(str(path).encode())
        assert h, path
        return cls(C.c_void_p(h))

    def __del__(self):
        try:
            if self._h:
                self._lib().orbo_vocab_destroy(self._h); self._h = None
        except Exception:
            pass

    def size(self):
        return self._lib().orbo_vocab_words(self._h)

    def nodes(self):
        return self._lib().orbo_vocab_nodes(self._h)

    def descend(self, desc, levelsup=4):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(d)
        w = np.zeros(n, np.uint32); nid = np.zeros(n, np.uint32); wt = np.zeros(n, np.float64)
        self._lib().orbo_vocab_descend(self._h, d.ctypes.data, n, levelsup, w.ctypes.data, nid.ctypes.data, wt.ctypes.data)
        return w, nid, wt

    def transform(self, desc, levelsup=4):
        return _flat_transform(self._lib().orbo_vocab_transform, self._h, desc, levelsup)


def oracle_distinctive_descriptors(desc, start):
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    start = np.ascontiguousarray(start, np.int32)
    best = np.zeros(max(len(start) - 1, 1), np.int32)
    OracleVocabulary._lib().orbo_distinctive_descriptors(d.ctypes.data, start.ctypes.data, len(start) - 1, best.ctypes.data)
    return best[:len(start) - 1]


class RefVocabulary:
    """The reference's own DBoW2 TemplatedVocabulary<FORB::TDescriptor, FORB> (oracle/_ref/libdbow2_ref.so)."""

    def __init__(self, path):
        global _dbow2
        if _dbow2 is None:
            L = C.CDLL(DBOW2_REF_SO)
            vp, i = C.c_void_p, C.c_int
            L.refvoc_load_text.restype = vp; L.refvoc_load_text.argtypes = [C.c_char_p]
            L.refvoc_free.argtypes = [vp]; L.refvoc_free.restype = None
            L.refvoc_size.argtypes = [vp]; L.refvoc_size.restype = C.c_uint
            L.refvoc_transform.argtypes = [vp, vp, i, i, vp, vp, C.POINTER(i), vp, vp, vp, C.POINTER(i)]
            _dbow2 = L
        self._L = _dbow2
        self._h = C.c_void_p(self._L.refvoc_load_text(str(path).encode()))
        assert self._h, path

    def __del__(self):
        try:
            if self._h:
                self._L.refvoc_free(self._h); self._h = None
        except Exception:
            pass

    def size(self):
        return self._L.refvoc_size(self._h)

    def transform(self, desc, levelsup=4):
        return _flat_transform(self._L.refvoc_transform, self._h, desc, levelsup)


# ---------------- the reference's own ORBmatcher.cc / Frame.cc / KeyFrame.cc / MapPoint.cc (oracle/_ref/liborbmatcher_ref.so)
MATCHER_REF_SO = os.path.join(ROOT, "oracle", "_ref", "liborbmatcher_ref.so")
_matcher_ref = None


def have_matcher_ref() -> bool:
    return os.path.exists(MATCHER_REF_SO)


class _RefProxy:
    """Looks up orbo_X as orbref_X in the reference build, with the oracle's argtypes (the signatures are identical)."""

    def __init__(self):
        global _matcher_ref
        if _matcher_ref is None:
            _matcher_ref = C.CDLL(MATCHER_REF_SO)
        self._ref, self._oracle = _matcher_ref, _bind_matcher(oracle())

    def __getattr__(self, name):
        fn = getattr(self._ref, name.replace("orbo_", "orbref_"))      # AttributeError when the wrapper lacks it
        fn.argtypes = getattr(self._oracle, name).argtypes
        return fn


class RefMatcher(OracleMatcher):
    """Same call shape as OracleMatcher, executed by the reference's own object code."""

    def __init__(self, nnratio=0.6, checkOri=True):
        self.L = _RefProxy()
        self.nnratio = float(np.float32(nnratio)); self.checkOri = bool(checkOri)


def ref_distinctive_descriptors(desc, start):
    """MapPoint::ComputeDistinctiveDescriptors of the reference build: (chosen descriptor bytes, has flag, observation order)."""
    global _matcher_ref
    if _matcher_ref is None:
        _matcher_ref = C.CDLL(MATCHER_REF_SO)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); start = np.ascontiguousarray(start, np.int32)
    npts = len(start) - 1
    chosen = np.zeros((max(npts, 1), 32), np.uint8); has = np.zeros(max(npts, 1), np.uint8); order = np.zeros(max(len(d), 1), np.int32)
    vp = C.c_void_p
    _matcher_ref.orbref_distinctive_descriptors.argtypes = [vp, vp, C.c_int, vp, vp, vp]
    _matcher_ref.orbref_distinctive_descriptors(d.ctypes.data, start.ctypes.data, npts, chosen.ctypes.data, has.ctypes.data, order.ctypes.data)
    return chosen[:npts], has[:npts], order[:len(d)]


def ref_stereo(imgL, imgR, nfeatures, scaleFactor, nlevels, iniTh, minTh, mbf, mb):
    """Reference extraction of both images + the reference's own Frame::ComputeStereoMatches.
    Returns (kL, dL, kR, dR, levelsL, levelsR, mvuRight, mvDepth); levels are lists of (h_l, w_l) arrays."""
    global _matcher_ref
    if _matcher_ref is None:
        _matcher_ref = C.CDLL(MATCHER_REF_SO)
    imgL = np.ascontiguousarray(imgL, np.uint8); imgR = np.ascontiguousarray(imgR, np.uint8)
    h, w = imgL.shape
    cap = nfeatures * 2 + 64
    kL = np.zeros(cap, KP); kR = np.zeros(cap, KP); dL = np.zeros((cap, 32), np.uint8); dR = np.zeros((cap, 32), np.uint8)
    lvL = np.zeros(w * h * 4, np.uint8); lvR = np.zeros(w * h * 4, np.uint8); dims = np.zeros(2 * nlevels, np.int32)
    u = np.zeros(cap, np.float32); z = np.zeros(cap, np.float32)
    nL, nR = C.c_int(), C.c_int()
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    fn = _matcher_ref.orbref_compute_stereo_matches
    fn.argtypes = [vp, vp, i, i, i, f, i, i, i, f, f, i, vp, vp, C.POINTER(i), vp, vp, C.POINTER(i), vp, vp, vp, vp, vp]
    rc = fn(imgL.ctypes.data, imgR.ctypes.data, w, h, nfeatures, scaleFactor, nlevels, iniTh, minTh, mbf, mb, cap, kL.ctypes.data,
            dL.ctypes.data, C.byref(nL), kR.ctypes.data, dR.ctypes.data, C.byref(nR), lvL.ctypes.data, lvR.ctypes.data, dims.ctypes.data,
            u.ctypes.data, z.ctypes.data)
    assert rc == 0
    levelsL, levelsR, off = [], [], 0
    for l in range(nlevels):
        wl, hl = int(dims[2 * l]), int(dims[2 * l + 1])
        levelsL.append(lvL[off:off + wl * hl].reshape(hl, wl).copy()); levelsR.append(lvR[off:off + wl * hl].reshape(hl, wl).copy())
        off += wl * hl
    return kL[:nL.value], dL[:nL.value], kR[:nR.value], dR[:nR.value], levelsL, levelsR, u[:nL.value], z[:nL.value]


def oracle_stereo_matches_on_levels(kL, dL, kR, dR, levelsL, levelsR, scale_factors, mbf, mb):
    """orbo_compute_stereo_matches on explicit (unbordered) level images."""
    L = oracle()
    n = len(levelsL)
    levelsL = [np.ascontiguousarray(a) for a in levelsL]; levelsR = [np.ascontiguousarray(a) for a in levelsR]
    pL = (C.c_void_p * n)(*[a.ctypes.data for a in levelsL]); pR = (C.c_void_p * n)(*[a.ctypes.data for a in levelsR])
    sL = (C.c_size_t * n)(*[a.strides[0] for a in levelsL]); sR = (C.c_size_t * n)(*[a.strides[0] for a in levelsR])
    cols = (C.c_int * n)(*[a.shape[1] for a in levelsL])
    sf = np.ascontiguousarray(scale_factors, np.float32); isf = (np.float32(1.0) / sf).astype(np.float32)
    kL = np.ascontiguousarray(kL, KP); kR = np.ascontiguousarray(kR, KP)
    dL = np.ascontiguousarray(dL, np.uint8); dR = np.ascontiguousarray(dR, np.uint8)
    u = np.zeros(len(kL), np.float32); z = np.zeros(len(kL), np.float32)
    L.orbo_compute_stereo_matches.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    kept = L.orbo_compute_stereo_matches(kL.ctypes.data, len(kL), dL.ctypes.data, kR.ctypes.data, len(kR), dR.ctypes.data,
                                         sf.ctypes.data, isf.ctypes.data, n, pL, pR, sL, sR, cols, levelsL[0].shape[0],
                                         mbf, mb, u.ctypes.data, z.ctypes.data)
    return u, z, kept


def _mref():
    global _matcher_ref
    if _matcher_ref is None:
        _matcher_ref = C.CDLL(MATCHER_REF_SO)
    return _matcher_ref


def ref_decompose_scw(Scw):
    S = np.ascontiguousarray(Scw, np.float32).reshape(4, 4)
    R = np.zeros((3, 3), np.float32); t = np.zeros(3, np.float32); O = np.zeros(3, np.float32)
    fn = _mref().orbref_decompose_scw
    fn.argtypes = [C.c_void_p] * 4; fn.restype = None
    fn(S.ctypes.data, R.ctypes.data, t.ctypes.data, O.ctypes.data)
    return R, t, O


def ref_sim3_projection(KF, Scw, points, excluded, normals, min_max_distance, matched_in, th, fuse=False):
    """The reference's SearchByProjection(pKF, Scw, ...) (fuse=False) / Fuse(pKF, Scw, ...) (fuse=True)."""
    from orb_slam_b200.matcher import MAPPOINT_DTYPE
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    S = f32(Scw).reshape(4, 4); p = np.ascontiguousarray(points, MAPPOINT_DTYPE); ex = np.ascontiguousarray(excluded, np.uint8)
    nr = f32(normals).reshape(-1, 3); mm = f32(min_max_distance).reshape(-1, 2); s = KF.struct()
    lsf = float(np.log(np.float32(KF.mvScaleFactors[1])))
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    if fuse:
        out = np.zeros(max(len(p), 1), np.int32)
        fn = _mref().orbref_fuse_select_sim3
        fn.argtypes = [vp, vp, vp, i, vp, vp, vp, f, f, vp]
        n = fn(C.byref(s), S.ctypes.data, p.ctypes.data, len(p), ex.ctypes.data, nr.ctypes.data, mm.ctypes.data, lsf, float(th), out.ctypes.data)
        return n, out[:len(p)]
    mi = np.ascontiguousarray(matched_in, np.uint8)
    out = np.zeros(max(KF.N, 1), np.int32)
    fn = _mref().orbref_search_by_projection_sim3
    fn.argtypes = [vp, vp, vp, i, vp, vp, vp, f, vp, i, vp]
    n = fn(C.byref(s), S.ctypes.data, p.ctypes.data, len(p), ex.ctypes.data, nr.ctypes.data, mm.ctypes.data, lsf, mi.ctypes.data, int(th),
           out.ctypes.data)
    return n, out[:KF.N]


def ref_search_by_sim3(KF1, KF2, points1, points2, skip1, skip2, mm1, mm2, s12, R12, t12, th):
    from orb_slam_b200.matcher import MAPPOINT_DTYPE
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    p1 = np.ascontiguousarray(points1, MAPPOINT_DTYPE); p2 = np.ascontiguousarray(points2, MAPPOINT_DTYPE)
    k1 = np.ascontiguousarray(skip1, np.uint8); k2 = np.ascontiguousarray(skip2, np.uint8)
    m1 = f32(mm1).reshape(-1, 2); m2 = f32(mm2).reshape(-1, 2); R = f32(R12).reshape(3, 3); t = f32(t12).ravel()
    s1, s2 = KF1.struct(), KF2.struct()
    out = np.zeros(max(KF1.N, 1), np.int32)
    vp, f = C.c_void_p, C.c_float
    fn = _mref().orbref_search_by_sim3
    fn.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, f, f, f, vp, vp, f, vp]
    n = fn(C.byref(s1), C.byref(s2), p1.ctypes.data, p2.ctypes.data, k1.ctypes.data, k2.ctypes.data, m1.ctypes.data, m2.ctypes.data,
           float(np.log(np.float32(KF1.mvScaleFactors[1]))), float(np.log(np.float32(KF2.mvScaleFactors[1]))), float(s12), R.ctypes.data,
           t.ctypes.data, float(th), out.ctypes.data)
    return n, out[:KF1.N]
