"""The matcher scenes behind tests/golden/ref_matcher_hashes.json, written once for three executors with the same call
shape: the REFERENCE's own object code (tests.oracle_lib.RefMatcher, only where /root/reference was compiled: it
produced the committed hashes, tests/golden/make_matcher_golden.py), the CPU oracle (OracleMatcher) and the CUDA
product (orb_slam_b200.matcher.ORBmatcher).  Every case returns a list of integer / float32 arrays; their sha256 is
what is compared, so a GPU box without /root/reference still checks oracle == reference and CUDA == reference."""
import hashlib

import numpy as np

from tests import scenes
from tests.oracle_lib import OracleExtractor


def digest(arrays):
    h = hashlib.sha256()
    for a in arrays:
        a = np.ascontiguousarray(a)
        h.update(str(a.dtype).encode()); h.update(str(a.shape).encode()); h.update(a.tobytes())
    return h.hexdigest()


def _grid(M):
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 752, 480, 1000, 30, (7, -4))
    F2 = scenes.make_frame(kb, db, 752, 480)
    m = M(0.6, True)
    out = list(m.grid_build(F2))
    rng = np.random.default_rng(3)
    for t in range(40):
        x, y = float(rng.uniform(-30, 790)), float(rng.uniform(-30, 510))
        rad = float(rng.choice([5, 15, 40, 100, 1000]))
        lv = [(-1, -1), (0, 0), (1, 3), (2, -1), (0, 7)][t % 5]
        out.append(np.asarray(m.features_in_area(F2, x, y, rad, *lv), np.int32))
    return out


def _init(ratio, ori, win):
    def run(M):
        (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 752, 480, 1000, 30, (7, -4))
        F1, F2 = scenes.make_frame(ka, da, 752, 480), scenes.make_frame(kb, db, 752, 480)
        prev = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
        n, m12 = M(ratio, ori).SearchForInitialization(F1, F2, prev, win)
        return [np.int32(n), m12, prev]
    return run


def _proj_frame(stereo, th, mono, tz):
    def run(M):
        (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 31, (3, 1))
        last = scenes.make_frame(ka, da, 1241, 376, stereo=stereo)
        Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [0.02, -0.01, tz]
        cur = scenes.make_frame(kb, db, 1241, 376, stereo=stereo, Tcw=Tcw, seed=1)
        pts = scenes.last_frame_points(last, seed=2)
        cur_obs = np.random.default_rng(4).choice([-1, -1, -1, 0, 2], cur.N).astype(np.int32)
        n, a = M(0.9, True).SearchByProjectionFrame(cur, last, pts, cur_obs, th, mono)
        # the reference leaves no trace of "assigned, then set to NULL by the rotation check" on a slot that was NULL before
        return [np.int32(n), np.where((cur_obs < 0) & (a == -1), -2, a).astype(np.int32)]
    return run


def _proj_points(th, ratio):
    def run(M):
        (ka, da), _ = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 32, (0, 0))
        F = scenes.make_frame(ka, da, 1241, 376, stereo=True)
        pts = scenes.local_map_points(F, seed=5)
        obs = np.random.default_rng(6).choice([-1, -1, 0, 3], F.N).astype(np.int32)
        n, a = M(ratio, True).SearchByProjectionPoints(F, pts, obs, th)
        return [np.int32(n), a]
    return run


def _bow(case, nnratio, ori):
    def run(M):
        s = scenes.bow_scene(**case)
        m = M(nnratio, ori)
        a = m.SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["fv2"])
        b = m.SearchByBoWKeyFrames(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["valid2"], s["fv2"])
        return [np.int32(a[0]), a[1], np.int32(b[0]), b[1]]
    return run


def _proj_keyframe(th, orbdist, tz):
    def run(M):
        (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 33, (4, -2))
        kf = scenes.make_frame(ka, da, 1241, 376)
        Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [0.03, 0.01, tz]
        cur = scenes.make_frame(kb, db, 1241, 376, Tcw=Tcw, seed=3)
        pts, mm, already = scenes.keyframe_points(kf, seed=11)
        has = (np.random.default_rng(12).random(cur.N) < 0.3).astype(np.uint8)
        n, a = M(0.9, True).SearchByProjectionKeyFrame(cur, pts, mm, already, kf.mvKeysUn["angle"], has, th, orbdist)
        return [np.int32(n), np.where((has == 0) & (a == -1), -2, a).astype(np.int32)]
    return run


def _fuse(stereo, th, tz):
    def run(M):
        (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 35, (-3, 2))
        src = scenes.make_frame(ka, da, 1241, 376)
        Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [-0.05, 0.02, tz]          # identity rotation: Ow = -tcw in any arithmetic
        kf = scenes.make_frame(kb, db, 1241, 376, stereo=stereo, Tcw=Tcw, seed=7)
        if not stereo:
            kf.mvuRight = np.full(kf.N, -1, np.float32)
        pts, mm, normals, in_kf = scenes.fuse_points(src, seed=21)
        sf = kf.mvScaleFactors
        inv_sigma2 = (1.0 / (sf * sf)).astype(np.float32)
        return [M(0.6, True).FuseSelect(kf, inv_sigma2, pts, in_kf, normals, mm, th)]
    return run


def _triang(seed, only_stereo, ori):
    def run(M):
        kw = dict(t=(0.02, 0.01, 0.6), stereo_frac=0.0) if seed == 4 else {}
        s = scenes.triangulation_scene(seed, **kw)
        n, m = M(0.6, ori).SearchForTriangulation(s["kf1"], s["kf2"], s["Cw1"], s["R2w"], s["t2w"], s["K2"], s["sf"], s["sig2"], s["F12"],
                                                  only_stereo)
        return [np.int32(n), m]
    return run


BOW_CASES = [dict(seed=1), dict(seed=3, n1=300, n2=1800, n_nodes=7, ties=200), dict(seed=8, n1=1200, n2=1100, flips=60, n_nodes=16, ties=50)]

CASES = {"grid_and_features_in_area": _grid}
for _r, _o, _w in [(0.9, True, 100), (0.9, False, 100), (0.6, True, 30)]:
    CASES["search_for_initialization[%g,%d,%d]" % (_r, _o, _w)] = _init(_r, _o, _w)
for _s, _th, _m, _tz in [(False, 15.0, True, 0.05), (True, 7.0, False, -0.9), (True, 14.0, False, 0.9)]:
    CASES["search_by_projection_frame[%d,%g,%d,%g]" % (_s, _th, _m, _tz)] = _proj_frame(_s, _th, _m, _tz)
for _th, _r in [(1.0, 0.8), (3.0, 0.8), (5.0, 0.6)]:
    CASES["search_by_projection_points[%g,%g]" % (_th, _r)] = _proj_points(_th, _r)
for _c in BOW_CASES:
    for _r, _o in [(0.7, True), (0.9, False)]:
        CASES["search_by_bow[s%d,%g,%d]" % (_c["seed"], _r, _o)] = _bow(_c, _r, _o)
for _th, _d, _tz in [(10.0, 100, 0.3), (3.0, 64, -0.4)]:
    CASES["search_by_projection_keyframe[%g,%d,%g]" % (_th, _d, _tz)] = _proj_keyframe(_th, _d, _tz)
for _s, _th, _tz in [(True, 3.0, 0.2), (False, 3.0, -0.3)]:
    CASES["fuse_select[%d,%g,%g]" % (_s, _th, _tz)] = _fuse(_s, _th, _tz)
for _sd, _os, _o in [(1, False, True), (2, True, True), (4, False, True)]:
    CASES["search_for_triangulation[%d,%d,%d]" % (_sd, _os, _o)] = _triang(_sd, _os, _o)

STEREO_CASE = dict(w=640, h=480, nf=800, disp=5, seed=6, mbf=386.1448, mb=0.5371)
