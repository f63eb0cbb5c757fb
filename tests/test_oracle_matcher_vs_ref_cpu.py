"""Pins the matcher oracle (oracle/orb_oracle_matcher.cc, orb_oracle_bow.cc, orb_oracle_vocab.cc) to the REFERENCE'S OWN
object code: src/ORBmatcher.cc, Frame.cc, KeyFrame.cc, MapPoint.cc compiled unmodified into oracle/_ref/liborbmatcher_ref.so
(`make -C oracle ref`).  Each scene runs through both; results must be identical."""
import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import (OracleExtractor, OracleMatcher, RefMatcher, have_matcher_ref, oracle_distinctive_descriptors,
                              ref_distinctive_descriptors)

pytestmark = pytest.mark.skipif(not have_matcher_ref(), reason="oracle/_ref/liborbmatcher_ref.so not built (needs /root/reference)")


@pytest.fixture(scope="module")
def pair():
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 752, 480, 1000, 30, (7, -4))
    return scenes.make_frame(ka, da, 752, 480), scenes.make_frame(kb, db, 752, 480)


def test_grid_and_features_in_area(pair):
    o, r = OracleMatcher(), RefMatcher()
    for F in pair:
        a, b = o.grid_build(F), r.grid_build(F)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    rng = np.random.default_rng(3)
    F2 = pair[1]
    for t in range(120):
        x, y = float(rng.uniform(-30, 790)), float(rng.uniform(-30, 510))
        rad = float(rng.choice([5, 15, 40, 100, 1000]))
        lv = [(-1, -1), (0, 0), (1, 3), (2, -1), (0, 7)][t % 5]
        assert np.array_equal(o.features_in_area(F2, x, y, rad, *lv), r.features_in_area(F2, x, y, rad, *lv)), (x, y, rad, lv)


@pytest.mark.parametrize("ratio,ori,win", [(0.9, True, 100), (0.9, False, 100), (0.6, True, 30), (0.75, True, 10)])
def test_search_for_initialization(pair, ratio, ori, win):
    F1, F2 = pair
    prev_o = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
    prev_r = prev_o.copy()
    n_o, m_o = OracleMatcher(ratio, ori).SearchForInitialization(F1, F2, prev_o, win)
    n_r, m_r = RefMatcher(ratio, ori).SearchForInitialization(F1, F2, prev_r, win)
    assert n_o == n_r and (win < 30 or n_r > 20)
    assert np.array_equal(m_o, m_r) and np.array_equal(prev_o, prev_r)


@pytest.mark.parametrize("stereo,th,mono", [(False, 15.0, True), (True, 7.0, False), (True, 14.0, False), (False, 30.0, True)])
@pytest.mark.parametrize("tz", [0.05, -0.9, 0.9])
def test_search_by_projection_frame(stereo, th, mono, tz):
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 31, (3, 1))
    last = scenes.make_frame(ka, da, 1241, 376, stereo=stereo)
    Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [0.02, -0.01, tz]
    cur = scenes.make_frame(kb, db, 1241, 376, stereo=stereo, Tcw=Tcw, seed=1)
    pts = scenes.last_frame_points(last, seed=2)
    cur_obs = np.random.default_rng(4).choice([-1, -1, -1, 0, 2], cur.N).astype(np.int32)
    n_o, a_o = OracleMatcher(0.9, True).SearchByProjectionFrame(cur, last, pts, cur_obs, th, mono)
    n_r, a_r = RefMatcher(0.9, True).SearchByProjectionFrame(cur, last, pts, cur_obs, th, mono)
    assert n_o == n_r
    # the reference leaves no trace of "assigned, then set to NULL by the rotation check" on a slot that was NULL before
    a_o = np.where((cur_obs < 0) & (a_o == -1), -2, a_o)
    assert np.array_equal(a_o, a_r)


@pytest.mark.parametrize("th", [1.0, 3.0, 5.0])
@pytest.mark.parametrize("ratio", [0.8, 0.6])
def test_search_by_projection_points(th, ratio):
    (ka, da), _ = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 32, (0, 0))
    F = scenes.make_frame(ka, da, 1241, 376, stereo=True)
    pts = scenes.local_map_points(F, seed=5)
    obs = np.random.default_rng(6).choice([-1, -1, 0, 3], F.N).astype(np.int32)
    n_o, a_o = OracleMatcher(ratio, True).SearchByProjectionPoints(F, pts, obs, th)
    n_r, a_r = RefMatcher(ratio, True).SearchByProjectionPoints(F, pts, obs, th)
    assert n_o == n_r and n_r > 50
    assert np.array_equal(a_o, a_r)


BOW_CASES = [dict(seed=1), dict(seed=2, n1=2000, n2=2000, n_nodes=100), dict(seed=3, n1=300, n2=1800, n_nodes=7, ties=200),
             dict(seed=4, n1=64, n2=5, n_nodes=1, flips=0, p_common=1.0), dict(seed=7, n1=900, n2=900, p_valid=0.0),
             dict(seed=8, n1=1200, n2=1100, flips=60, n_nodes=16, ties=50)]


@pytest.mark.parametrize("case", BOW_CASES, ids=lambda c: "s%d" % c["seed"])
@pytest.mark.parametrize("nnratio,ori", [(0.7, True), (0.9, False)])
def test_search_by_bow(case, nnratio, ori):
    s = scenes.bow_scene(**case)
    o, r = OracleMatcher(nnratio, ori), RefMatcher(nnratio, ori)
    a = o.SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["fv2"])
    b = r.SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["fv2"])
    assert a[0] == b[0] and np.array_equal(a[1], b[1])
    a = o.SearchByBoWKeyFrames(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["valid2"], s["fv2"])
    b = r.SearchByBoWKeyFrames(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["valid2"], s["fv2"])
    assert a[0] == b[0] and np.array_equal(a[1], b[1])


def test_distinctive_descriptors():
    D, start = scenes.observation_lists(80, seed=4)
    chosen, has, order = ref_distinctive_descriptors(D, start)
    # feed the restatement each point's list in the order the reference's std::map<KeyFrame*, size_t> iterated it
    perm = np.concatenate([start[p] + order[start[p]:start[p + 1]] for p in range(len(start) - 1)]) if len(D) else np.zeros(0, int)
    Dp = D[perm]
    best = oracle_distinctive_descriptors(Dp, start)
    for p in range(len(start) - 1):
        assert bool(has[p]) == (best[p] >= 0)
        if best[p] >= 0:
            assert np.array_equal(Dp[start[p] + best[p]], chosen[p]), p


@pytest.mark.parametrize("w,h,nf,disp,seed", [(752, 480, 1000, 14, 3), (1241, 376, 2000, 23, 4), (640, 480, 800, 5, 6)])
def test_compute_stereo_matches(w, h, nf, disp, seed):
    """Frame::ComputeStereoMatches: the reference's own Frame.cc on reference-extracted key points and pyramids against the
    restatement (oracle/orb_oracle_stereo.cc) on the same key points and pyramid levels."""
    from orb_slam_b200.synth import synth_frame
    from tests.oracle_lib import oracle_stereo_matches_on_levels, ref_stereo
    left, right = synth_frame(w, h, seed), synth_frame(w, h, seed, shift=(-disp, 0))
    mbf, mb = 386.1448, 0.5371
    kL, dL, kR, dR, lvL, lvR, u_r, z_r = ref_stereo(left, right, nf, 1.2, 8, 20, 7, mbf, mb)
    u_o, z_o, kept = oracle_stereo_matches_on_levels(kL, dL, kR, dR, lvL, lvR, scenes.scale_factors(8, 1.2), mbf, mb)
    assert (u_r >= 0).sum() > 0.4 * len(kL) and kept == (u_r >= 0).sum()
    assert np.array_equal(u_o, u_r) and np.array_equal(z_o, z_r)


def keyframe_camera_center(Tcw):
    """KeyFrame::SetPose (src/KeyFrame.cc:70-85): Ow = -Rwc*tcw with Rwc = Rcw.t() materialised -> OpenCV's small-matrix gemm
    branch: float products and sums left to right, times alpha = -1."""
    T = np.asarray(Tcw, np.float32)
    Rwc, t = T[:3, :3].T.copy(), T[:3, 3]
    acc = Rwc[:, 0] * t[0]
    acc = (acc + Rwc[:, 1] * t[1]).astype(np.float32)
    acc = (acc + Rwc[:, 2] * t[2]).astype(np.float32)
    return (-acc).astype(np.float32)


@pytest.mark.parametrize("th,orbdist,tz", [(10.0, 100, 0.3), (3.0, 64, -0.4), (10.0, 100, 4.0)])
def test_search_by_projection_keyframe(th, orbdist, tz):
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 33, (4, -2))
    kf = scenes.make_frame(ka, da, 1241, 376)
    Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [0.03, 0.01, tz]
    cur = scenes.make_frame(kb, db, 1241, 376, Tcw=Tcw, seed=3)
    pts, mm, already = scenes.keyframe_points(kf, seed=11)
    has = (np.random.default_rng(12).random(cur.N) < 0.3).astype(np.uint8)
    args = (cur, pts, mm, already, kf.mvKeysUn["angle"], has, th, orbdist)
    n_o, a_o = OracleMatcher(0.9, True).SearchByProjectionKeyFrame(*args)
    n_r, a_r = RefMatcher(0.9, True).SearchByProjectionKeyFrame(*args)
    assert n_o == n_r and (tz > 1 or n_r > 20)
    a_o = np.where((has == 0) & (a_o == -1), -2, a_o)       # "assigned, then NULLed" leaves no trace in the reference
    assert np.array_equal(a_o, a_r)


@pytest.mark.parametrize("stereo,th,tz", [(True, 3.0, 0.2), (False, 3.0, -0.3), (True, 6.0, 1.5)])
def test_fuse_select(stereo, th, tz):
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 35, (-3, 2))
    src = scenes.make_frame(ka, da, 1241, 376)
    Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [-0.05, 0.02, tz]
    kf = scenes.make_frame(kb, db, 1241, 376, stereo=stereo, Tcw=Tcw, seed=7)
    if not stereo:
        kf.mvuRight = np.full(kf.N, -1, np.float32)
    pts, mm, normals, in_kf = scenes.fuse_points(src, seed=21)
    sf = kf.mvScaleFactors
    inv_sigma2 = (1.0 / (sf * sf)).astype(np.float32)
    cc = keyframe_camera_center(Tcw)
    o = OracleMatcher(0.6, True).FuseSelect(kf, inv_sigma2, pts, in_kf, normals, mm, th, camera_center=cc)
    r = RefMatcher(0.6, True).FuseSelect(kf, inv_sigma2, pts, in_kf, normals, mm, th, camera_center=cc)
    assert (r >= 0).sum() > 20
    assert np.array_equal(o, r)


@pytest.mark.parametrize("seed,only_stereo,ori", [(1, False, True), (2, True, True), (3, False, False), (4, False, True)])
def test_search_for_triangulation(seed, only_stereo, ori):
    kw = dict(t=(0.02, 0.01, 0.6), stereo_frac=0.0) if seed == 4 else {}
    s = scenes.triangulation_scene(seed, **kw)
    args = (s["kf1"], s["kf2"], s["Cw1"], s["R2w"], s["t2w"], s["K2"], s["sf"], s["sig2"], s["F12"], only_stereo)
    n_o, m_o = OracleMatcher(0.6, ori).SearchForTriangulation(*args)
    n_r, m_r = RefMatcher(0.6, ori).SearchForTriangulation(*args)
    assert n_o == n_r and n_r > 50
    assert np.array_equal(m_o, m_r)


@pytest.mark.parametrize("scale,tz,th", [(1.0, 0.2, 10), (1.15, -0.3, 10), (0.9, 1.0, 4)])
def test_sim3_projection_and_fuse(scale, tz, th):
    """Loop closing's SearchByProjection(pKF, Scw, ...) (:290-403) and Fuse(pKF, Scw, ...) (:1044-1168)"""
    from tests.oracle_lib import ref_decompose_scw, ref_sim3_projection
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 37, (2, 3))
    src = scenes.make_frame(ka, da, 1241, 376)
    kf = scenes.make_frame(kb, db, 1241, 376)
    pts, mm, normals, in_kf = scenes.fuse_points(src, seed=31)
    pts["valid"] = 1
    excluded = (pts["bad"].astype(bool) | in_kf.astype(bool)).astype(np.uint8)
    a = np.float32(0.01)
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]], np.float32)
    Scw = np.eye(4, dtype=np.float32)
    Scw[:3, :3] = np.float32(scale) * R; Scw[:3, 3] = np.array([0.03, -0.02, tz], np.float32)
    Rcw, tcw, Ow = ref_decompose_scw(Scw)                      # what the reference computes at :299-303
    matched_in = (np.random.default_rng(32).random(kf.N) < 0.3).astype(np.uint8)
    n_o, m_o = OracleMatcher(0.75, True).SearchByProjectionSim3(kf, Rcw, tcw, Ow, pts, excluded, normals, mm, matched_in, th)
    n_r, m_r = ref_sim3_projection(kf, Scw, pts, excluded, normals, mm, matched_in, th)
    assert n_o == n_r and n_r > 20
    assert np.array_equal(m_o, m_r)
    b_o = OracleMatcher(0.75, True).FuseSelectSim3(kf, Rcw, tcw, Ow, pts, excluded, normals, mm, float(th))
    n_f, b_r = ref_sim3_projection(kf, Scw, pts, excluded, normals, mm, None, float(th), fuse=True)
    assert np.array_equal(b_o, b_r) and n_f == (b_r >= 0).sum() > 20


@pytest.mark.parametrize("seed,s12,th", [(0, 1.08, 7.5), (1, 0.93, 7.5), (2, 1.0, 3.0)])
def test_search_by_sim3(seed, s12, th):
    from tests.oracle_lib import ref_search_by_sim3
    sc = scenes.sim3_scene(seed=seed, s12=s12)
    n_o, m_o = OracleMatcher(0.75, True).SearchBySim3(sc["kf1"], sc["kf2"], sc["pts1"], sc["pts2"], sc["skip1"], sc["skip2"], sc["mm1"],
                                                      sc["mm2"], sc["sR12"], sc["t12"], sc["sR21"], sc["t21"], th)
    n_r, m_r = ref_search_by_sim3(sc["kf1"], sc["kf2"], sc["pts1"], sc["pts2"], sc["skip1"], sc["skip2"], sc["mm1"], sc["mm2"],
                                  sc["s12"], sc["R12"], sc["t12"], th)
    assert n_o == n_r and n_r > 300
    assert np.array_equal(m_o, m_r)
