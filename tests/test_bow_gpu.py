"""GPU parity of orbm_search_by_bow (ORBmatcher::SearchByBoW, src/ORBmatcher.cc:159-288 and :590-723) against the CPU
oracle: the match arrays and the returned counts are identical."""
import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import OracleMatcher

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from orb_slam_b200.matcher import MatcherContext
    c = MatcherContext(0)
    yield c
    c.close()


CASES = [dict(seed=1), dict(seed=2, n1=2000, n2=2000, n_nodes=100), dict(seed=3, n1=300, n2=1800, n_nodes=7, ties=200),
         dict(seed=4, n1=64, n2=5, n_nodes=1, flips=0, p_common=1.0), dict(seed=5, n1=0, n2=40), dict(seed=6, n1=40, n2=0),
         dict(seed=7, n1=900, n2=900, p_valid=0.0), dict(seed=8, n1=1200, n2=1100, flips=60, n_nodes=16, ties=50)]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "s%d" % c["seed"])
@pytest.mark.parametrize("nnratio,ori", [(0.7, True), (0.9, False), (0.6, True)])
def test_bow_keyframe_frame(ctx, case, nnratio, ori):
    from orb_slam_b200.matcher import ORBmatcher
    s = scenes.bow_scene(**case)
    g = ORBmatcher(nnratio, ori, ctx).SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["fv2"])
    o = OracleMatcher(nnratio, ori).SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["fv2"])
    assert g[0] == o[0]
    np.testing.assert_array_equal(g[1], o[1])
    assert g[0] == int((g[1] >= 0).sum())
    if case["seed"] in (1, 2) and nnratio == 0.7:
        assert g[0] > 100           # the scene does produce matches


@pytest.mark.parametrize("case", CASES, ids=lambda c: "s%d" % c["seed"])
@pytest.mark.parametrize("nnratio,ori", [(0.8, True), (0.6, False)])
def test_bow_keyframe_keyframe(ctx, case, nnratio, ori):
    from orb_slam_b200.matcher import ORBmatcher
    s = scenes.bow_scene(**case)
    args = (s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["valid2"], s["fv2"])
    g = ORBmatcher(nnratio, ori, ctx).SearchByBoWKeyFrames(*args)
    o = OracleMatcher(nnratio, ori).SearchByBoWKeyFrames(*args)
    assert g[0] == o[0]
    np.testing.assert_array_equal(g[1], o[1])
    m = g[1][g[1] >= 0]
    assert len(np.unique(m)) == len(m)                 # vbMatched2: a map point of pKF2 is handed out once
    assert np.all(s["valid2"][m] == 1) and np.all(s["valid1"][np.nonzero(g[1] >= 0)[0]] == 1)


def test_bow_rejects_bad_arguments(ctx):
    from orb_slam_b200.matcher import FeatureVector, ORBmatcher
    from orb_slam_b200 import FrontendError
    s = scenes.bow_scene(seed=9, n1=50, n2=50)
    bad = FeatureVector({3: [0, 1, 99]})                 # feature index past n1
    with pytest.raises(FrontendError):
        ORBmatcher(0.7, True, ctx).SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], bad, s["desc2"], s["angle2"], s["fv2"])


@pytest.mark.parametrize("seed,only_stereo,ori", [(1, False, True), (2, True, True), (3, False, False), (4, False, True)])
def test_search_for_triangulation(ctx, seed, only_stereo, ori):
    """SearchForTriangulation (src/ORBmatcher.cc:725-891): the whole inner loop runs on the device"""
    from orb_slam_b200.matcher import ORBmatcher
    kw = dict(t=(0.02, 0.01, 0.6), stereo_frac=0.0) if seed == 4 else {}         # seed 4: monocular, epipole inside the image
    s = scenes.triangulation_scene(seed, **kw)
    args = (s["kf1"], s["kf2"], s["Cw1"], s["R2w"], s["t2w"], s["K2"], s["sf"], s["sig2"], s["F12"], only_stereo)
    n_g, m_g = ORBmatcher(0.6, ori, ctx).SearchForTriangulation(*args)
    n_o, m_o = OracleMatcher(0.6, ori).SearchForTriangulation(*args)
    assert n_g == n_o
    np.testing.assert_array_equal(m_g, m_o)
    assert n_g == int((m_g >= 0).sum()) and n_g > 50
    i = np.nonzero(m_g >= 0)[0]
    assert not np.any(s["kf1"].has_point[i]) and not np.any(s["kf2"].has_point[m_g[i]])
    if only_stereo:
        assert np.all(s["kf1"].uRight[i] >= 0) and np.all(s["kf2"].uRight[m_g[i]] >= 0)
    assert (m_g[i] == s["true12"][i]).mean() > 0.9                              # and they are the planted correspondences


def test_search_for_triangulation_empty(ctx):
    from orb_slam_b200.matcher import FeatureVector, KeyFrameView, ORBmatcher
    from orb_slam_b200._lib import KEYPOINT_DTYPE
    s = scenes.triangulation_scene(5, n=50, extra=10)
    empty = KeyFrameView(np.zeros(0, KEYPOINT_DTYPE), np.zeros((0, 32), np.uint8), np.zeros(0, np.float32), np.zeros(0, np.uint8), FeatureVector({}))
    n, m = ORBmatcher(0.6, True, ctx).SearchForTriangulation(empty, s["kf2"], s["Cw1"], s["R2w"], s["t2w"], s["K2"], s["sf"], s["sig2"], s["F12"])
    assert n == 0 and len(m) == 0
    n, m = ORBmatcher(0.6, True, ctx).SearchForTriangulation(s["kf1"], empty, s["Cw1"], s["R2w"], s["t2w"], s["K2"], s["sf"], s["sig2"], s["F12"])
    assert n == 0 and np.all(m == -1)
