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
