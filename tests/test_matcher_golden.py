"""Golden hashes produced by the REFERENCE's own ORBmatcher.cc / Frame.cc object code (tests/golden/make_matcher_golden.py,
run where /root/reference exists) against (a) the CPU oracle - this is what pins the matcher oracle on a machine without
the reference tree - and (b) the CUDA product path through the C ABI (-m gpu)."""
import json
import os

import numpy as np
import pytest

from tests.matcher_cases import CASES, STEREO_CASE, digest
from tests.oracle_lib import OracleExtractor, OracleMatcher

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "ref_matcher_hashes.json")))


def _stereo_inputs():
    from orb_slam_b200.synth import synth_frame
    c = STEREO_CASE
    left, right = synth_frame(c["w"], c["h"], c["seed"]), synth_frame(c["w"], c["h"], c["seed"], shift=(-c["disp"], 0))
    g = np.load(os.path.join(HERE, "golden", "ref_stereo_640x480.npz"))
    return c, left, right, g


def test_every_case_has_a_golden():
    assert set(CASES) | {"compute_stereo_matches[640x480]"} == set(GOLD)


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_equals_reference_hash(name):
    assert digest(CASES[name](OracleMatcher)) == GOLD[name]


def test_oracle_stereo_equals_reference_hash():
    """Frame::ComputeStereoMatches restated, on the reference's own key points (committed) and the oracle's pyramid levels"""
    from tests import scenes
    from tests.oracle_lib import oracle_stereo_matches_on_levels
    c, left, right, g = _stereo_inputs()
    exL, exR = OracleExtractor(c["nf"], 1.2, 8, 20, 7), OracleExtractor(c["nf"], 1.2, 8, 20, 7)
    exL(left); exR(right)
    lvL = [exL.get_level(l) for l in range(8)]; lvR = [exR.get_level(l) for l in range(8)]
    u, z, kept = oracle_stereo_matches_on_levels(g["kL"], g["dL"], g["kR"], g["dR"], lvL, lvR, scenes.scale_factors(8, 1.2), c["mbf"], c["mb"])
    assert digest([u, z]) == GOLD["compute_stereo_matches[640x480]"] and kept == (g["u"] >= 0).sum() > 300


@pytest.fixture(scope="module")
def ctx():
    from orb_slam_b200.matcher import MatcherContext
    c = MatcherContext(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_equals_reference_hash(ctx, name):
    from orb_slam_b200.matcher import ORBmatcher
    assert digest(CASES[name](lambda ratio, ori: _Gpu(ORBmatcher(ratio, ori, ctx), ctx))) == GOLD[name]


class _Gpu:
    """ORBmatcher plus the two grid entry points of the context under the oracle's names"""

    def __init__(self, m, ctx):
        self._m, self._ctx = m, ctx

    def __getattr__(self, k):
        return getattr(self._ctx if k in ("grid_build", "features_in_area") else self._m, k)


@pytest.mark.gpu
def test_cuda_stereo_equals_reference_hash():
    from orb_slam_b200 import ORBextractor
    c, left, right, g = _stereo_inputs()
    exL, exR = ORBextractor(c["nf"], 1.2, 8, 20, 7, device=0), ORBextractor(c["nf"], 1.2, 8, 20, 7, device=0)
    try:
        exL(left); exR(right)                                   # the pyramids stay resident; the key points are the reference's
        u, z, n = exL.stereo_match(exR, g["kL"], g["dL"], g["kR"], g["dR"], c["mbf"], c["mb"])
        assert digest([np.asarray(u, np.float32), np.asarray(z, np.float32)]) == GOLD["compute_stereo_matches[640x480]"]
        assert n == (g["u"] >= 0).sum()
    finally:
        exL.close(); exR.close()
