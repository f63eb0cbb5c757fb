"""GPU parity of Frame::ComputeStereoMatches on the device (orbx_stereo_match) against the CPU oracle: mvuRight and
mvDepth bit-exact, on a synthetic rectified pair (right = left scene shifted by a disparity) at the KITTI shape."""
import numpy as np
import pytest

from orb_slam_b200.synth import synth_frame
from tests.oracle_lib import OracleExtractor, oracle_stereo_matches
from tests.scenes import KITTI_BF, KITTI_K

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h,nf,disp", [(1241, 376, 2000, 23), (752, 480, 1000, 9), (1241, 376, 2000, 0)])
def test_stereo_matches_equal_oracle(w, h, nf, disp):
    from orb_slam_b200 import ORBextractor
    left = synth_frame(w, h, 77)
    right = synth_frame(w, h, 77, shift=(-disp, 0))
    rng = np.random.default_rng(3)
    right = np.clip(right.astype(np.int16) + rng.integers(-3, 4, right.shape), 0, 255).astype(np.uint8)   # sensor noise
    mbf = float(KITTI_BF); mb = float(np.float32(KITTI_BF) / KITTI_K[0, 0])
    gl, gr = ORBextractor(nf, 1.2, 8, 20, 7), ORBextractor(nf, 1.2, 8, 20, 7)
    kl, dl = gl(left); kr, dr = gr(right)
    ol, orr = OracleExtractor(nf, 1.2, 8, 20, 7), OracleExtractor(nf, 1.2, 8, 20, 7)
    okl, odl = ol(left); okr, odr = orr(right)
    assert kl.tobytes() == okl.tobytes() and kr.tobytes() == okr.tobytes()
    # same keypoints and (oracle) descriptors on both sides: the comparison isolates the stereo matcher
    u, z, n = gl.stereo_match(gr, okl, odl, okr, odr, mbf, mb)
    ou, oz, on = oracle_stereo_matches(ol, orr, okl, odl, okr, odr, mbf, mb)
    assert n == on and on > 0.3 * len(okl), (n, on, len(okl))
    assert u.tobytes() == ou.tobytes()
    assert z.tobytes() == oz.tobytes()
    good = ou >= 0
    if disp:                                           # the recovered disparity is the one the scene was built with
        assert np.median(okl["x"][good] - ou[good]) == pytest.approx(disp, abs=0.6)


def test_stereo_edge_cases():
    from orb_slam_b200 import ORBextractor
    img = synth_frame(640, 480, 5)
    a, b = ORBextractor(500, 1.2, 8, 20, 7), ORBextractor(500, 1.2, 8, 20, 7)
    ka, da = a(img); kb, db = b(np.full_like(img, 90))            # textureless right image: no right keypoints
    assert len(kb) == 0
    u, z, n = a.stereo_match(b, ka, da, kb, db.reshape(-1, 32), 40.0, 0.08)
    assert n == 0 and (u == -1).all() and (z == -1).all()
