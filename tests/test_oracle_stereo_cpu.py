"""CPU: the stereo-matching oracle (oracle/orb_oracle_stereo.cc, restating src/Frame.cc:630-872) behaves like stereo
matching: on a synthetic rectified pair built with a known disparity it recovers that disparity to sub-pixel accuracy,
depth = mbf / disparity, unmatched entries are -1, and the median cut removes only the SAD tail."""
import numpy as np

from orb_slam_b200.synth import synth_frame
from tests.oracle_lib import OracleExtractor, oracle_stereo_matches


def test_known_disparity_is_recovered():
    w, h, nf, disp = 752, 480, 1000, 14
    left, right = synth_frame(w, h, 3), synth_frame(w, h, 3, shift=(-disp, 0))
    exl, exr = OracleExtractor(nf, 1.2, 8, 20, 7), OracleExtractor(nf, 1.2, 8, 20, 7)
    kl, dl = exl(left); kr, dr = exr(right)
    mbf, mb = 47.9, 0.11
    u, z, kept = oracle_stereo_matches(exl, exr, kl, dl, kr, dr, mbf, mb)
    good = u >= 0
    assert kept == good.sum() and kept > 0.5 * len(kl)
    d = kl["x"][good] - u[good]
    assert abs(np.median(d) - disp) < 0.25 and np.percentile(np.abs(d - disp), 90) < 1.0
    assert np.allclose(z[good], np.float32(mbf) / d.astype(np.float32), rtol=1e-6)
    assert (z[~good] == -1).all() and (u[~good] == -1).all()
    # disparity gate: nothing outside [0, mbf/mb)
    assert (d >= 0).all() and (d < mbf / mb).all()


def test_no_right_keypoints():
    img = synth_frame(640, 480, 5)
    exl, exr = OracleExtractor(500, 1.2, 8, 20, 7), OracleExtractor(500, 1.2, 8, 20, 7)
    kl, dl = exl(img); kr, dr = exr(np.full_like(img, 90))
    u, z, kept = oracle_stereo_matches(exl, exr, kl, dl, kr, dr.reshape(-1, 32), 40.0, 0.08)
    assert kept == 0 and (u == -1).all() and (z == -1).all()
