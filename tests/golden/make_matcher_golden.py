"""Generates tests/golden/ref_matcher_hashes.json and ref_stereo_640x480.npz by running tests/matcher_cases.py through the
REFERENCE's own object code (oracle/_ref/liborbmatcher_ref.so = /root/reference/src/ORBmatcher.cc, Frame.cc, KeyFrame.cc,
MapPoint.cc compiled where they lie by `make -C oracle ref`).  Run in the build container: python tests/golden/make_matcher_golden.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from orb_slam_b200.synth import synth_frame                      # noqa: E402
from tests.matcher_cases import CASES, STEREO_CASE, digest       # noqa: E402
from tests.oracle_lib import RefMatcher, ref_stereo              # noqa: E402

out = {name: digest(fn(RefMatcher)) for name, fn in CASES.items()}
c = STEREO_CASE
left, right = synth_frame(c["w"], c["h"], c["seed"]), synth_frame(c["w"], c["h"], c["seed"], shift=(-c["disp"], 0))
kL, dL, kR, dR, lvL, lvR, u, z = ref_stereo(left, right, c["nf"], 1.2, 8, 20, 7, c["mbf"], c["mb"])
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_stereo_640x480.npz"), kL=kL, dL=dL, kR=kR, dR=dR, u=u, z=z)
out["compute_stereo_matches[640x480]"] = digest([u, z])
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "ref_matcher_hashes.json"), "w"), indent=1, sort_keys=True)
print(len(out), "cases written")
