"""Generates golden outputs of the REFERENCE's own ORBextractor (oracle/_ref, built from
/root/reference/src/ORBextractor.cc by `make -C oracle ref`; canonical tie-break = bump allocator).
Run HERE only (needs /root/reference):

    python tests/golden/make_ref_golden.py

Outputs (committed): tests/golden/ref_extract_640x480_n1000.npz  - keypoints + descriptors, config 1
                     tests/golden/ref_extract_hashes.json         - sha256 of the outputs for the other shapes
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from orb_slam_b200.synth import synth_frame  # noqa: E402
from tests.oracle_lib import RefExtractor  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    ex = RefExtractor(1000, 1.2, 8, 20, 7)
    img = synth_frame(640, 480, 0)
    k, d = ex(img)
    lv = {f"level{l}_sha": sha(ex.get_level(l, border=19)) for l in range(8)}
    np.savez_compressed(os.path.join(HERE, "ref_extract_640x480_n1000.npz"), keypoints=k, descriptors=d,
                        input_sha=sha(img), **lv)
    cases = []
    for w, h, nf, sf, nl, ini, mn, kind, seed in [
        (1241, 376, 2000, 1.2, 8, 20, 7, "synth", 1), (752, 480, 1000, 1.2, 8, 20, 7, "synth", 2),
        (640, 480, 1000, 1.2, 8, 20, 7, "noise", 3), (1241, 376, 2000, 1.2, 8, 20, 7, "noise", 4),
        (1920, 1080, 4000, 1.2, 8, 20, 7, "synth", 7), (321, 243, 500, 1.2, 8, 20, 7, "synth", 5),
        (200, 150, 300, 1.2, 8, 20, 7, "synth", 6), (640, 480, 1500, 1.5, 5, 30, 10, "synth", 20),
        (640, 480, 800, 1.1, 10, 12, 5, "synth", 20), (3840, 2160, 4000, 1.2, 8, 20, 7, "synth", 9),
    ]:
        img = synth_frame(w, h, seed) if kind == "synth" else np.random.default_rng(seed).integers(0, 256, (h, w), dtype=np.uint8)
        e = RefExtractor(nf, sf, nl, ini, mn)
        k, d = e(img)
        cases.append({"w": w, "h": h, "nfeatures": nf, "scaleFactor": sf, "nlevels": nl, "iniThFAST": ini, "minThFAST": mn,
                      "kind": kind, "seed": seed, "n": int(len(k)), "keypoints_sha": sha(k), "descriptors_sha": sha(d),
                      "per_level": np.bincount(k["octave"], minlength=nl).tolist()})
        print(w, h, nf, kind, len(k))
    json.dump({"source": "reference src/ORBextractor.cc via oracle/_ref (canonical tie-break), cvprim pinned to OpenCV 4.13.0",
               "cases": cases}, open(os.path.join(HERE, "ref_extract_hashes.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
