"""CPU: the oracle's OpenCV-primitive restatements (oracle/cvprim) against the committed cv2 4.13.0
golden vectors (tests/golden/make_cv_golden.py).  Bit-exact."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

from tests.oracle_lib import oracle
from orb_slam_b200.synth import synth_frame

G = os.path.join(os.path.dirname(__file__), "golden")
SMALL = np.load(os.path.join(G, "cv_golden_small.npz"))
HASHES = json.load(open(os.path.join(G, "cv_golden_hashes.json")))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def p(a):
    return C.c_void_p(a.ctypes.data)


def resize(img, w, h):
    out = np.empty((h, w), np.uint8)
    oracle().cvprim_resize_linear_u8(p(img), img.shape[1], img.shape[0], img.strides[0], p(out), w, h, out.strides[0])
    return out


def border(img):
    out = np.empty((img.shape[0] + 38, img.shape[1] + 38), np.uint8)
    oracle().cvprim_border_reflect101_u8(p(img), img.shape[1], img.shape[0], img.strides[0], p(out), 19, out.strides[0])
    return out


def blur(img, inplace=False):
    src = img.copy()
    out = src if inplace else np.empty_like(img)
    oracle().cvprim_gaussian7x7_s2_u8(p(src), img.shape[1], img.shape[0], src.strides[0], p(out), out.strides[0])
    return out


def fast(img, th):
    cap = img.size
    out = np.zeros((cap, 3), np.int32)
    n = oracle().cvprim_fast9_16(p(img), img.shape[1], img.shape[0], img.strides[0], th, 1, p(out), cap)
    return out[:n]


def atan2(y, x):
    out = np.empty(len(y), np.float32)
    oracle().cvprim_fast_atan2(p(y), p(x), p(out), len(y))
    return out


def test_small_vectors():
    prev = SMALL["resize_in"]
    for l in range(1, 4):
        ref = SMALL[f"resize_l{l}"]
        prev = resize(prev, ref.shape[1], ref.shape[0])
        assert np.array_equal(prev, ref)
    assert np.array_equal(border(SMALL["border_in"]), SMALL["border_out"])
    assert np.array_equal(blur(SMALL["blur_in"]), SMALL["blur_out"])
    assert np.array_equal(blur(SMALL["blur_in"], inplace=True), SMALL["blur_out"])
    assert np.array_equal(blur(SMALL["blur_tiny_in"]), SMALL["blur_tiny_out"])
    for i in range(6):
        img = np.ascontiguousarray(SMALL[f"fast_in{i}"])
        assert np.array_equal(fast(img, 20), SMALL[f"fast_t20_{i}"])
        assert np.array_equal(fast(img, 7), SMALL[f"fast_t7_{i}"])
    got = atan2(np.ascontiguousarray(SMALL["atan_y"]), np.ascontiguousarray(SMALL["atan_x"]))
    assert got.tobytes() == SMALL["atan_out"].tobytes()


@pytest.mark.parametrize("case", HASHES["cases"], ids=lambda c: f"{c['kind']}{c['w']}x{c['h']}")
def test_config_size_hashes(case):
    w, h, seed = case["w"], case["h"], case["seed"]
    img = synth_frame(w, h, seed) if case["kind"] == "synth" else \
        np.random.default_rng(seed).integers(0, 256, (h, w), dtype=np.uint8)
    assert sha(img) == case["input"], "input generator changed"
    from tests.golden.make_ref_golden import sha as _  # noqa: F401  (keeps the generator importable)
    s = np.float32(1.0)
    prev = img
    for l in range(8):
        inv = np.float32(1.0) / s
        lw, lh = int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))
        if l:
            prev = resize(prev, lw, lh)
        assert sha(prev) == case["levels"][l], f"resize level {l}"
        assert sha(border(prev)) == case["bordered"][l], f"border level {l}"
        assert sha(blur(prev)) == case["blur"][l], f"blur level {l}"
        if l < 3:
            assert sha(fast(prev, 20)) == case["fast20"][l], f"FAST th20 level {l}"
            assert sha(fast(prev, 7)) == case["fast7"][l], f"FAST th7 level {l}"
        s = np.float32(np.float64(s) * np.float64(np.float32(1.2)))


def test_atan2_hash():
    m = np.random.default_rng(77).integers(-749 * 15 * 255, 749 * 15 * 255, (2, 200000)).astype(np.float32)
    assert sha(atan2(np.ascontiguousarray(m[0]), np.ascontiguousarray(m[1]))) == HASHES["atan2_200k_seed77"]


def test_mat_algebra_golden():
    """The CV_32F matrix algebra of oracle/ref_build/cvstub (what the reference's ORBmatcher.cc / Frame.cc / KeyFrame.cc are
    compiled against in oracle/_ref) reproduces cv2 4.13 bit for bit: A*x+c, A*x, -A.t()*x, (-A)*x, norm(x)."""
    import ctypes as C
    import pytest
    from tests.oracle_lib import MATCHER_REF_SO, have_matcher_ref
    if not have_matcher_ref():
        pytest.skip("oracle/_ref/liborbmatcher_ref.so not built (needs /root/reference)")
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv_algebra.npz"))
    L = C.CDLL(MATCHER_REF_SO)
    L.orbref_algebra_probe.argtypes = [C.c_void_p] * 4; L.orbref_algebra_probe.restype = None
    A, x, c, gold = z["A"], z["x"], z["c"], z["out"]
    got = np.zeros_like(gold)
    for i in range(len(A)):
        a, xi, ci = np.ascontiguousarray(A[i]), np.ascontiguousarray(x[i]), np.ascontiguousarray(c[i])
        L.orbref_algebra_probe(a.ctypes.data, xi.ctypes.data, ci.ctypes.data, got[i].ctypes.data)
    for name, sl in (("A*x+c", slice(0, 3)), ("A*x", slice(3, 6)), ("-A.t()*x", slice(6, 9)), ("(-A)*x", slice(9, 12)), ("norm", slice(12, 13))):
        assert np.array_equal(got[:, sl].view(np.uint32), gold[:, sl].view(np.uint32)), name
