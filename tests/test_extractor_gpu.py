"""GPU parity: the CUDA extractor (through the C ABI) against the CPU oracle on
the same inputs — pyramid, blur, FAST candidates, quadtree survivors, angles,
descriptors.  Bit-exact for everything integer; angles within 1e-3 deg;
descriptors bit-exact on >= 99.9 % of keypoints (BASELINE.json north_star)."""
import numpy as np
import pytest

from tests.oracle_lib import OracleExtractor
from orb_slam_b200.synth import synth_frame

pytestmark = pytest.mark.gpu

CASES = [
    # (w, h, nfeatures, kind, index)
    (640, 480, 1000, "synth", 0),
    (1241, 376, 2000, "synth", 1),
    (752, 480, 1000, "synth", 2),
    (640, 480, 1000, "noise", 3),
    (1241, 376, 2000, "noise", 4),
    (321, 243, 500, "synth", 5),
    (200, 150, 300, "synth", 6),
    (1920, 1080, 4000, "synth", 7),
]


def make_image(w, h, kind, idx):
    if kind == "synth":
        return synth_frame(w, h, idx)
    if kind == "noise":
        return np.random.default_rng(idx).integers(0, 256, (h, w), dtype=np.uint8)
    if kind == "flat":
        return np.full((h, w), 77, np.uint8)
    raise ValueError(kind)


def compare(gk, gd, ok, od):
    assert len(gk) == len(ok), (len(gk), len(ok))
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(gk[f], ok[f]), f
    assert np.max(np.abs(gk["angle"] - ok["angle"]), initial=0) <= 1e-3
    same = (gd == od).all(axis=1)
    assert same.mean() >= 0.999 if len(same) else True, f"descriptor rows differing: {(~same).sum()} of {len(same)}"
    return int((~same).sum()), int((gk["angle"] != ok["angle"]).sum())


@pytest.mark.parametrize("w,h,nf,kind,idx", CASES)
def test_extract_matches_oracle(w, h, nf, kind, idx):
    from orb_slam_b200 import ORBextractor
    img = make_image(w, h, kind, idx)
    gpu = ORBextractor(nf, 1.2, 8, 20, 7)
    cpu = OracleExtractor(nf, 1.2, 8, 20, 7)
    gk, gd = gpu(img)
    ok, od = cpu(img)
    for l in range(8):
        assert np.array_equal(gpu.get_level(l, border=19), cpu.get_level(l, border=19)), f"pyramid level {l}"
        ob = cpu.get_level(l, blurred=True)
        if ob is not None:
            assert np.array_equal(gpu.get_level(l, blurred=True), ob), f"blurred level {l}"
        assert np.array_equal(gpu.get_candidates(l), cpu.get_candidates(l)), f"candidates level {l}"
    compare(gk, gd, ok, od)
    gpu.close()


def test_tables_match_oracle():
    from orb_slam_b200 import ORBextractor
    for nf, sf, nl in [(1000, 1.2, 8), (2000, 1.2, 8), (1200, 1.2, 8), (4000, 1.2, 8), (500, 1.5, 5), (300, 1.1, 12)]:
        gpu = ORBextractor(nf, sf, nl, 20, 7)
        cpu = OracleExtractor(nf, sf, nl, 20, 7)
        t = cpu.tables()
        assert np.array_equal(gpu.GetScaleFactors(), t[0])
        assert np.array_equal(gpu.GetInverseScaleFactors(), t[1])
        assert np.array_equal(gpu.GetScaleSigmaSquares(), t[2])
        assert np.array_equal(gpu.GetInverseScaleSigmaSquares(), t[3])
        assert np.array_equal(gpu.mnFeaturesPerLevel, t[4])
        gpu.close()


@pytest.mark.parametrize("w,h,nf,kind,idx", CASES)
def test_batch_tiles_match_oracle(w, h, nf, kind, idx):
    """Calls of one or two frames take K1's 16-row tiles (TH_LATENCY), larger batches the 32-row tiles: the same shapes
    again as a three-frame batch, planes of the last frame included."""
    from orb_slam_b200 import ORBextractor
    img = make_image(w, h, kind, idx)
    imgs = np.stack([img[::-1].copy(), np.ascontiguousarray(img[:, ::-1]), img])
    gpu = ORBextractor(nf, 1.2, 8, 20, 7)
    cpu = OracleExtractor(nf, 1.2, 8, 20, 7)
    res = gpu.extract_batch(imgs)
    for i in (0, 1):
        compare(*res[i], *cpu(imgs[i]))
    ok, od = cpu(img)                                    # the oracle's planes now belong to imgs[2]
    compare(*res[2], ok, od)
    for l in range(8):
        assert np.array_equal(gpu.get_level(l, frame=2, border=19), cpu.get_level(l, border=19)), f"pyramid level {l}"
        ob = cpu.get_level(l, blurred=True)
        if ob is not None:
            assert np.array_equal(gpu.get_level(l, frame=2, blurred=True), ob), f"blurred level {l}"
    gpu.close()


def test_batch_equals_single_and_is_repeatable():
    from orb_slam_b200 import ORBextractor
    imgs = np.stack([synth_frame(752, 480, 10 + i) for i in range(5)])
    gpu = ORBextractor(1000, 1.2, 8, 20, 7)
    cpu = OracleExtractor(1000, 1.2, 8, 20, 7)
    res = gpu.extract_batch(imgs)
    res2 = gpu.extract_batch(imgs)
    for i in range(5):
        ok, od = cpu(imgs[i])
        compare(res[i][0], res[i][1], ok, od)
        assert res[i][0].tobytes() == res2[i][0].tobytes() and np.array_equal(res[i][1], res2[i][1])
    gpu.close()


def test_edge_cases():
    from orb_slam_b200 import ORBextractor
    gpu = ORBextractor(1000, 1.2, 8, 20, 7)
    cpu = OracleExtractor(1000, 1.2, 8, 20, 7)
    # flat image: no keypoints at all (descriptors released, reference :1274-1276)
    flat = make_image(640, 480, "flat", 0)
    gk, gd = gpu(flat)
    assert len(gk) == 0 and gd.shape == (0, 32)
    # non-contiguous rows (step > width)
    big = synth_frame(800, 480, 3)
    view = big[:, 24:24 + 640]
    gk, gd = gpu(view)
    ok, od = cpu(np.ascontiguousarray(view))
    compare(gk, gd, ok, od)
    # empty image: outputs untouched
    assert gpu(np.zeros((0, 0), np.uint8)) == (None, None)
    # other thresholds / scale factors
    # (one and two levels: the few-frames schedule without / with its level-0 fork)
    for nf, sf, nl, ini, mn in [(1500, 1.5, 5, 30, 10), (800, 1.1, 10, 12, 5), (500, 1.2, 1, 20, 7), (600, 1.3, 2, 20, 7)]:
        g2 = ORBextractor(nf, sf, nl, ini, mn)
        c2 = OracleExtractor(nf, sf, nl, ini, mn)
        img = synth_frame(640, 480, 20)
        a, b = g2(img); c, d = c2(img)
        compare(a, b, c, d)
        g2.close()
    gpu.close()


def test_quadtree_general_passes_and_radix_sort():
    """quadtree_kernel keeps two forms of its list passes and of its sort: the register-scan passes and the bucket + rank sort
    (taken for the list lengths and bucket sizes of ordinary frames) and the general passes with the radix sort (long lists,
    crowded buckets).  ORBX_QT_GENERAL=1 (read once per process) sends every list through the general form: the same parity
    tests again in a child process: single-frame blocks (512 threads), batch blocks (128 threads) and the cluster pre-sort of
    large levels (qt_presort_kernel, the 4K frame)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, ORBX_QT_GENERAL="1")
    r = subprocess.run([sys.executable, "-m", "pytest", "tests/test_extractor_gpu.py", "tests/test_configs_gpu.py", "-q", "-x", "-m", "gpu",
                        "-p", "no:cacheprovider", "-k",
                        "test_extract_matches_oracle or test_batch_tiles_match_oracle or test_batch_across_chunks or test_other_extractor_settings "
                        "or test_4k_stress_frame"],
                       cwd=root, env=env, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and "failed" not in r.stdout
