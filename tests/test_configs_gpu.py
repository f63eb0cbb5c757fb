"""GPU parity at the BASELINE.json config shapes and on the edge cases of the host-buffer path: batches crossing the
pipeline chunk size, padded row pitch, pageable vs page-locked buffers, empty / flat / too-small images, 4K frames,
EuRoC-shape batches followed by SearchForInitialization on consecutive frames."""
import ctypes as C

import numpy as np
import pytest

from orb_slam_b200.synth import synth_frame
from tests import scenes
from tests.oracle_lib import OracleExtractor, OracleMatcher

pytestmark = pytest.mark.gpu


def same(gk, gd, ok, od):
    assert len(gk) == len(ok)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(gk[f], ok[f]), f
    assert np.max(np.abs(gk["angle"] - ok["angle"]), initial=0) <= 1e-3
    assert (gd == od).all(axis=1).mean() >= 0.999 if len(gd) else True


@pytest.mark.parametrize("n,pinned", [(70, False), (150, False), (260, True)])
def test_batch_across_chunks_equals_single_frame_calls(n, pinned):
    """Batches that cross the host-buffer pipeline's chunking: 70 pageable frames = 8 + 16 + 32 + a ragged chunk; 150
    pageable (32-frame chunks) and 260 page-locked frames (64-frame chunks) take the ramp-up / ramp-down schedule over
    the four compute streams and wrap around the six buffer sets.  Every frame must equal its own single-frame call."""
    import torch
    from orb_slam_b200 import KEYPOINT_DTYPE, ORBextractor
    from orb_slam_b200._lib import check
    w, h, nf = 321, 243, 300
    imgs = np.stack([synth_frame(w, h, 200 + (i % 9), shift=(i % 40, -(i % 40))) for i in range(n)])
    ex = ORBextractor(nf, 1.2, 8, 20, 7)
    if pinned:
        cap = ex.max_keypoints(w, h)
        pin = torch.from_numpy(imgs).pin_memory()
        kp = torch.empty((n, cap, 28), dtype=torch.uint8, pin_memory=True)
        dp = torch.empty((n, cap, 32), dtype=torch.uint8, pin_memory=True)
        cnt = np.zeros(n, np.int32)
        ptrs = (C.c_void_p * n)(*[pin.data_ptr() + i * w * h for i in range(n)])
        check(ex._L.orbx_extract_batch(ex.handle, ptrs, n, w, h, w, kp.data_ptr(), dp.data_ptr(), cap, cnt.ctypes.data),
              "orbx_extract_batch")
        batch = [(kp.numpy()[i, : cnt[i]].copy().view(KEYPOINT_DTYPE).reshape(-1), dp.numpy()[i, : cnt[i]].copy()) for i in range(n)]
    else:
        batch = ex.extract_batch(imgs)
    one = ORBextractor(nf, 1.2, 8, 20, 7)
    for i in sorted({0, 1, 7, 8, 15, 16, 31, 32, 47, 48, 63, 64, 65, 69, n // 2, n - 17, n - 9, n - 8, n - 1}):
        k, d = one(imgs[i])
        assert k.tobytes() == batch[i][0].tobytes() and np.array_equal(d, batch[i][1]), i
    cpu = OracleExtractor(nf, 1.2, 8, 20, 7)
    for i in (0, 64, n - 1):
        same(*batch[i], *cpu(imgs[i]))


def test_padded_pitch_and_pinned_buffers_give_identical_results():
    import torch
    from orb_slam_b200 import KEYPOINT_DTYPE, ORBextractor
    from orb_slam_b200._lib import check
    w, h, nf = 1241, 376, 2000
    img = synth_frame(w, h, 8)
    ex = ORBextractor(nf, 1.2, 8, 20, 7)
    k0, d0 = ex(img)
    padded = np.zeros((h, 1300), np.uint8); padded[:, :w] = img; padded[:, w:] = 255      # step > width, junk in the padding
    k1, d1 = ex(padded[:, :w])
    assert k1.tobytes() == k0.tobytes() and np.array_equal(d1, d0)
    sparse = np.zeros((h, 3000), np.uint8); sparse[:, :w] = img                             # step > 2*width: staged path
    k2, d2 = ex(sparse[:, :w])
    assert k2.tobytes() == k0.tobytes() and np.array_equal(d2, d0)
    # page-locked input and outputs: direct DMA in both directions
    cap = ex.max_keypoints(w, h)
    pin = torch.from_numpy(img).pin_memory()
    kp = torch.empty((cap, 28), dtype=torch.uint8, pin_memory=True); dp = torch.empty((cap, 32), dtype=torch.uint8, pin_memory=True)
    n = C.c_int()
    check(ex._L.orbx_extract(ex.handle, pin.data_ptr(), w, h, w, kp.data_ptr(), dp.data_ptr(), cap, C.byref(n)), "orbx_extract")
    assert kp.numpy()[: n.value].tobytes() == k0.tobytes() and np.array_equal(dp.numpy()[: n.value], d0)


def test_empty_flat_and_unsupported_images():
    from orb_slam_b200 import FrontendError, ORBextractor
    ex = ORBextractor(500, 1.2, 8, 20, 7)
    assert ex(np.zeros((0, 0), np.uint8)) == (None, None)          # outputs untouched, src/ORBextractor.cc:1252
    k, d = ex(np.full((240, 320), 128, np.uint8))                    # no corners anywhere: 0 keypoints, descriptors released
    assert len(k) == 0 and d.shape == (0, 32)
    with pytest.raises(FrontendError):                               # level 7 of a 60x60 image is narrower than the 19-px border
        ex(np.zeros((60, 60), np.uint8))
    n = C.c_int(5)
    rc = ex._L.orbx_extract(ex.handle, None, 0, 0, 0, None, None, 0, C.byref(n))
    assert rc == 0 and n.value == -1
    # capacity too small is reported, never overrun
    img = synth_frame(640, 480, 1)
    kps = np.zeros((10, 28), np.uint8); desc = np.zeros((10, 32), np.uint8)
    rc = ex._L.orbx_extract(ex.handle, img.ctypes.data, 640, 480, 640, kps.ctypes.data, desc.ctypes.data, 10, C.byref(n))
    assert rc == -3 and n.value > 10


@pytest.mark.parametrize("w,h,nf", [(3840, 2160, 4000)])
def test_4k_stress_frame(w, h, nf):
    from orb_slam_b200 import ORBextractor
    img = synth_frame(w, h, 21)
    ex = ORBextractor(nf, 1.2, 8, 20, 7)
    gk, gd = ex(img)                                     # one frame: 16-row K1 tiles
    ok, od = OracleExtractor(nf, 1.2, 8, 20, 7)(img)
    assert len(ok) >= nf
    same(gk, gd, ok, od)
    for bk, bd in ex.extract_batch(np.stack([img] * 3)):  # a batch: 32-row tiles
        assert bk.tobytes() == gk.tobytes() and np.array_equal(bd, gd)


def test_other_extractor_settings():
    """scaleFactor 1.5 / 4 levels, 2x nFeatures init extractor (Tracking.cc:133), FAST 12/5"""
    from orb_slam_b200 import ORBextractor
    img = synth_frame(752, 480, 33)
    for args in [(1000, 1.5, 4, 20, 7), (2000, 1.2, 8, 20, 7), (800, 1.2, 6, 12, 5), (50, 1.2, 8, 20, 7)]:
        same(*ORBextractor(*args)(img), *OracleExtractor(*args)(img))


def test_euroc_batch_then_search_for_initialization():
    """config 3 in miniature: a batch of EuRoC-shape frames, then ORBmatcher(0.9,true).SearchForInitialization on
    consecutive frames (window 100), all against the oracle"""
    from orb_slam_b200 import ORBextractor
    from orb_slam_b200.matcher import MatcherContext, ORBmatcher
    w, h, nf, n = 752, 480, 1000, 6
    imgs = np.stack([synth_frame(w, h, 60, shift=(3 * i, -2 * i)) for i in range(n)])
    res = ORBextractor(nf, 1.2, 8, 20, 7).extract_batch(imgs)
    cpu = OracleExtractor(nf, 1.2, 8, 20, 7)
    ores = [cpu(im) for im in imgs]
    for a, b in zip(res, ores):
        same(*a, *b)
    ctx = MatcherContext(0)
    for i in range(n - 1):
        F1, F2 = scenes.make_frame(*ores[i], w, h), scenes.make_frame(*ores[i + 1], w, h)
        pg = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32); po = pg.copy()
        ng, mg = ORBmatcher(0.9, True, ctx).SearchForInitialization(F1, F2, pg, 100)
        no, mo = OracleMatcher(0.9, True).SearchForInitialization(F1, F2, po, 100)
        assert ng == no and no > 50 and np.array_equal(mg, mo) and np.array_equal(pg, po)
    ctx.close()


def test_hamming_ragged_sizes():
    import torch
    from orb_slam_b200.matcher import MatcherContext
    from orb_slam_b200.synth import synth_descriptors
    ctx = MatcherContext(0)
    db = synth_descriptors(5000, 3)
    d_db = torch.from_numpy(db).cuda()
    o = OracleMatcher()
    for nq, ndb in [(1, 5000), (1025, 129), (33, 2), (7, 128), (2049, 4097)]:
        q = synth_descriptors(nq, 100 + nq)
        assert ctx.bf2nn(q, d_db.data_ptr(), ndb).tobytes() == o.bf2nn(q, db[:ndb]).tobytes(), (nq, ndb)
    ctx.close()
