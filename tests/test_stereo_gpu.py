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


@pytest.mark.parametrize("w,h,nf,n", [(1241, 376, 2000, 5), (752, 480, 1000, 2), (640, 480, 500, 1)])
def test_stereo_batch_device_equals_oracle_frame_by_frame(w, h, nf, n):
    """orbx_stereo_match_batch_device: n stereo pairs extracted and matched without leaving the device; mvuRight / mvDepth of every
    pair bit-exact against the oracle run on that pair's own key points, descriptors and pyramids."""
    import torch
    from orb_slam_b200 import KEYPOINT_DTYPE, ORBextractor
    from orb_slam_b200._lib import check
    disps = [23, 9, 0, 41, 15][:n]
    lefts = np.stack([synth_frame(w, h, 80 + i) for i in range(n)])
    rights = np.stack([synth_frame(w, h, 80 + i, shift=(-disps[i], 0)) for i in range(n)])
    rng = np.random.default_rng(3)
    rights = np.clip(rights.astype(np.int16) + rng.integers(-3, 4, rights.shape), 0, 255).astype(np.uint8)   # sensor noise
    if n > 1:
        rights[1] = 90                                                 # a textureless right image: no right key points
    mbf = float(KITTI_BF); mb = float(np.float32(KITTI_BF) / KITTI_K[0, 0])
    gl, gr = ORBextractor(nf, 1.2, 8, 20, 7), ORBextractor(nf, 1.2, 8, 20, 7)
    cap = gl.max_keypoints(w, h)
    dev = {}
    for name, ex, imgs in (("L", gl, lefts), ("R", gr, rights)):
        d_img = torch.from_numpy(imgs).cuda()
        d_k = torch.zeros((n, cap, 28), dtype=torch.uint8, device="cuda"); d_d = torch.zeros((n, cap, 32), dtype=torch.uint8, device="cuda")
        d_n = torch.zeros(n, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        check(ex._L.orbx_extract_batch_device(ex.handle, d_img.data_ptr(), n, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap,
                                              d_n.data_ptr()), "orbx_extract_batch_device")
        dev[name] = (d_img, d_k, d_d, d_n)
    d_u = torch.full((n, cap), 7.0, dtype=torch.float32, device="cuda"); d_z = torch.full((n, cap), 7.0, dtype=torch.float32, device="cuda")
    d_nm = torch.full((n,), -5, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    gl.stereo_match_batch_device(gr, n, dev["L"][1].data_ptr(), dev["L"][2].data_ptr(), dev["L"][3].data_ptr(), dev["R"][1].data_ptr(),
                                 dev["R"][2].data_ptr(), dev["R"][3].data_ptr(), cap, mbf, mb, d_u.data_ptr(), d_z.data_ptr(), d_nm.data_ptr())
    gl.synchronize()
    u, z, nm = d_u.cpu().numpy(), d_z.cpu().numpy(), d_nm.cpu().numpy()
    kL = dev["L"][1].cpu().numpy().view(KEYPOINT_DTYPE).reshape(n, cap); dL = dev["L"][2].cpu().numpy(); nL = dev["L"][3].cpu().numpy()
    kR = dev["R"][1].cpu().numpy().view(KEYPOINT_DTYPE).reshape(n, cap); dR = dev["R"][2].cpu().numpy(); nR = dev["R"][3].cpu().numpy()
    for i in range(n):
        ol, orr = OracleExtractor(nf, 1.2, 8, 20, 7), OracleExtractor(nf, 1.2, 8, 20, 7)
        ol(lefts[i]); orr(rights[i])                                   # the oracle's pyramids (bit-exact with the device's)
        ou, oz, on = oracle_stereo_matches(ol, orr, kL[i, :nL[i]], dL[i, :nL[i]], kR[i, :nR[i]], dR[i, :nR[i]].reshape(-1, 32), mbf, mb)
        assert nm[i] == on, i
        assert u[i, :nL[i]].tobytes() == ou.tobytes() and z[i, :nL[i]].tobytes() == oz.tobytes(), i
        assert (u[i, nL[i]:] == 7.0).all()                             # entries past the frame's count untouched
        if i != 1:
            assert on > 0.3 * nL[i]
        else:
            assert on == 0 and nR[i] == 0
    gl.close(); gr.close()


@pytest.mark.parametrize("sf,nlevels", [(1.5, 8), (1.2, 12)])
def test_stereo_wide_row_bands_take_the_scan_path(sf, nlevels):
    """Extractor settings whose top-level row band (2 * ceil(2 * scaleFactor[top]) + 3 rows per right key point) exceeds the row
    list's per-point capacity of 20: scaleFactor 1.5 with 8 levels gives 71 rows, 1.2 with 12 levels 33.  The row list would
    overflow (advisor finding, round 1); such configurations scan every right key point instead.  Results against the oracle,
    single pair and batch."""
    import torch
    from orb_slam_b200 import KEYPOINT_DTYPE, ORBextractor
    from orb_slam_b200._lib import check
    w, h, nf, disp = 1280, 960, 1500, 17
    left = synth_frame(w, h, 91)
    right = synth_frame(w, h, 91, shift=(-disp, 0))
    mbf = float(KITTI_BF); mb = float(np.float32(KITTI_BF) / KITTI_K[0, 0])
    gl, gr = ORBextractor(nf, sf, nlevels, 20, 7), ORBextractor(nf, sf, nlevels, 20, 7)
    ol, orr = OracleExtractor(nf, sf, nlevels, 20, 7), OracleExtractor(nf, sf, nlevels, 20, 7)
    try:
        kl, dl = gl(left); kr, dr = gr(right)
        okl, odl = ol(left); okr, odr = orr(right)
        assert kl.tobytes() == okl.tobytes() and kr.tobytes() == okr.tobytes()
        wide = [l for l in range(nlevels) if 2 * int(np.ceil(2 * sf ** l)) + 3 > 20]
        assert np.isin(okr["octave"], wide).sum() > 10               # right key points whose band exceeds the row list's per-point capacity
        u, z, n = gl.stereo_match(gr, okl, odl, okr, odr, mbf, mb)
        ou, oz, on = oracle_stereo_matches(ol, orr, okl, odl, okr, odr, mbf, mb)
        assert n == on and on > 0.2 * len(okl)
        assert u.tobytes() == ou.tobytes() and z.tobytes() == oz.tobytes()
        # the batch entry point on the same pair (twice, so a frame's rows lie next to another frame's)
        cap = gl.max_keypoints(w, h)
        bufs = {}
        for name, ex, img in (("L", gl, left), ("R", gr, right)):
            d_img = torch.from_numpy(np.stack([img, img])).cuda()
            d_k = torch.zeros((2, cap, 28), dtype=torch.uint8, device="cuda"); d_d = torch.zeros((2, cap, 32), dtype=torch.uint8, device="cuda")
            d_n = torch.zeros(2, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            check(ex._L.orbx_extract_batch_device(ex.handle, d_img.data_ptr(), 2, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap,
                                                  d_n.data_ptr()), "orbx_extract_batch_device")
            bufs[name] = (d_img, d_k, d_d, d_n)
        d_u = torch.zeros((2, cap), dtype=torch.float32, device="cuda"); d_z = torch.zeros((2, cap), dtype=torch.float32, device="cuda")
        d_nm = torch.zeros(2, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        gl.stereo_match_batch_device(gr, 2, bufs["L"][1].data_ptr(), bufs["L"][2].data_ptr(), bufs["L"][3].data_ptr(), bufs["R"][1].data_ptr(),
                                     bufs["R"][2].data_ptr(), bufs["R"][3].data_ptr(), cap, mbf, mb, d_u.data_ptr(), d_z.data_ptr(), d_nm.data_ptr())
        gl.synchronize()
        nL = int(bufs["L"][3][0].item())
        kLd = bufs["L"][1].cpu().numpy().view(KEYPOINT_DTYPE).reshape(2, cap); dLd = bufs["L"][2].cpu().numpy()
        kRd = bufs["R"][1].cpu().numpy().view(KEYPOINT_DTYPE).reshape(2, cap); dRd = bufs["R"][2].cpu().numpy()
        nR = int(bufs["R"][3][0].item())
        ou2, oz2, on2 = oracle_stereo_matches(ol, orr, kLd[0, :nL], dLd[0, :nL], kRd[0, :nR], dRd[0, :nR].reshape(-1, 32), mbf, mb)
        for f in range(2):
            assert int(d_nm[f].item()) == on2
            assert d_u[f, :nL].cpu().numpy().tobytes() == ou2.tobytes() and d_z[f, :nL].cpu().numpy().tobytes() == oz2.tobytes()
    finally:
        gl.close(); gr.close()
