"""GPU parity of the matcher / grid C ABI (include/orbm.h) against the CPU oracle: candidate
enumeration order, distances, match indices and counts are bit-exact."""
import numpy as np
import pytest

from orb_slam_b200.synth import synth_descriptors
from tests import scenes
from tests.oracle_lib import OracleExtractor, OracleMatcher

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from orb_slam_b200.matcher import MatcherContext
    c = MatcherContext(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def pair():
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 752, 480, 1000, 30, (7, -4))
    return scenes.make_frame(ka, da, 752, 480), scenes.make_frame(kb, db, 752, 480)


def test_descriptor_distance_and_three_maxima():
    from orb_slam_b200.matcher import ORBmatcher
    o = OracleMatcher()
    d = synth_descriptors(400, 5)
    for i in range(0, 400, 2):
        assert ORBmatcher.DescriptorDistance(d[i], d[i + 1]) == o.DescriptorDistance(d[i], d[i + 1])
    assert ORBmatcher.DescriptorDistance(d[0], d[0]) == 0
    assert ORBmatcher.DescriptorDistance(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256
    rng = np.random.default_rng(1)
    for t in range(200):
        c = rng.integers(0, 40 if t % 2 else 3, 30).astype(np.int32)
        if t % 7 == 0:
            c[:] = 0
        assert ORBmatcher.ComputeThreeMaxima(c) == o.ComputeThreeMaxima(c)


def test_grid_and_features_in_area(ctx, pair):
    F1, F2 = pair
    o = OracleMatcher()
    for F in (F1, F2):
        cs, it = ctx.grid_build(F)
        ocs, oit = o.grid_build(F)
        assert np.array_equal(cs, ocs) and np.array_equal(it, oit)
    rng = np.random.default_rng(3)
    for t in range(60):
        x, y = float(rng.uniform(-30, 790)), float(rng.uniform(-30, 510))
        r = float(rng.choice([5, 15, 40, 100, 1000]))
        lv = [(-1, -1), (0, 0), (1, 3), (2, -1), (0, 7)][t % 5]
        assert np.array_equal(ctx.features_in_area(F2, x, y, r, *lv), o.features_in_area(F2, x, y, r, *lv)), (x, y, r, lv)


@pytest.mark.parametrize("ratio,ori,win", [(0.9, True, 100), (0.9, False, 100), (0.6, True, 30)])
def test_search_for_initialization(ctx, pair, ratio, ori, win):
    from orb_slam_b200.matcher import ORBmatcher
    F1, F2 = pair
    prev_g = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
    prev_o = prev_g.copy()
    n_g, m_g = ORBmatcher(ratio, ori, ctx).SearchForInitialization(F1, F2, prev_g, win)
    n_o, m_o = OracleMatcher(ratio, ori).SearchForInitialization(F1, F2, prev_o, win)
    assert n_g == n_o and n_o > 20
    assert np.array_equal(m_g, m_o)
    assert np.array_equal(prev_g, prev_o)


@pytest.mark.parametrize("stereo,th,mono", [(False, 15.0, True), (True, 7.0, False), (True, 14.0, False)])
def test_search_by_projection_frame(ctx, stereo, th, mono):
    from orb_slam_b200.matcher import ORBmatcher
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 31, (3, 1))
    last = scenes.make_frame(ka, da, 1241, 376, stereo=stereo)
    Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [0.02, -0.01, -0.9 if stereo else 0.05]
    cur = scenes.make_frame(kb, db, 1241, 376, stereo=stereo, Tcw=Tcw, seed=1)
    pts = scenes.last_frame_points(last, seed=2)
    cur_obs = np.random.default_rng(4).choice([-1, -1, -1, 0, 2], cur.N).astype(np.int32)
    n_g, a_g = ORBmatcher(0.9, True, ctx).SearchByProjectionFrame(cur, last, pts, cur_obs, th, mono)
    n_o, a_o = OracleMatcher(0.9, True).SearchByProjectionFrame(cur, last, pts, cur_obs, th, mono)
    assert n_g == n_o and (a_o >= 0).sum() > 20
    assert np.array_equal(a_g, a_o)


@pytest.mark.parametrize("th", [1.0, 3.0, 5.0])
def test_search_by_projection_points(ctx, th):
    from orb_slam_b200.matcher import ORBmatcher
    (ka, da), _ = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 32, (0, 0))
    F = scenes.make_frame(ka, da, 1241, 376, stereo=True)
    pts = scenes.local_map_points(F, seed=5)
    obs = np.random.default_rng(6).choice([-1, -1, 0, 3], F.N).astype(np.int32)
    n_g, a_g = ORBmatcher(0.8, True, ctx).SearchByProjectionPoints(F, pts, obs, th)
    n_o, a_o = OracleMatcher(0.8, True).SearchByProjectionPoints(F, pts, obs, th)
    assert n_g == n_o and n_o > 50
    assert np.array_equal(a_g, a_o)


@pytest.mark.parametrize("th,orbdist,tz", [(10.0, 100, 0.3), (3.0, 64, -0.4), (10.0, 100, 4.0)])
def test_search_by_projection_keyframe(ctx, th, orbdist, tz):
    """Tracking::Relocalization's SearchByProjection(CurrentFrame, pKF, sFound, th, ORBdist) (:1540-1667)"""
    from orb_slam_b200.matcher import ORBmatcher
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 33, (4, -2))
    kf = scenes.make_frame(ka, da, 1241, 376)
    Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [0.03, 0.01, tz]
    cur = scenes.make_frame(kb, db, 1241, 376, Tcw=Tcw, seed=3)
    pts, mm, already = scenes.keyframe_points(kf, seed=11)
    has = (np.random.default_rng(12).random(cur.N) < 0.3).astype(np.uint8)
    args = (cur, pts, mm, already, kf.mvKeysUn["angle"], has, th, orbdist)
    n_g, a_g = ORBmatcher(0.9, True, ctx).SearchByProjectionKeyFrame(*args)
    n_o, a_o = OracleMatcher(0.9, True).SearchByProjectionKeyFrame(*args)
    assert n_g == n_o
    assert np.array_equal(a_g, a_o)
    if tz < 1:
        assert (a_o >= 0).sum() > 20
    m = a_g[a_g >= 0]
    assert not np.any(has[a_g >= 0]) and not np.any(already[m]) and np.all(pts["valid"][m] == 1) and not np.any(pts["bad"][m])


@pytest.mark.parametrize("stereo,th,tz", [(True, 3.0, 0.2), (False, 3.0, -0.3), (True, 6.0, 1.5)])
def test_fuse_select(ctx, stereo, th, tz):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th) (:893-1040): the keypoint every map point fuses with"""
    from orb_slam_b200.matcher import ORBmatcher
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 35, (-3, 2))
    src = scenes.make_frame(ka, da, 1241, 376)
    Tcw = np.eye(4, dtype=np.float32); Tcw[:3, 3] = [-0.05, 0.02, tz]
    kf = scenes.make_frame(kb, db, 1241, 376, stereo=stereo, Tcw=Tcw, seed=7)
    if not stereo:
        kf.mvuRight = np.full(kf.N, -1, np.float32)
    pts, mm, normals, in_kf = scenes.fuse_points(src, seed=21)
    sf = kf.mvScaleFactors
    inv_sigma2 = (1.0 / (sf * sf)).astype(np.float32)
    args = (kf, inv_sigma2, pts, in_kf, normals, mm, th)
    g = ORBmatcher(0.6, True, ctx).FuseSelect(*args)
    o = OracleMatcher(0.6, True).FuseSelect(*args)
    assert np.array_equal(g, o)
    sel = g >= 0
    assert sel.sum() > 20
    assert not np.any(in_kf[sel]) and np.all(pts["valid"][sel] == 1) and not np.any(pts["bad"][sel])


@pytest.mark.parametrize("scale,tz,th", [(1.0, 0.2, 10), (1.15, -0.3, 10), (0.9, 1.0, 4)])
def test_sim3_matchers(ctx, scale, tz, th):
    """Loop closing's SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:290-403) and Fuse(pKF, Scw, ...) (:1044-1168)"""
    from orb_slam_b200.matcher import ORBmatcher
    (ka, da), (kb, db) = scenes.extract_pair(OracleExtractor, 1241, 376, 2000, 37, (2, 3))
    src = scenes.make_frame(ka, da, 1241, 376)
    kf = scenes.make_frame(kb, db, 1241, 376)
    pts, mm, normals, in_kf = scenes.fuse_points(src, seed=31)
    pts["valid"] = 1
    excluded = (pts["bad"].astype(bool) | in_kf.astype(bool)).astype(np.uint8)
    # a similarity with scale `scale`: the functions receive it decomposed (rotation without the scale, tcw / s, camera centre)
    a = np.float32(0.01)
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]], np.float32)
    t = (np.array([0.03, -0.02, tz], np.float32) / np.float32(scale)).astype(np.float32)
    Ow = (-(R.T.astype(np.float64) @ t.astype(np.float64))).astype(np.float32)
    matched_in = (np.random.default_rng(32).random(kf.N) < 0.3).astype(np.uint8)
    n_g, m_g = ORBmatcher(0.75, True, ctx).SearchByProjectionSim3(kf, R, t, Ow, pts, excluded, normals, mm, matched_in, th)
    n_o, m_o = OracleMatcher(0.75, True).SearchByProjectionSim3(kf, R, t, Ow, pts, excluded, normals, mm, matched_in, th)
    assert n_g == n_o and n_o > 20
    assert np.array_equal(m_g, m_o)
    hit = m_g >= 0
    assert not np.any(matched_in[hit]) and not np.any(excluded[m_g[hit]]) and len(np.unique(m_g[hit])) == hit.sum()
    b_g = ORBmatcher(0.75, True, ctx).FuseSelectSim3(kf, R, t, Ow, pts, excluded, normals, mm, float(th))
    b_o = OracleMatcher(0.75, True).FuseSelectSim3(kf, R, t, Ow, pts, excluded, normals, mm, float(th))
    assert np.array_equal(b_g, b_o) and (b_o >= 0).sum() > 20
    assert not np.any(excluded[b_g >= 0])


@pytest.mark.parametrize("seed,s12,th", [(0, 1.08, 7.5), (1, 0.93, 7.5), (2, 1.0, 3.0)])
def test_search_by_sim3(ctx, seed, s12, th):
    """ORBmatcher::SearchBySim3 (:1170-1394): both projection directions and the agreement check"""
    from orb_slam_b200.matcher import ORBmatcher
    sc = scenes.sim3_scene(seed=seed, s12=s12)
    args = (sc["kf1"], sc["kf2"], sc["pts1"], sc["pts2"], sc["skip1"], sc["skip2"], sc["mm1"], sc["mm2"], sc["sR12"], sc["t12"],
            sc["sR21"], sc["t21"], th)
    n_g, m_g = ORBmatcher(0.75, True, ctx).SearchBySim3(*args)
    n_o, m_o = OracleMatcher(0.75, True).SearchBySim3(*args)
    assert n_g == n_o and np.array_equal(m_g, m_o)
    assert n_o > 300
    i1 = np.nonzero(m_o >= 0)[0]
    assert not np.any(sc["skip1"][i1]) and not np.any(sc["skip2"][m_o[i1]]) and len(np.unique(m_o[i1])) == len(i1)
    shared = i1[i1 < len(sc["i2"])]
    assert (m_o[shared] == sc["i2"][shared]).mean() > 0.95              # the planted correspondences are what it finds


def test_bruteforce_2nn_and_sharded_merge(ctx):
    import ctypes as C
    import torch
    from orb_slam_b200.matcher import TOP2_DTYPE
    ndb, nq = 150001, 300
    db = synth_descriptors(ndb, 0xDB)
    rng = np.random.default_rng(9)
    q = db[rng.integers(0, ndb, nq)].copy()
    for i in range(nq):                                    # k in [0,40] random bit flips
        for b in rng.integers(0, 256, rng.integers(0, 41)):
            q[i, b >> 3] ^= 1 << (b & 7)
    q[:5] = db[7]                                          # duplicates: ties must resolve to the lowest index
    db[1000] = db[7]
    ref = OracleMatcher().bf2nn(q, db)
    d_db = torch.from_numpy(db).cuda()
    got = ctx.bf2nn(q, d_db.data_ptr(), ndb)
    assert got.tobytes() == ref.tobytes()
    assert (got["i1"][:5] == 7).all() and (got["d1"][:5] == 0).all() and (got["d2"][:5] == 0).all()
    # three uneven shards with global indices, merged exactly (what the NCCL all-gather path does)
    cuts = [0, 40000, 40001, ndb]
    parts = np.concatenate([ctx.bf2nn(q, d_db.data_ptr() + 32 * cuts[i], cuts[i + 1] - cuts[i], cuts[i]) for i in range(3)])
    d_parts = torch.from_numpy(parts.view(np.uint8)).cuda()
    d_out = torch.zeros(nq * 8, dtype=torch.uint8, device="cuda")
    L = ctx._L
    assert L.orbm_merge_top2_device(ctx.h, d_parts.data_ptr(), 3, nq, d_out.data_ptr()) == 0
    assert L.orbm_synchronize(ctx.h) == 0
    merged = d_out.cpu().numpy().view(TOP2_DTYPE)
    assert merged.tobytes() == ref.tobytes()
    # edge: fewer than two rows
    one = ctx.bf2nn(q[:3], d_db.data_ptr(), 1)
    assert (one["d2"] == 256).all() and (one["i1"] == 0).all()


def test_sharded_database_single_rank(ctx):
    """ShardedDatabase at world size 1 (no collective) equals the oracle scan."""
    from orb_slam_b200.sharded import ShardedDatabase
    db = synth_descriptors(70001, 0xDB)
    q = db[::997].copy(); q[:, 5] ^= 0x5A
    got = ShardedDatabase(db, 0, context=ctx).query(q)
    assert got.tobytes() == OracleMatcher().bf2nn(q, db).tobytes()


def test_sharded_database_nccl_two_ranks():
    """the product's distributed path (NCCL all-gather + merge kernel) on 2 GPUs, when the box has them"""
    import os, socket, subprocess, sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(root, "tests", "dist_worker.py"), "nccl", "300001", "512"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0 and "DIST_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_sharded_window_matching_nccl_two_ranks():
    """SearchForInitialization over a frame sequence split across 2 GPUs (pair_slice: the boundary pair stays with the
    lower rank), product extraction + batch matcher on each rank, per-pair results against the oracle"""
    import os, socket, subprocess, sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(root, "tests", "dist_worker.py"), "nccl", "pairs", "9"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0 and "DIST_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
