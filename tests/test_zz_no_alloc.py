"""orbm_bf2nn_device_noalloc (include/orbm.h): the brute-force 2-NN with its scratch in a fixed file-scope device array,
for callers that cannot call cudaMalloc.  It has no counterpart in the reference;
its results must equal orbm_bf2nn_device's bit for bit (integer work), here on two calls in a row that
share the array.  Collected last on purpose (file name)."""
import numpy as np
import pytest

from orb_slam_b200.synth import synth_descriptors

pytestmark = pytest.mark.gpu

MAX_Q = 4096           # ORBM_NOALLOC_MAX_QUERIES


def _both(ctx, d_q, nq, d_db, ndb, base):
    import torch
    from orb_slam_b200.matcher import TOP2_DTYPE
    L = ctx._L
    a = torch.zeros(nq * 8, dtype=torch.uint8, device="cuda")
    b = torch.full((nq * 8,), 0xAB, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    assert L.orbm_bf2nn_device(ctx.h, d_q.data_ptr(), nq, d_db.data_ptr(), ndb, base, a.data_ptr()) == 0
    assert L.orbm_bf2nn_device_noalloc(ctx.h, d_q.data_ptr(), nq, d_db.data_ptr(), ndb, base, b.data_ptr()) == 0
    assert L.orbm_synchronize(ctx.h) == 0
    return a.cpu().numpy().view(TOP2_DTYPE), b.cpu().numpy().view(TOP2_DTYPE)


def test_noalloc_bf2nn_equals_the_allocating_entry_point_twice_in_a_row():
    import torch
    from orb_slam_b200.matcher import MatcherContext
    from tests.oracle_lib import OracleMatcher
    ctx = MatcherContext(0)
    try:
        # call 1: the largest query count the entry point takes, ties included
        ndb1 = 150_001
        db1 = synth_descriptors(ndb1, 0xA1)
        q1 = db1[(np.arange(MAX_Q) * 31) % ndb1].copy()
        q1[:, 7] ^= (np.arange(MAX_Q) % 251).astype(np.uint8)
        q1[:4] = db1[9]; db1[140_000] = db1[9]                      # duplicates: lowest row wins
        d_db1, d_q1 = torch.from_numpy(db1).cuda(), torch.from_numpy(q1).cuda()
        want, got = _both(ctx, d_q1, MAX_Q, d_db1, ndb1, 1000)
        assert got.tobytes() == want.tobytes()
        assert (got["i1"][:4] == 1009).all() and (got["d2"][:4] == 0).all()
        # call 2, straight after: fewer queries (more slices per query tile, other records of the shared array), other rows
        ndb2, nq2 = 70_001, 300
        db2 = synth_descriptors(ndb2, 0xB2)
        q2 = db2[::233][:nq2].copy(); q2[:, 0] ^= 0x81
        d_db2, d_q2 = torch.from_numpy(db2).cuda(), torch.from_numpy(q2).cuda()
        want2, got2 = _both(ctx, d_q2, nq2, d_db2, ndb2, 0)
        assert got2.tobytes() == want2.tobytes()
        assert got2.tobytes() == OracleMatcher().bf2nn(q2, db2).tobytes()      # and both equal the oracle's scan
        # call 3: the first case again - nothing of call 2 may leak through the shared scratch
        want3, got3 = _both(ctx, d_q1, MAX_Q, d_db1, ndb1, 1000)
        assert got3.tobytes() == want.tobytes() == want3.tobytes()
        # edge: a single row
        w1, g1 = _both(ctx, d_q2, 3, d_db2, 1, 0)
        assert g1.tobytes() == w1.tobytes() and (g1["d2"] == 256).all()
    finally:
        ctx.close()


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from orb_slam_b200.matcher import MatcherContext
    ctx = MatcherContext(0)
    try:
        nq = MAX_Q + 1
        d_q = torch.zeros((nq, 32), dtype=torch.uint8, device="cuda")
        d_db = torch.zeros((256, 32), dtype=torch.uint8, device="cuda")
        out = torch.zeros(nq * 8, dtype=torch.uint8, device="cuda")
        rc = ctx._L.orbm_bf2nn_device_noalloc(ctx.h, d_q.data_ptr(), nq, d_db.data_ptr(), 256, 0, out.data_ptr())
        assert rc == -3                                                # ORBX_E_CAPACITY
        assert ctx._L.orbm_bf2nn_device_noalloc(ctx.h, d_q.data_ptr(), 0, d_db.data_ptr(), 256, 0, out.data_ptr()) == -1
    finally:
        ctx.close()
