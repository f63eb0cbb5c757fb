"""torchrun worker for the sharded brute-force 2-NN (tests only).

backend gloo (CPU, no GPU needed): exercises the product's partitioning + all-gather plumbing
(orb_slam_b200.sharded.shard_range / gather_parts); the per-shard scan and the merge - CUDA kernels in the
product - are played by the oracle and a numpy merge.
backend nccl (>= 2 GPUs): the product path end to end (ShardedDatabase.query: orbm_shard_* = bf2nn_kernel +
ncclAllGather + merge_parts_kernel inside the C library) against the oracle.
mode `pairs` (argv[2]): SearchForInitialization over the consecutive pairs of a short frame sequence split with
orb_slam_b200.sharded.pair_slice - the pair straddling two ranks' slices belongs to the lower rank, which extracts one
frame of overlap.  gloo: extraction and matching by the oracle; nccl: by the product (batch extraction + batch matcher)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from orb_slam_b200.matcher import TOP2_DTYPE                     # noqa: E402
from orb_slam_b200.sharded import gather_parts, shard_range      # noqa: E402
from orb_slam_b200.synth import synth_descriptors                # noqa: E402
from tests.oracle_lib import OracleMatcher                       # noqa: E402


def numpy_merge(parts):
    """[world, nq] TOP2 -> global top-2 (best = min (d1, i1); second = next smallest distance)."""
    world, nq = parts.shape
    out = np.zeros(nq, TOP2_DTYPE)
    for q in range(nq):
        cands = []
        for p in range(world):
            r = parts[p, q]
            cands.append((int(r["d1"]), int(r["i1"])))
            cands.append((int(r["d2"]), 1 << 40))
        cands.sort()
        out[q] = (cands[0][1], cands[0][0], cands[1][0])
    return out


def run_pairs(backend):
    """every rank: its slice of the sequence; rank 0 also the whole sequence in one go; per-pair results must agree"""
    from orb_slam_b200.sharded import pair_slice
    from orb_slam_b200.synth import synth_frame
    from tests import scenes
    from tests.oracle_lib import OracleExtractor
    nframes, w, h, nf = int(sys.argv[3]), 320, 240, 500
    rank, world = dist.get_rank(), dist.get_world_size()
    frames = [synth_frame(w, h, 900, shift=(2 * i, -i)) for i in range(nframes)]
    lo, hi, npairs = pair_slice(nframes, world, rank)

    def oracle_pairs(a, b, n):
        ex = OracleExtractor(nf, 1.2, 8, 20, 7)
        kd = [ex(frames[i]) for i in range(a, b)]
        om = OracleMatcher(0.9, True)
        out = []
        for p in range(n):
            F1 = scenes.make_frame(*kd[p], w, h); F2 = scenes.make_frame(*kd[p + 1], w, h)
            prev = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
            nm, m12 = om.SearchForInitialization(F1, F2, prev, 100)
            out.append((nm, int(np.sum((m12.astype(np.int64) + 2) * (np.arange(len(m12)) + 1)) % 1000003)))
        return out

    if backend == "gloo":
        mine = oracle_pairs(lo, hi, npairs)
    else:
        import ctypes as C
        from orb_slam_b200 import ORBextractor
        from orb_slam_b200._lib import check
        from orb_slam_b200.matcher import MatcherContext
        local_rank = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local_rank)
        ex = ORBextractor(nf, 1.2, 8, 20, 7, device=local_rank); ctx = MatcherContext(local_rank)
        cap = ex.max_keypoints(w, h); n = hi - lo
        d_img = torch.from_numpy(np.stack(frames[lo:hi])).cuda()
        d_k = torch.zeros((n, cap, 28), dtype=torch.uint8, device="cuda"); d_d = torch.zeros((n, cap, 32), dtype=torch.uint8, device="cuda")
        d_n = torch.zeros(n, dtype=torch.int32, device="cuda")
        d_m = torch.full((max(npairs, 1), cap), -1, dtype=torch.int32, device="cuda"); d_nm = torch.zeros(max(npairs, 1), dtype=torch.int32, device="cuda")
        check(ex._L.orbx_extract_batch_device(ex.handle, d_img.data_ptr(), n, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap,
                                              d_n.data_ptr()), "orbx_extract_batch_device")
        check(ex._L.orbx_synchronize(ex.handle), "sync")
        if npairs:
            ctx.search_for_initialization_batch_device(d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), cap, n, w, h, npairs=npairs,
                                                       d_matches12=d_m.data_ptr(), d_nmatches=d_nm.data_ptr())
            ctx.synchronize()
        nm = d_nm.cpu().numpy(); m = d_m.cpu().numpy(); nn = d_n.cpu().numpy()
        mine = [(int(nm[p]), int(np.sum((m[p, :nn[p]].astype(np.int64) + 2) * (np.arange(nn[p]) + 1)) % 1000003)) for p in range(npairs)]
    # gather (pair index, count, checksum) rows, padded to the longest slice
    longest = (nframes + world - 1) // world + 1
    t = torch.full((longest, 3), -1, dtype=torch.int64)
    for p, (a, b) in enumerate(mine):
        t[p] = torch.tensor([lo + p, a, b])
    if backend == "nccl":
        t = t.cuda()
    parts = gather_parts(t).cpu().numpy().reshape(-1, 3)
    got = {int(r[0]): (int(r[1]), int(r[2])) for r in parts if r[0] >= 0}
    ok = True
    if rank == 0:
        ref = oracle_pairs(0, nframes, nframes - 1)
        ok = len(parts[parts[:, 0] >= 0]) == nframes - 1 and got == {p: ref[p] for p in range(nframes - 1)}
        print("DIST_OK" if ok else "DIST_MISMATCH pairs %s vs %s" % (got, ref), world, flush=True)
    flag = torch.tensor([1 if ok else 0])
    if backend == "nccl":
        flag = flag.cuda()
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


def main():
    backend = sys.argv[1]
    dist.init_process_group(backend)
    if sys.argv[2] == "pairs":
        return run_pairs(backend)
    ndb, nq = int(sys.argv[2]), int(sys.argv[3])
    rank, world = dist.get_rank(), dist.get_world_size()
    db = synth_descriptors(ndb, 0xDB)
    rng = np.random.default_rng(5)
    q = db[rng.integers(0, ndb, nq)].copy()
    q[:, 3] ^= rng.integers(0, 256, nq).astype(np.uint8)
    if ndb > 8:
        q[:4] = db[2]; db[ndb - 1] = db[2]                           # a tie across shards: lowest global index wins
    lo, hi = shard_range(ndb, world, rank)
    if ndb > 64 * world:
        for r in range(world - 1):                                   # every shard boundary: last row of r == first row of r + 1
            b = shard_range(ndb, world, r)[1]
            db[b] = db[b - 1]
            q[4 + r] = db[b - 1]                                     # best = (0, b - 1), second distance 0 from the next shard
    ref = OracleMatcher().bf2nn(q, db)
    if backend == "gloo":
        local = OracleMatcher().bf2nn(q, db[lo:hi], idx_base=lo) if hi > lo else None
        if local is None:
            local = np.zeros(nq, TOP2_DTYPE); local["d1"] = 256; local["d2"] = 256; local["i1"] = 0xFFFFFFFF
        t = torch.from_numpy(local.view(np.uint8).reshape(nq, 8).copy())
        parts = gather_parts(t).numpy().view(TOP2_DTYPE).reshape(world, nq)
        got = numpy_merge(parts)
    else:
        from orb_slam_b200.sharded import ShardedDatabase
        local_rank = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local_rank)
        sd = ShardedDatabase(db[lo:hi], lo, device=local_rank)
        got = sd.query(q)
    ok = got.tobytes() == ref.tobytes()
    flag = torch.tensor([1 if ok else 0])
    if backend == "nccl":
        flag = flag.cuda()
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("DIST_OK" if int(flag.item()) == 1 else "DIST_MISMATCH", world, lo, hi, flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
