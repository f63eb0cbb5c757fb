"""torchrun worker for the sharded brute-force 2-NN (tests only).

backend gloo (CPU, no GPU needed): exercises the product's partitioning + all-gather plumbing
(orb_slam_b200.sharded.shard_range / gather_parts); the per-shard scan and the merge - CUDA kernels in the
product - are played by the oracle and a numpy merge.
backend nccl (>= 2 GPUs): the product path end to end (ShardedDatabase.query) against the oracle."""
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


def main():
    backend = sys.argv[1]
    ndb, nq = int(sys.argv[2]), int(sys.argv[3])
    dist.init_process_group(backend)
    rank, world = dist.get_rank(), dist.get_world_size()
    db = synth_descriptors(ndb, 0xDB)
    rng = np.random.default_rng(5)
    q = db[rng.integers(0, ndb, nq)].copy()
    q[:, 3] ^= rng.integers(0, 256, nq).astype(np.uint8)
    if ndb > 8:
        q[:4] = db[2]; db[ndb - 1] = db[2]                           # a tie across shards: lowest global index wins
    lo, hi = shard_range(ndb, world, rank)
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
