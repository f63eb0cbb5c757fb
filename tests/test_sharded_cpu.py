"""CPU (gloo, world size 2 and 3): partitioning and all-gather plumbing of the sharded brute-force 2-NN."""
import os
import socket
import subprocess
import sys

import pytest

from orb_slam_b200.sharded import shard_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_shard_range_is_a_contiguous_balanced_partition():
    for n in (0, 1, 7, 8, 1000, 10_000_000):
        for world in (1, 2, 3, 4, 8):
            cuts = [shard_range(n, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.parametrize("world,ndb", [(2, 20001), (3, 5000), (2, 1)])
def test_gloo_allgather_merge_equals_global_scan(world, ndb):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "tests", "dist_worker.py"), "gloo", str(ndb), "64"]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0 and "DIST_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_pair_slice_covers_every_consecutive_pair_once():
    from orb_slam_b200.sharded import pair_slice
    for n, world in ((1024, 1), (1024, 2), (1024, 8), (1023, 4), (10, 3), (3, 8), (1, 2), (0, 2)):
        pairs = []
        for r in range(world):
            lo, hi, npairs = pair_slice(n, world, r)
            slo, shi = shard_range(n, world, r)
            assert lo == slo and hi in (shi, shi + 1)                  # at most one frame of overlap, taken from the next slice
            assert npairs == 0 or lo + npairs < hi
            pairs += [lo + p for p in range(npairs)]                   # pair (i, i + 1) is named by its first frame
            assert all(slo <= i < shi for i in pairs[len(pairs) - npairs:])      # ... which lies in this rank's own slice
        assert sorted(pairs) == list(range(max(n - 1, 0))) and len(set(pairs)) == len(pairs)


def test_gloo_boundary_pair_belongs_to_the_lower_rank():
    """SearchForInitialization over a 7-frame sequence on 2 ranks: results per pair equal the single-process run"""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "tests", "dist_worker.py"), "gloo", "pairs", "7"]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0 and "DIST_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
