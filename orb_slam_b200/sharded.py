"""Multi-GPU modes of the front end: one process per GPU.

* frames-parallel extraction: `shard_range` gives every rank a contiguous frame slice; no collective.
* windowed matching of consecutive frames (SearchForInitialization over a frame sequence, BASELINE.json configs[2]):
  `pair_slice` - pairs stay with the rank that extracted their first frame; the pair that straddles two slices
  belongs to the LOWER rank, which therefore extracts one extra frame (the first frame of the next slice).  No
  collective either.
* brute-force Hamming 2-NN over a keyframe-descriptor database split across GPUs: `ShardedDatabase` is a thin
  binding over the C ABI (`orbm_shard_*`, include/orbm.h): bf2nn over this rank's rows with global row indices,
  `ncclAllGather` of the per-rank (i1, d1, d2) records on the matcher stream (nq x 8 bytes per rank over NVLink), exact
  merge kernel.  Scan, collective and merge are all issued by the C++ library; torch.distributed only carries the
  128-byte NCCL unique id from rank 0 to the others at construction.

The compute is always the CUDA library; there is no CPU path here.  `shard_range`, `pair_slice` and `gather_parts`
are backend-agnostic (they are exercised with gloo at world size 2 by the CPU tests).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .matcher import TOP2_DTYPE, MatcherContext


def shard_range(n: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous, balanced split of n units (frames or database rows): first `n % world` ranks get one extra."""
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def pair_slice(nframes: int, world: int, rank: int) -> tuple[int, int, int]:
    """Consecutive-pair matching over a frame sequence split with shard_range: returns (lo, hi, npairs) - this rank
    extracts frames [lo, hi) and matches the pairs (lo, lo+1) ... (lo+npairs-1, lo+npairs).  Pair (i, i+1) belongs to
    the rank whose frame slice holds i, so every rank but the last extracts one frame of overlap (its hi is the next
    rank's lo + 1) and the nframes - 1 pairs are covered exactly once."""
    lo, hi = shard_range(nframes, world, rank)
    if hi <= lo:
        return lo, lo, 0
    if hi < nframes:
        return lo, hi + 1, hi - lo
    return lo, hi, hi - lo - 1


def gather_parts(local, group=None):
    """all_gather of equally-shaped per-rank partial results -> tensor [world, ...] through torch.distributed (gloo for
    CPU tensors, NCCL for CUDA tensors).  World size 1: a view, no collective.  The product's 2-NN path does this with
    ncclAllGather inside the C library; this helper serves host-side gathers of per-rank results and the CPU tests."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local.unsqueeze(0)
    world = dist.get_world_size(group)
    local = local.contiguous()
    out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local, group=group)          # rank-major concatenation along dim 0
    return out.view((world,) + tuple(local.shape))


def exchange_nccl_id(ctx: MatcherContext, group=None) -> tuple[bytes | None, int, int]:
    """(unique id, rank, world) for orbm_shard_create: rank 0 draws the id (ncclGetUniqueId through the C ABI) and
    torch.distributed broadcasts its 128 bytes."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None, 0, 1
    from ._lib import check
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [None]
    if rank == 0:
        buf = (C.c_uint8 * 128)()
        check(ctx._L.orbm_nccl_unique_id(buf), "orbm_nccl_unique_id")
        box[0] = bytes(buf)
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return box[0], rank, world


class ShardedDatabase:
    """This rank's shard of a descriptor database, resident in HBM, plus the exact distributed 2-NN query
    (orbm_shard_* of include/orbm.h)."""

    def __init__(self, shard_rows, idx_base: int, device: int = 0, context: MatcherContext | None = None, group=None):
        """shard_rows: (n, 32) uint8 numpy array or CUDA tensor holding global rows [idx_base, idx_base + n).
        COLLECTIVE over the process group when it has more than one rank (ncclCommInitRank)."""
        import torch
        from ._lib import check
        self.torch = torch
        self.device = torch.device("cuda", device)
        if isinstance(shard_rows, np.ndarray):
            shard_rows = torch.from_numpy(np.ascontiguousarray(shard_rows, np.uint8).reshape(-1, 32))
        self.db = shard_rows.to(self.device).contiguous()
        assert self.db.dtype == torch.uint8 and self.db.shape[1] == 32
        self.idx_base = int(idx_base)
        self.ctx = context or MatcherContext(device)
        self.stream = torch.cuda.ExternalStream(self.ctx._L.orbm_stream(self.ctx.h), device=self.device)
        uid, self.rank, self.world = exchange_nccl_id(self.ctx, group)
        h = C.c_void_p()
        idbuf = (C.c_uint8 * 128).from_buffer_copy(uid) if uid is not None else None
        check(self.ctx._L.orbm_shard_create(C.byref(h), self.ctx.h, idbuf, self.rank, self.world), "orbm_shard_create")
        self.h = h
        torch.cuda.synchronize(self.device)              # the rows were copied on torch's stream
        check(self.ctx._L.orbm_shard_set_rows(self.h, self.db.data_ptr(), self.db.shape[0], self.idx_base), "orbm_shard_set_rows")
        self._merged = None

    def close(self):
        if getattr(self, "h", None):
            self.ctx._L.orbm_shard_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def query_device(self, d_queries):
        """d_queries: (nq, 32) uint8 CUDA tensor, identical on every rank.  Returns the merged (nq, 8) uint8 CUDA
        tensor (orbm_top2 records) - enqueued on the matcher stream, not synchronised."""
        from ._lib import check
        t = self.torch
        nq = d_queries.shape[0]
        if self._merged is None or self._merged.shape[0] != nq:
            self._merged = t.empty((nq, 8), dtype=t.uint8, device=self.device)
        self.stream.wait_stream(t.cuda.current_stream(self.device))      # queries produced on torch's stream
        check(self.ctx._L.orbm_shard_bf2nn_device(self.h, d_queries.data_ptr(), nq, self._merged.data_ptr()), "orbm_shard_bf2nn_device")
        return self._merged

    def query(self, queries: np.ndarray) -> np.ndarray:
        """Host-buffer call: queries (nq, 32) uint8 -> TOP2_DTYPE array with global row indices."""
        from ._lib import check
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
        out = np.zeros(len(q), TOP2_DTYPE)
        check(self.ctx._L.orbm_shard_bf2nn(self.h, q.ctypes.data, len(q), out.ctypes.data), "orbm_shard_bf2nn")
        return out
