"""Multi-GPU modes of the front end: one process per GPU, torch.distributed for the plumbing.

* frames-parallel extraction: `shard_range` gives every rank a contiguous frame slice; no collective.
* brute-force Hamming 2-NN over a keyframe-descriptor database split across GPUs: every rank scans its
  rows with `orbm_bf2nn_device` (global row indices via idx_base), the per-rank (i1, d1, d2) triples are
  all-gathered (NCCL over NVLink: nq x 8 bytes per rank) and merged exactly by `orbm_merge_top2_device`.

The compute is always the CUDA library; there is no CPU path here.  `shard_range` and `gather_parts` are
backend-agnostic (they are exercised with gloo at world size 2 by the CPU tests).
"""
from __future__ import annotations

import numpy as np

from .matcher import TOP2_DTYPE, MatcherContext


def shard_range(n: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous, balanced split of n units (frames or database rows): first `n % world` ranks get one extra."""
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_parts(local, group=None):
    """all_gather of equally-shaped per-rank partial results -> tensor [world, ...] (NCCL for CUDA tensors,
    gloo for CPU tensors).  World size 1: a view, no collective."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local.unsqueeze(0)
    world = dist.get_world_size(group)
    local = local.contiguous()
    out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local, group=group)          # rank-major concatenation along dim 0
    return out.view((world,) + tuple(local.shape))


class ShardedDatabase:
    """This rank's shard of a descriptor database, resident in HBM, plus the exact distributed 2-NN query."""

    def __init__(self, shard_rows, idx_base: int, device: int = 0, context: MatcherContext | None = None, group=None):
        """shard_rows: (n, 32) uint8 numpy array or CUDA tensor holding global rows [idx_base, idx_base + n)."""
        import torch
        self.torch = torch
        self.device = torch.device("cuda", device)
        if isinstance(shard_rows, np.ndarray):
            shard_rows = torch.from_numpy(np.ascontiguousarray(shard_rows, np.uint8).reshape(-1, 32))
        self.db = shard_rows.to(self.device).contiguous()
        assert self.db.dtype == torch.uint8 and self.db.shape[1] == 32
        self.idx_base = int(idx_base)
        self.ctx = context or MatcherContext(device)
        self.group = group
        self.stream = torch.cuda.ExternalStream(self.ctx._L.orbm_stream(self.ctx.h), device=self.device)
        self._q = self._local = self._merged = None

    def _buffers(self, nq):
        t = self.torch
        if self._q is None or self._q.shape[0] != nq:
            self._q = t.empty((nq, 32), dtype=t.uint8, device=self.device)
            self._local = t.empty((nq, 8), dtype=t.uint8, device=self.device)
            self._merged = t.empty((nq, 8), dtype=t.uint8, device=self.device)

    def query_device(self, d_queries):
        """d_queries: (nq, 32) uint8 CUDA tensor, identical on every rank.  Returns the merged (nq, 8) uint8 CUDA
        tensor (orbm_top2 records) - enqueued, not synchronised."""
        from ._lib import check
        t = self.torch
        nq = d_queries.shape[0]
        self._buffers(nq)
        L, h = self.ctx._L, self.ctx.h
        cur = t.cuda.current_stream(self.device)
        self.stream.wait_stream(cur)                     # queries produced on torch's stream
        check(L.orbm_bf2nn_device(h, d_queries.data_ptr(), nq, self.db.data_ptr(), self.db.shape[0], self.idx_base,
                                  self._local.data_ptr()), "orbm_bf2nn_device")
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            cur.wait_stream(self.stream)
            parts = gather_parts(self._local, self.group)        # NCCL all-gather on torch's stream
            self.stream.wait_stream(cur)
            check(L.orbm_merge_top2_device(h, parts.data_ptr(), parts.shape[0], nq, self._merged.data_ptr()),
                  "orbm_merge_top2_device")
            self._parts = parts                                   # keep alive until the merge has run
            return self._merged
        return self._local

    def query(self, queries: np.ndarray) -> np.ndarray:
        """Host-buffer call: queries (nq, 32) uint8 -> TOP2_DTYPE array with global row indices."""
        t = self.torch
        q = t.from_numpy(np.ascontiguousarray(queries, np.uint8).reshape(-1, 32))
        with t.cuda.device(self.device):
            d_q = q.to(self.device, non_blocking=True)
            out = self.query_device(d_q)
            self.stream.synchronize()
            return out.cpu().numpy().view(TOP2_DTYPE).reshape(-1)
