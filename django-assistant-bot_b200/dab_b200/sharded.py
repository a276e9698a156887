"""Row-sharded store over the GPUs of one box: one process per GPU, one exchange step.

The reference keeps every embedding in a single PostgreSQL table
(assistant/storage/models.py:19-58); here rows are sharded contiguously over the ranks,
shard ``g`` holding rows ``[g*N/G, (g+1)*N/G)`` with ids = global row index.  A search runs the
exact local top-k on every GPU (``ChunkStore.search``), all-gathers the ``(Q, k)`` score and id
lists (``Q*k*12`` bytes per rank, NCCL over NVLink) and merges them with ``dab_merge_topk``:
global top-k is a subset of the union of local top-k lists, so the result equals the
single-table ``ORDER BY distance LIMIT n`` (assistant/rag/services/search_service.py:192-194).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(total_rows: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous row range of ``rank``; the same formula bench.py and the ingest side use, so a
    pooled chunk lands on the GPU that owns its row without any exchange."""
    return rank * total_rows // world, (rank + 1) * total_rows // world


def owner_of_row(row: int, total_rows: int, world: int) -> int:
    """Inverse of shard_bounds."""
    g = min(world - 1, (row * world) // max(total_rows, 1))
    while row < shard_bounds(total_rows, world, g)[0]:
        g -= 1
    while row >= shard_bounds(total_rows, world, g)[1]:
        g += 1
    return g


def gather_topk(scores: torch.Tensor, ids: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """All-gather the local ``[Q,k]`` lists into ``[R,Q,k]`` (rank-major) on every rank.
    Works on any backend (NCCL for CUDA tensors; gloo in the CPU tests of the host logic)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return scores[None].contiguous(), ids[None].contiguous()
    scores = scores.contiguous()
    ids = ids.contiguous()
    # output concatenated along dim 0 (the layout every backend accepts), viewed as [R,Q,k]
    Q = scores.shape[0]
    gs = torch.empty((world * Q,) + tuple(scores.shape[1:]), dtype=scores.dtype, device=scores.device)
    gi = torch.empty((world * Q,) + tuple(ids.shape[1:]), dtype=ids.dtype, device=ids.device)
    dist.all_gather_into_tensor(gs, scores, group=group)
    dist.all_gather_into_tensor(gi, ids, group=group)
    return gs.view((world,) + tuple(scores.shape)), gi.view((world,) + tuple(ids.shape))


class ShardedStore:
    """This rank's shard of a corpus of ``total_rows`` rows + the merged search."""

    def __init__(self, total_rows: int, dim: int, dtype: torch.dtype = torch.bfloat16, device=None, group=None):
        from . import ops
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.total_rows = int(total_rows)
        self.lo, self.hi = shard_bounds(self.total_rows, self.world, self.rank)
        self.local = ops.ChunkStore(max(1, self.hi - self.lo), dim, dtype, device, id_base=self.lo)

    def append_local(self, rows: torch.Tensor) -> None:
        """Rows of THIS shard, in global row order."""
        self.local.append(rows)

    def search(self, queries: torch.Tensor, k: int, allow_bitmap: Optional[torch.Tensor] = None):
        """Replicated queries in, identical merged ``(scores [Q,k], ids [Q,k])`` out on every rank."""
        from . import ops
        s, i = self.local.search(queries, k, allow_bitmap)
        if self.world == 1:
            return s, i
        gs, gi = gather_topk(s, i, self.group)
        return ops.merge_topk(gs, gi)

    def close(self) -> None:
        self.local.close()
