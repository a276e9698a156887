"""Row-sharded store over the GPUs of one box: one process per GPU, one exchange step.

The reference keeps every embedding in a single PostgreSQL table
(assistant/storage/models.py:19-58); here rows are sharded contiguously over the ranks,
shard ``g`` holding rows ``[g*N/G, (g+1)*N/G)`` with ids = global row index.  A search runs the
exact local top-k on every GPU (``ChunkStore``), exchanges the ``(Q, k)`` score and id lists
(``Q*k*12`` bytes per rank) and merges them: global top-k is a subset of the union of local top-k
lists, so the result equals the single-table ``ORDER BY distance LIMIT n``
(assistant/rag/services/search_service.py:192-194).

Two transports for the exchange:

* ``"p2p"`` (default on NCCL/CUDA when torch symmetric memory can be set up): every rank's search
  writes its lists straight into a peer-mapped staging block and ONE kernel
  (``dab_exchange_merge_topk``) publishes, waits for the peers, pulls their lists over NVLink and
  merges - no NCCL call on the data path; the ranks' search verdicts are OR-ed in the same kernel so
  that every rank takes the same retry decision.
* ``"nccl"``: ``all_gather_into_tensor`` of scores and ids, then ``dab_merge_topk``.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Tuple

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


class PeerExchange:
    """Peer-mapped staging block + the fused publish/wait/gather/merge kernel."""

    def __init__(self, device: torch.device, group=None, max_q: int = 256, max_k: int = 256):
        import torch.distributed._symmetric_memory as symm
        from ._lib import lib
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.max_q, self.max_k = int(max_q), int(max_k)
        self.device = device
        nbytes = int(lib().dab_exchange_bytes(self.world, self.max_q, self.max_k))
        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, self.group)
        torch.cuda.synchronize(device)
        self.hdl.barrier()                       # every rank's block is zeroed before anyone publishes
        self.bases_dev = int(self.hdl.buffer_ptrs_dev)
        self.epoch = 0
        self.status_dev = torch.zeros(2, dtype=torch.int32, device=device)
        self.status_host = torch.zeros(2, dtype=torch.int32).pin_memory()

    def fits(self, Q: int, k: int) -> bool:
        return Q <= self.max_q and k <= self.max_k and self.world * k <= 8192

    def next_slot(self, Q: int, k: int):
        """Tensors viewing the slot this rank's next search must write (scores [Q,k] f32, ids [Q,k] i64)."""
        from ._lib import check, lib
        self.epoch += 1
        ps, pi = ctypes.c_void_p(), ctypes.c_void_p()
        check(lib().dab_exchange_slot(self.buf.data_ptr(), self.world, self.max_q, self.max_k, self.epoch,
                                      ctypes.byref(ps), ctypes.byref(pi)))
        off_s, off_i = ps.value - self.buf.data_ptr(), pi.value - self.buf.data_ptr()
        s = self.buf[off_s: off_s + Q * k * 4].view(torch.float32).view(Q, k)
        i = self.buf[off_i: off_i + Q * k * 8].view(torch.int64).view(Q, k)
        return s, i

    def _packed(self, Q: int, k: int):
        """One device block [scores | ids | verdict] and its pinned twin, so that a checked search costs ONE
        device-to-host copy and ONE stream synchronisation."""
        key = (Q, k)
        if getattr(self, "_packed_key", None) != key:
            off_i = (Q * k * 4 + 15) // 16 * 16
            off_st = off_i + Q * k * 8
            total = off_st + 16
            self._packed_dev = torch.empty(total, dtype=torch.uint8, device=self.device)
            self._packed_host = torch.empty(total, dtype=torch.uint8).pin_memory()
            self._packed_off = (off_i, off_st, total)
            self._packed_key = key
        return self._packed_dev, self._packed_host, self._packed_off

    def merge(self, Q: int, k: int, local_status_ptr: Optional[int], packed: bool = False):
        """Launch the exchange kernel for the slot handed out last; returns merged (scores, ids) and
        leaves the OR-ed verdict in ``status_dev`` (``packed``: everything lands in one device block whose
        pinned copy ``fetch_packed`` reads back with a single transfer)."""
        from ._lib import check, lib
        if packed:
            dev_blk, _, (off_i, off_st, _) = self._packed(Q, k)
            out_s = dev_blk[:Q * k * 4].view(torch.float32).view(Q, k)
            out_i = dev_blk[off_i:off_i + Q * k * 8].view(torch.int64).view(Q, k)
            st_ptr = dev_blk.data_ptr() + off_st
        else:
            out_s = torch.empty((Q, k), dtype=torch.float32, device=self.device)
            out_i = torch.empty((Q, k), dtype=torch.int64, device=self.device)
            st_ptr = self.status_dev.data_ptr()
        with torch.cuda.device(self.device):
            check(lib().dab_exchange_merge_topk(self.bases_dev, self.rank, self.world, self.max_q, self.max_k, Q, k,
                                                self.epoch, local_status_ptr, out_s.data_ptr(), out_i.data_ptr(),
                                                st_ptr, torch.cuda.current_stream(self.device).cuda_stream))
        return out_s, out_i

    def fetch_packed(self, Q: int, k: int):
        """After ``merge(packed=True)``: one D2H copy + one synchronise.  Returns host views
        (scores [Q,k], ids [Q,k], incomplete, inexact)."""
        dev_blk, host_blk, (off_i, off_st, _) = self._packed(Q, k)
        host_blk.copy_(dev_blk, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        s = host_blk[:Q * k * 4].view(torch.float32).view(Q, k)
        i = host_blk[off_i:off_i + Q * k * 8].view(torch.int64).view(Q, k)
        st = host_blk[off_st:off_st + 8].view(torch.int32)
        return s, i, bool(st[0]), bool(st[1])

    def verdict(self) -> Tuple[bool, bool]:
        """(incomplete, inexact) of the last exchange, identical on every rank.  Synchronises."""
        self.status_host.copy_(self.status_dev, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return bool(self.status_host[0]), bool(self.status_host[1])


class ShardedStore:
    """This rank's shard of a corpus of ``total_rows`` rows + the merged search."""

    def __init__(self, total_rows: int, dim: int, dtype: torch.dtype = torch.bfloat16, device=None, group=None,
                 exchange: str = "auto", max_q: int = 256, max_k: int = 256):
        from . import ops
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.total_rows = int(total_rows)
        self.lo, self.hi = shard_bounds(self.total_rows, self.world, self.rank)
        self.local = ops.ChunkStore(max(1, self.hi - self.lo), dim, dtype, device, id_base=self.lo)
        self.peer: Optional[PeerExchange] = None
        if exchange not in ("auto", "p2p", "nccl"):
            raise ValueError("exchange must be 'auto', 'p2p' or 'nccl'")
        if self.world > 1 and exchange in ("auto", "p2p"):
            try:
                self.peer = PeerExchange(self.local.device, group, max_q, max_k)
            except Exception:
                if exchange == "p2p":
                    raise
                self.peer = None                 # symmetric memory unavailable: NCCL all-gather instead

    def append_local(self, rows: torch.Tensor) -> None:
        """Rows of THIS shard, in global row order."""
        self.local.append(rows)

    def ingest(self, hidden: torch.Tensor, mask: Optional[torch.Tensor] = None, ids: Optional[torch.Tensor] = None) -> int:
        """The write side of the sharded store, data-parallel and without any collective: this rank pools and
        normalises the encoder output of ITS chunks (``dab_pool_l2norm``) and appends the rows to its own shard -
        the chunk -> GPU map is the row -> shard map (``shard_bounds``), so an embedding is born on the GPU that
        will scan it.  What ``embedder.embeddings(texts)`` + ``bulk_update(['embedding'])`` do through the
        embedding service and PostgreSQL (assistant/processing/documents/steps/embeddings.py:27-39,57-69).
        ``hidden`` [B,T,D] fp32/bf16 on this rank's device, ``mask`` [B,T].  Returns the rows appended."""
        if ids is None:
            return self.local.ingest(hidden, mask)      # one kernel: the pool's outputs are the shard's own rows
        from . import ops
        if self.local.dtype == torch.bfloat16:
            _, rows = ops.pool_l2norm(hidden, mask, want_bf16=True)
        else:
            rows = ops.pool_l2norm(hidden, mask)
        self.local.append(rows, ids)
        return int(rows.shape[0])

    def _search_p2p(self, queries: torch.Tensor, k: int, allow_bitmap):
        from ._lib import DabError, lib
        if queries.dim() == 1:
            queries = queries[None, :]
        Q = queries.shape[0]
        spare, path = -1, -1
        kept = None
        while True:
            slot = self.peer.next_slot(Q, k)
            self.local.search_async(queries, k, allow_bitmap, out=slot, path=path, spare_rounds=spare)
            s, i = self.peer.merge(Q, k, lib().dab_store_status_dev(self.local._h))
            incomplete, inexact = self.peer.verdict()          # same answer on every rank
            if inexact and path != 1:
                # Some rank's tensor pass could not keep every row inside the rescoring margin for SOME
                # queries; that rank's lists for them are poisoned (NaN in front).  The union over the ranks
                # (every rank is here: the verdict is OR-ed) is repeated on the exact CUDA-core path by all
                # ranks together; the other queries' merged answers are final.
                path = 1
                bad = torch.isnan(slot[0][:, 0]).to(torch.int32)
                dist.all_reduce(bad, op=dist.ReduceOp.MAX, group=self.group)
                bad = bad.bool()
                n_bad = int(bad.sum())
                if 0 < n_bad < Q:
                    kept = (s, i, bad)
                    queries = queries[bad].contiguous()
                    Q = n_bad
                continue
            if not incomplete:
                if kept is not None:
                    kept[0][kept[2]] = s
                    kept[1][kept[2]] = i
                    return kept[0], kept[1]
                return s, i
            spare = 4 if spare <= 0 else spare * 2
            if spare > 64:
                raise DabError(5, "sharded search did not complete within 64 spare rounds")

    def search(self, queries: torch.Tensor, k: int, allow_bitmap: Optional[torch.Tensor] = None):
        """Replicated queries in, identical merged ``(scores [Q,k], ids [Q,k])`` out on every rank."""
        from . import ops
        if self.world == 1:
            return self.local.search(queries, k, allow_bitmap)
        Q = queries.shape[0] if queries.dim() == 2 else 1
        if self.peer is not None and self.peer.fits(Q, k):
            return self._search_p2p(queries, k, allow_bitmap)
        s, i = self.local.search(queries, k, allow_bitmap)
        gs, gi = gather_topk(s, i, self.group)
        return ops.merge_topk(gs, gi)

    def search_host(self, queries_host, k: int, allow_bitmap: Optional[torch.Tensor] = None, out=None):
        """The reference-facing form on every rank: HOST queries in (a float32 tensor, ideally pinned; or anything
        ``torch.as_tensor`` accepts, e.g. the Python list ``_objects_embedding_search`` receives,
        assistant/rag/services/search_service.py:186), HOST ``(scores [Q,k], ids [Q,k])`` out.  On the peer-memory
        transport the merged lists and the OR-ed verdict come back in one device-to-host copy and one
        synchronise; the rare verdicts (inexact / incomplete) fall back to the checked device search."""
        q = torch.as_tensor(queries_host, dtype=torch.float32)
        if q.dim() == 1:
            q = q[None, :]
        Q = q.shape[0]
        qd = q.to(self.local.device, non_blocking=True)
        if self.world > 1 and self.peer is not None and self.peer.fits(Q, k):
            from ._lib import lib
            slot = self.peer.next_slot(Q, k)
            self.local.search_async(qd, k, allow_bitmap, out=slot)
            self.peer.merge(Q, k, lib().dab_store_status_dev(self.local._h), packed=True)
            s, i, incomplete, inexact = self.peer.fetch_packed(Q, k)
            if not (incomplete or inexact):
                if out is not None:
                    out[0].copy_(s)
                    out[1].copy_(i)
                    return out
                return s.clone(), i.clone()
        s, i = self.search(qd, k, allow_bitmap)
        if out is None:
            out = (torch.empty(s.shape, dtype=s.dtype).pin_memory(), torch.empty(i.shape, dtype=i.dtype).pin_memory())
        out[0].copy_(s, non_blocking=True)
        out[1].copy_(i, non_blocking=True)
        torch.cuda.current_stream(self.local.device).synchronize()
        return out

    def close(self) -> None:
        self.local.close()


class LocalShardedStore:
    """The row-sharded store driven by ONE process (``dab_sharded_*``): what a Django / Celery / FastAPI worker can
    use without ``torchrun`` - the reference searches from an asgiref worker thread of an ordinary web process
    (assistant/rag/services/search_service.py:191).  ``devices`` lists the GPU of every shard; rows
    ``[g*N/G, (g+1)*N/G)`` live on ``devices[g]`` with ids = global row index.  ``search_host`` takes host queries
    (NumPy / list) and returns NumPy ``(scores [Q,k], ids [Q,k])``: the shards scan concurrently, the library's own
    exchange kernel merges over NVLink peer memory, one copy brings the answer back."""

    def __init__(self, total_rows: int, dim: int, dtype: torch.dtype = torch.bfloat16, devices: Optional[Sequence[int]] = None,
                 max_q: int = 256, max_k: int = 256):
        from . import _lib, ops
        from ._lib import check, lib
        if devices is None:
            devices = list(range(torch.cuda.device_count()))
        self.devices = [int(d) for d in devices]
        self.total_rows, self.dim, self.dtype = int(total_rows), int(dim), dtype
        self._h = ctypes.c_void_p()
        arr = (ctypes.c_int * len(self.devices))(*self.devices)
        check(lib().dab_sharded_create(ctypes.byref(self._h), arr, len(self.devices), self.total_rows, self.dim,
                                       ops._DTYPES[dtype], int(max_q), int(max_k)))
        self.shards = []
        self.bounds = []
        self.streams = []
        for g in range(len(self.devices)):
            dev, lo, hi, st = ctypes.c_int(), ctypes.c_int64(), ctypes.c_int64(), ctypes.c_void_p()
            h = lib().dab_sharded_shard(self._h, g, ctypes.byref(dev), ctypes.byref(lo), ctypes.byref(hi), ctypes.byref(st))
            self.shards.append(ops.ChunkStore._view(h, torch.device("cuda", dev.value), self.dim, dtype))
            self.bounds.append((lo.value, hi.value))
            self.streams.append(st.value)

    def append(self, g: int, rows: torch.Tensor) -> None:
        """Rows of shard ``g`` in global row order (a tensor on ``devices[g]``)."""
        self.shards[g].append(rows)

    def ingest(self, g: int, hidden: torch.Tensor, mask: Optional[torch.Tensor] = None) -> int:
        """Pool + normalise + append on the GPU that owns the rows (see ``ShardedStore.ingest``)."""
        return self.shards[g].ingest(hidden, mask)

    def search_host(self, queries, k: int):
        import numpy as np
        from ._lib import DabError, check, lib
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        if q.ndim != 2 or q.shape[1] != self.dim:
            raise DabError(3, f"different vector dimensions {q.shape[-1]} and {self.dim}")
        Q = q.shape[0]
        scores = np.empty((Q, k), dtype=np.float32)
        ids = np.empty((Q, k), dtype=np.int64)
        check(lib().dab_sharded_search_host(self._h, q.ctypes.data, Q, int(k), scores.ctypes.data, ids.ctypes.data))
        return scores, ids

    def close(self) -> None:
        from ._lib import lib
        if self._h.value:
            for s in self.shards:
                s.close()
            lib().dab_sharded_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass
