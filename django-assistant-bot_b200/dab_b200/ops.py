"""Torch-facing wrappers over the C-ABI (tensors in, tensors out; torch only supplies device
memory and streams).  Everything here requires CUDA tensors and raises otherwise."""
from __future__ import annotations

import threading
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import DabError, check, lib

_DTYPES = {torch.float32: _lib.DAB_F32, torch.bfloat16: _lib.DAB_BF16}
_MASK_DTYPES = {torch.uint8: _lib.DAB_MASK_U8, torch.bool: _lib.DAB_MASK_U8, torch.int32: _lib.DAB_MASK_I32,
                torch.int64: _lib.DAB_MASK_I64}


def _stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _need_cuda(t: torch.Tensor, what: str) -> None:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise DabError(_lib.DAB_E_INVALID, f"{what} must be a CUDA tensor (the B200 path has no CPU fallback)")


def pool_l2norm(hidden: torch.Tensor, mask: Optional[torch.Tensor] = None, *, want_bf16: bool = False,
                want_inv_norm: bool = False):
    """Masked mean over tokens then L2 normalise: the batched form of
    ``outputs.last_hidden_state.mean(dim=1)`` (assistant/ai/embedders/transformers.py:25).

    hidden [B,T,D] fp32/bf16, mask [B,T] (bool/uint8/int32/int64; the HF attention_mask) or None.
    Returns ``out`` [B,D] fp32, plus the bf16 copy and/or the applied 1/|mean| when asked.
    """
    _need_cuda(hidden, "hidden")
    if hidden.dim() != 3:
        raise DabError(_lib.DAB_E_INVALID, "hidden must be [B,T,D]")
    if hidden.dtype not in _DTYPES:
        raise DabError(_lib.DAB_E_INVALID, f"hidden dtype {hidden.dtype} unsupported (fp32 or bf16)")
    hidden = hidden.contiguous()
    B, T, D = hidden.shape
    mptr, mdt = None, 0
    if mask is not None:
        _need_cuda(mask, "mask")
        if tuple(mask.shape) != (B, T):
            raise DabError(_lib.DAB_E_INVALID, f"mask shape {tuple(mask.shape)} != {(B, T)}")
        if mask.dtype not in _MASK_DTYPES:
            raise DabError(_lib.DAB_E_INVALID, f"mask dtype {mask.dtype} unsupported")
        mask = mask.contiguous()
        mptr, mdt = mask.data_ptr(), _MASK_DTYPES[mask.dtype]
    dev = hidden.device
    out = torch.empty((B, D), dtype=torch.float32, device=dev)
    out_bf16 = torch.empty((B, D), dtype=torch.bfloat16, device=dev) if want_bf16 else None
    out_inv = torch.empty((B,), dtype=torch.float32, device=dev) if want_inv_norm else None
    with torch.cuda.device(dev):
        check(lib().dab_pool_l2norm(hidden.data_ptr(), _DTYPES[hidden.dtype], mptr, mdt, out.data_ptr(),
                                    out_bf16.data_ptr() if want_bf16 else None,
                                    out_inv.data_ptr() if want_inv_norm else None, B, T, D, _stream(dev)))
    res = (out,)
    if want_bf16:
        res += (out_bf16,)
    if want_inv_norm:
        res += (out_inv,)
    return res[0] if len(res) == 1 else res


def pack_allow_bitmap(allow: torch.Tensor) -> torch.Tensor:
    """bool [n] row filter -> the uint32 bitmap dab_store_search takes (bit r of word r//32)."""
    _need_cuda(allow, "allow")
    a = allow.to(torch.bool).flatten()
    n = a.numel()
    words = (n + 31) // 32
    pad = words * 32 - n
    if pad:
        a = torch.cat([a, torch.zeros(pad, dtype=torch.bool, device=a.device)])
    w = (a.view(words, 32).to(torch.int64) << torch.arange(32, device=a.device, dtype=torch.int64)).sum(dim=1)
    # same 32 bits, held in an int32 tensor (values >= 2^31 wrap to negative)
    return (w - ((w >> 31) << 32)).to(torch.int32)


class ChunkStore:
    """One device-resident shard of the chunk-embedding table
    (assistant/storage/models.py:19-29: ``embedding = VectorField(dimensions=768)``).

    ``search`` is the device half of ``_objects_embedding_search``
    (assistant/rag/services/search_service.py:185-196).
    """

    def __init__(self, capacity: int, dim: int, dtype: torch.dtype = torch.float32, device=None, id_base: int = 0):
        if not torch.cuda.is_available():
            raise DabError(_lib.DAB_E_CUDA, "no CUDA device: the B200 store has no CPU fallback")
        if dtype not in _DTYPES:
            raise DabError(_lib.DAB_E_INVALID, f"store dtype {dtype} unsupported (fp32 or bf16)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.dim = int(dim)
        self.dtype = dtype
        self._h = _lib.c_void_p()
        self._path = 0
        self._lock = threading.Lock()
        check(lib().dab_store_create(_lib.ctypes.byref(self._h), self.device.index, int(capacity), int(dim), _DTYPES[dtype]))
        if id_base:
            check(lib().dab_store_set_id_base(self._h, int(id_base)))

    @classmethod
    def _view(cls, handle: int, device: torch.device, dim: int, dtype: torch.dtype) -> "ChunkStore":
        """A non-owning wrapper around a dab_store that something else created and will destroy
        (the shards of ``dab_sharded_create``)."""
        self = cls.__new__(cls)
        self.device, self.dim, self.dtype = device, int(dim), dtype
        self._h = _lib.c_void_p(handle)
        self._path = 0
        self._lock = threading.Lock()
        self._borrowed = True
        return self

    # -- lifecycle
    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            if not getattr(self, "_borrowed", False):
                lib().dab_store_destroy(self._h)
            self._h = _lib.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def __len__(self) -> int:
        return int(lib().dab_store_size(self._h))

    @property
    def capacity(self) -> int:
        return int(lib().dab_store_capacity(self._h))

    def reset(self) -> None:
        check(lib().dab_store_reset(self._h))

    def set_path(self, path: str) -> None:
        code = {"auto": 0, "simt": 1, "tcgen05": 2}[path]
        check(lib().dab_store_set_path(self._h, code))
        self._path = code

    def set_spare_rounds(self, n: int) -> None:
        check(lib().dab_store_set_spare_rounds(self._h, int(n)))

    def describe_plan(self, Q: int, k: int) -> dict:
        """How a search of Q queries / top-k would run on the rows held now (path, tile, levels)."""
        c = _lib.c_int
        path, tile, levels, growth = c(), c(), c(), c()
        strides = (c * 8)()
        ref = _lib.ctypes.byref
        check(lib().dab_store_describe_plan(self._h, int(Q), int(k), ref(path), ref(tile), ref(levels), ref(growth), strides))
        return {"path": {1: "simt", 2: "tcgen05", 3: "tcgen05-cluster"}[path.value], "tile_rows": tile.value,
                "num_levels": levels.value, "growth": growth.value, "strides": list(strides)[:levels.value]}

    def set_profile(self, on: bool) -> None:
        check(lib().dab_store_set_profile(self._h, 1 if on else 0))

    def get_profile(self) -> dict:
        """Summed cudaEvent times since the last call: scan kernels and whole searches (ms)."""
        sm, sn = _lib.ctypes.c_float(), _lib.c_int()
        tm, tn = _lib.ctypes.c_float(), _lib.c_int()
        check(lib().dab_store_get_profile(self._h, _lib.ctypes.byref(sm), _lib.ctypes.byref(sn), _lib.ctypes.byref(tm),
                                          _lib.ctypes.byref(tn)))
        return {"scan_ms": sm.value, "scan_launches": sn.value, "search_ms": tm.value, "searches": tn.value}

    # -- write side
    def append(self, rows: torch.Tensor, ids: Optional[torch.Tensor] = None) -> None:
        _need_cuda(rows, "rows")
        if rows.dim() != 2 or rows.shape[1] != self.dim:
            raise DabError(_lib.DAB_E_DIM, f"different vector dimensions {rows.shape[-1]} and {self.dim}")
        if rows.dtype not in _DTYPES:
            raise DabError(_lib.DAB_E_INVALID, f"rows dtype {rows.dtype} unsupported")
        rows = rows.contiguous()
        iptr = None
        if ids is not None:
            _need_cuda(ids, "ids")
            ids = ids.to(torch.int64).contiguous()
            if ids.numel() != rows.shape[0]:
                raise DabError(_lib.DAB_E_INVALID, "ids and rows disagree on the row count")
            iptr = ids.data_ptr()
        with torch.cuda.device(self.device):
            check(lib().dab_store_append(self._h, rows.data_ptr(), _DTYPES[rows.dtype], iptr, rows.shape[0],
                                         _stream(self.device)))

    def ingest(self, hidden: torch.Tensor, mask: Optional[torch.Tensor] = None) -> int:
        """Pool + normalise ``hidden`` [B,T,dim] (fp32/bf16; ``mask`` [B,T] or None) and append the B rows, in one
        kernel whose outputs are the store's own rows, inverse norms and ids (``dab_store_ingest``).  Returns B."""
        _need_cuda(hidden, "hidden")
        if hidden.dim() != 3 or hidden.shape[2] != self.dim:
            raise DabError(_lib.DAB_E_DIM, f"hidden must be [B,T,{self.dim}]")
        if hidden.dtype not in _DTYPES:
            raise DabError(_lib.DAB_E_INVALID, f"hidden dtype {hidden.dtype} unsupported (fp32 or bf16)")
        hidden = hidden.contiguous()
        B, T, _ = hidden.shape
        mptr, mdt = None, 0
        if mask is not None:
            _need_cuda(mask, "mask")
            if tuple(mask.shape) != (B, T) or mask.dtype not in _MASK_DTYPES:
                raise DabError(_lib.DAB_E_INVALID, "mask must be [B,T] bool/uint8/int32/int64")
            mask = mask.contiguous()
            mptr, mdt = mask.data_ptr(), _MASK_DTYPES[mask.dtype]
        with torch.cuda.device(self.device):
            check(lib().dab_store_ingest(self._h, hidden.data_ptr(), _DTYPES[hidden.dtype], mptr, mdt, B, T, _stream(self.device)))
        return int(B)

    def delete_rows(self, rows: torch.Tensor) -> None:
        """Tombstone store rows (row indices, not ids): they are never returned again."""
        _need_cuda(rows, "rows")
        rows = rows.to(torch.int64).contiguous()
        with torch.cuda.device(self.device):
            check(lib().dab_store_delete_rows(self._h, rows.data_ptr(), rows.numel(), _stream(self.device)))

    def compact(self) -> int:
        """Physically drop the tombstoned rows (survivors keep their order; implicit ids are renumbered
        to ``id_base + row``).  Synchronous; returns how many rows were removed."""
        removed = _lib.c_int64(0)
        with self._lock, torch.cuda.device(self.device):
            check(lib().dab_store_compact(self._h, _lib.ctypes.byref(removed), _stream(self.device)))
        return int(removed.value)

    def export(self, row0: int, n: int):
        """(rows [n,dim] in the store's dtype, inv_norm [n] fp32, ids [n] int64) copied out of the store."""
        rows = torch.empty((n, self.dim), dtype=self.dtype, device=self.device)
        inv = torch.empty((n,), dtype=torch.float32, device=self.device)
        ids = torch.empty((n,), dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            check(lib().dab_store_export(self._h, int(row0), int(n), rows.data_ptr(), inv.data_ptr(), ids.data_ptr(),
                                         _stream(self.device)))
        return rows, inv, ids

    # -- queryset filters on the device
    def set_column(self, column: int, values: torch.Tensor, row0: int = 0) -> None:
        """int64 filter column ``column`` for store rows ``[row0, row0 + len(values))`` (document id, bot id, a
        status code ...): what the WHERE clauses of the reference's querysets test, kept in HBM beside the rows."""
        _need_cuda(values, "values")
        values = values.to(torch.int64).contiguous()
        with torch.cuda.device(self.device):
            check(lib().dab_store_set_column(self._h, int(column), values.data_ptr(), int(row0), values.numel(),
                                             _stream(self.device)))

    def build_allow(self, predicates) -> torch.Tensor:
        """The allow bitmap of a conjunction of predicates, built by one kernel on the device.
        ``predicates``: up to 8 tuples ``(column, op, operand)`` - column index or ``_lib.DAB_COL_ID``; ``op`` one of
        ``_lib.DAB_OP_*``; operand an int, or for IN / NOT_IN an ASCENDING int64 CUDA tensor.  Tombstoned rows are
        never allowed.  Returns the int32 tensor ``search`` takes as ``allow_bitmap``."""
        preds = list(predicates)
        if len(preds) > _lib.DAB_MAX_PREDICATES:
            raise DabError(_lib.DAB_E_INVALID, "at most 8 predicates per call")
        arr = (_lib.Predicate * max(1, len(preds)))()
        keep = []
        for i, (col, op, operand) in enumerate(preds):
            arr[i].column, arr[i].op = int(col), int(op)
            if op in (_lib.DAB_OP_IN, _lib.DAB_OP_NOT_IN):
                _need_cuda(operand, "IN operand")
                t = operand.to(torch.int64).contiguous()
                keep.append(t)
                arr[i].set_dev, arr[i].set_n, arr[i].value = t.data_ptr() if t.numel() else None, t.numel(), 0
            else:
                arr[i].value, arr[i].set_dev, arr[i].set_n = int(operand), None, 0
        words = (len(self) + 31) // 32
        bitmap = torch.empty(max(words, 1), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib().dab_store_build_allow(self._h, arr, len(preds), bitmap.data_ptr(), _stream(self.device)))
        # `keep` (the IN sets) must outlive the kernel: same stream as the caller's next use, so returning is safe
        bitmap._dab_keep = keep
        return bitmap

    def allow_from_rows(self, rows: torch.Tensor) -> torch.Tensor:
        """The allow bitmap with one bit per listed STORE ROW (int64 CUDA tensor): a device scatter, O(len(rows))."""
        _need_cuda(rows, "rows")
        rows = rows.to(torch.int64).contiguous()
        bitmap = torch.empty(max((len(self) + 31) // 32, 1), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            check(lib().dab_store_allow_from_rows(self._h, rows.data_ptr() if rows.numel() else None, rows.numel(),
                                                  bitmap.data_ptr(), _stream(self.device)))
        return bitmap

    # -- read side
    def search_async(self, queries: torch.Tensor, k: int, allow_bitmap: Optional[torch.Tensor] = None,
                     out: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, *, path: int = -1, spare_rounds: int = -1):
        """Enqueue the search on the current stream; returns device tensors (scores [Q,k] fp32
        descending, ids [Q,k] int64).  No host synchronisation.  See ``search`` for the checked form.
        ``path`` / ``spare_rounds`` override the store's settings for this call only (-1 = keep)."""
        _need_cuda(queries, "queries")
        if queries.dim() == 1:
            queries = queries[None, :]
        if queries.shape[-1] != self.dim:
            raise DabError(_lib.DAB_E_DIM, f"different vector dimensions {queries.shape[-1]} and {self.dim}")
        queries = queries.to(torch.float32).contiguous()
        Q = queries.shape[0]
        if out is None:
            scores = torch.empty((Q, k), dtype=torch.float32, device=self.device)
            ids = torch.empty((Q, k), dtype=torch.int64, device=self.device)
        else:
            scores, ids = out
        aptr = None
        if allow_bitmap is not None:
            _need_cuda(allow_bitmap, "allow_bitmap")
            if allow_bitmap.numel() * 32 < len(self):
                raise DabError(_lib.DAB_E_INVALID, "allow bitmap shorter than the store")
            aptr = allow_bitmap.data_ptr()
        with torch.cuda.device(self.device):
            check(lib().dab_store_search_ex(self._h, queries.data_ptr(), Q, int(k), aptr, scores.data_ptr(), ids.data_ptr(),
                                            int(path), int(spare_rounds), _stream(self.device)))
        return scores, ids

    def search_host(self, queries, k: int, allow_bitmap: Optional[torch.Tensor] = None):
        """Host buffers in, host buffers out: ``queries`` is a float32 NumPy array (or anything
        ``np.asarray`` accepts, e.g. the Python list the reference passes) of shape [Q, dim] or [dim];
        returns NumPy ``(scores [Q,k] float32, ids [Q,k] int64)``.  One C call does the copies, the
        search, the verdict check and any retry."""
        import numpy as np
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        if q.ndim != 2 or q.shape[1] != self.dim:
            raise DabError(_lib.DAB_E_DIM, f"different vector dimensions {q.shape[-1]} and {self.dim}")
        Q = q.shape[0]
        scores = np.empty((Q, k), dtype=np.float32)
        ids = np.empty((Q, k), dtype=np.int64)
        aptr = None
        if allow_bitmap is not None:
            _need_cuda(allow_bitmap, "allow_bitmap")
            if allow_bitmap.numel() * 32 < len(self):
                raise DabError(_lib.DAB_E_INVALID, "allow bitmap shorter than the store")
            aptr = allow_bitmap.data_ptr()
        # (the GIL is released inside the C call; concurrent calls on one store are serialised by the library)
        with torch.cuda.device(self.device):
            check(lib().dab_store_search_host(self._h, q.ctypes.data, Q, int(k), aptr, scores.ctypes.data, ids.ctypes.data,
                                              _stream(self.device)))
        return scores, ids

    def search(self, queries: torch.Tensor, k: int, allow_bitmap: Optional[torch.Tensor] = None):
        """Checked search: runs ``search_async``, waits, and if a scan round overflowed its
        candidate buffer (adversarial row order) repeats with spare rounds until complete.
        Safe to call from several threads on one store: the verdict read belongs to this thread's
        own call, and the retries pass their path / spare rounds per call instead of changing the store."""
        spare, path = -1, -1
        kept = None
        while True:
            scores, ids = self.search_async(queries, k, allow_bitmap, path=path, spare_rounds=spare)
            with torch.cuda.device(self.device):
                rc = lib().dab_store_search_status(self._h, _stream(self.device))
            if rc == _lib.DAB_E_INEXACT and path != 1:
                # more rows inside the rescoring margin than the tensor pass can keep, for SOME queries:
                # those came back poisoned (NaN score in front), the others are final.  Answer the
                # poisoned ones on the exact CUDA-core path and splice them in.
                path = 1
                bad = torch.isnan(scores[:, 0])
                n_bad = int(bad.sum())
                if 0 < n_bad < scores.shape[0]:
                    kept = (scores, ids, bad)
                    queries = (queries if queries.dim() == 2 else queries[None, :])[bad].contiguous()
                continue
            if rc != _lib.DAB_E_INCOMPLETE:
                check(rc)
                if kept is not None:
                    kept[0][kept[2]] = scores
                    kept[1][kept[2]] = ids
                    return kept[0], kept[1]
                return scores, ids
            spare = 4 if spare <= 0 else spare * 2
            if spare > 64:
                raise DabError(_lib.DAB_E_INCOMPLETE, "search did not complete within 64 spare rounds")


def doc_aggregate(scores: torch.Tensor, rows: torch.Tensor, doc_of_row: torch.Tensor, max_scores_n: int, top_n: int):
    """Device form of the grouping in ``embedding_search`` (assistant/rag/services/search_service.py:133-150).
    ``scores``/``rows`` [Q,n] as returned by ``ChunkStore.search`` on a store with implicit ids;
    ``doc_of_row`` [N] int64.  Returns (doc_ids [Q,top_n] int64 with -1 padding, doc_scores [Q,top_n] fp32)."""
    for t, name in ((scores, "scores"), (rows, "rows"), (doc_of_row, "doc_of_row")):
        _need_cuda(t, name)
    if scores.dim() != 2 or scores.shape != rows.shape:
        raise DabError(_lib.DAB_E_INVALID, "scores and rows must both be [Q,n]")
    scores = scores.to(torch.float32).contiguous()
    rows = rows.to(torch.int64).contiguous()
    doc_of_row = doc_of_row.to(torch.int64).contiguous()
    Q, n = scores.shape
    dev = scores.device
    out_docs = torch.empty((Q, top_n), dtype=torch.int64, device=dev)
    out_scores = torch.empty((Q, top_n), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib().dab_doc_aggregate(scores.data_ptr(), rows.data_ptr(), doc_of_row.data_ptr(), doc_of_row.numel(), Q, n,
                                      int(max_scores_n), int(top_n), out_docs.data_ptr(), out_scores.data_ptr(), _stream(dev)))
    return out_docs, out_scores


def merge_topk(gathered_scores: torch.Tensor, gathered_ids: torch.Tensor):
    """k-way merge of per-shard results: [R,Q,k] -> [Q,k] under (score desc, id asc)."""
    _need_cuda(gathered_scores, "gathered_scores")
    _need_cuda(gathered_ids, "gathered_ids")
    if gathered_scores.dim() != 3 or gathered_scores.shape != gathered_ids.shape:
        raise DabError(_lib.DAB_E_INVALID, "gathered tensors must both be [R,Q,k]")
    gs = gathered_scores.to(torch.float32).contiguous()
    gi = gathered_ids.to(torch.int64).contiguous()
    R, Q, k = gs.shape
    dev = gs.device
    scores = torch.empty((Q, k), dtype=torch.float32, device=dev)
    ids = torch.empty((Q, k), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        check(lib().dab_merge_topk(gs.data_ptr(), gi.data_ptr(), R, Q, k, scores.data_ptr(), ids.data_ptr(), _stream(dev)))
    return scores, ids
