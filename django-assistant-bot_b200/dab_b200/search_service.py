"""RAG similarity search with the reference's API, executed on the B200 store.

Mirror of ``assistant/rag/services/search_service.py``: the same function names, arguments and
return shapes - ``_objects_embedding_search`` (:185-196), ``embedding_search_documents /
_questions / _sentences`` (:161-182), ``embedding_search`` (:111-152), ``get_embedding``
(:155-158), ``cosine_similarity`` / ``embeddings_similarity`` (:17-21, :32-35).  The one thing
that changes is where ``ORDER BY embedding <=> q LIMIT n`` runs: not in PostgreSQL+pgvector but in
``dab_store_search`` over a device-resident copy of the embedding column (``DeviceTable``).

Two kinds of ``qs`` are accepted:

* ``ChunkQuerySet`` - the stand-in used where Django is not installed (this repository's tests and
  benchmarks): a filterable view over a ``DeviceTable``;
* a Django ``QuerySet`` (duck-typed: ``.model``, ``.values_list('pk', flat=True)``,
  ``.model.objects.in_bulk``) - the table registered for ``(model label, field)`` is searched with
  the queryset's pks as the allow-set, and real model instances come back with ``.distance`` set.

There is no CPU path: a search without a CUDA device raises ``DabError``.
"""
from __future__ import annotations

import asyncio
import logging
import time
from collections import defaultdict
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops
from ._lib import DabError

logger = logging.getLogger(__name__)

EMBEDDING_DIM = 768  # assistant/storage/models.py:13,23 ("for RuBert")


# --------------------------------------------------------------------------- small NumPy helpers
def cosine_similarity(a, b):
    """search_service.py:17-21 - two-vector utility used by the emb_test command, not the scan."""
    a = np.asarray(a)
    b = np.asarray(b)
    return np.dot(a, b) / (np.linalg.norm(a) * np.linalg.norm(b))


def embeddings_similarity(a: List[float], b: List[float]) -> float:
    """search_service.py:32-35."""
    return cosine_similarity(np.array(a), np.array(b))


# --------------------------------------------------------------------------- the device table
class ChunkRow:
    """One search hit when no ORM is present; attribute-compatible with a
    ``BaseEmbeddingModel`` row (assistant/storage/models.py:19-29) plus ``.distance``."""
    __slots__ = ("id", "pk", "document_id", "text", "order", "distance")

    def __init__(self, pk, document_id, text, order, distance):
        self.id = self.pk = pk
        self.document_id = document_id
        self.text = text
        self.order = order
        self.distance = distance

    def __repr__(self):
        return f"ChunkRow(id={self.id}, document_id={self.document_id}, distance={self.distance:.6f})"


class _Grow:
    """A NumPy array that grows by doubling: the host-side columns of a table with 10^8 rows cannot be Python lists."""

    def __init__(self, dtype):
        self.a = np.empty(1024, dtype=dtype)
        self.n = 0

    def extend(self, values) -> None:
        v = np.asarray(values, dtype=self.a.dtype).reshape(-1)
        need = self.n + v.shape[0]
        if need > self.a.shape[0]:
            cap = self.a.shape[0]
            while cap < need:
                cap *= 2
            b = np.empty(cap, dtype=self.a.dtype)
            b[:self.n] = self.a[:self.n]
            self.a = b
        self.a[self.n:need] = v
        self.n = need

    def view(self) -> np.ndarray:
        return self.a[:self.n]

    def keep(self, idx: np.ndarray) -> None:
        kept = self.a[:self.n][idx]
        self.a, self.n = kept.copy(), kept.shape[0]
        if self.a.shape[0] == 0:
            self.a = np.empty(1024, dtype=kept.dtype)


class DeviceTable:
    """Device-resident copy of one embedding column: ``Sentence.embedding``,
    ``Question.embedding`` (assistant/storage/models.py:32-58) or ``Document.content_embedding``
    (:13).  Rows with a NULL embedding are simply not added (the column is nullable, :23).

    Host side keeps what the ORM would - pk, document_id, order, the columns the reference filters on
    (e.g. ``document__wiki__bot``; non-numeric values dictionary-encoded) - as growable NumPy int64 arrays aligned
    with the store rows (8 bytes per row and column, not a Python object), and the texts only if any were given.
    """

    def __init__(self, capacity: int, dim: int = EMBEDDING_DIM, dtype: torch.dtype = torch.float32, device=None):
        self.store = ops.ChunkStore(capacity, dim, dtype, device)
        self.dim = dim
        self.device = self.store.device
        self._pks = _Grow(np.int64)
        self._doc = _Grow(np.int64)
        self._order = _Grow(np.int64)
        self._dead = _Grow(np.bool_)                      # store rows turned into tombstones
        self._n_dead = 0
        self._text: Optional[List[Optional[str]]] = None  # allocated when the first text arrives
        self._meta: Dict[str, _Grow] = {}                 # filter columns: int64 values or dictionary codes
        self._codes: Dict[str, Dict[Any, int]] = {}       # value -> code of the dictionary-encoded columns
        self._decode: Dict[str, List[Any]] = {}           # code -> value
        self._pk_sorted: Optional[Tuple[np.ndarray, np.ndarray]] = None   # (sorted pks, their rows), built on demand
        self._doc_dev: Optional[torch.Tensor] = None      # device copy of the document_id column (lazy)
        # filter columns in HBM (dab_store_set_column): name -> [device column index, rows uploaded so far]
        self._dev_cols: Dict[str, List[int]] = {}
        self.documents: Dict[int, Any] = {}

    def __len__(self):
        return self._pks.n

    def _encode(self, name: str, values) -> np.ndarray:
        """Column values as int64: integers as they are, anything else through the column's dictionary."""
        arr = np.asarray(values)
        if arr.dtype.kind in "iub" and name not in self._codes:
            return arr.astype(np.int64)
        if len(self) and name in self._meta and name not in self._codes:
            raise DabError(1, f"filter column {name!r} held integers so far; it cannot switch to other values")
        codes, decode = self._codes.setdefault(name, {}), self._decode.setdefault(name, [])
        out = np.empty(arr.shape[0], dtype=np.int64)
        for i, v in enumerate(arr.tolist()):
            c = codes.get(v)
            if c is None:
                c = codes[v] = len(decode)
                decode.append(v)
            out[i] = c
        return out

    def add(self, pks: Sequence[int], embeddings, document_ids: Optional[Sequence[int]] = None,
            texts: Optional[Sequence[str]] = None, orders: Optional[Sequence[int]] = None, **meta) -> None:
        """The write side: what ``bulk_update(rows, fields=['embedding'])`` does to the column
        (assistant/processing/documents/steps/embeddings.py:35-39,65-69)."""
        n = len(pks)
        if n == 0:
            return
        emb = embeddings if isinstance(embeddings, torch.Tensor) else torch.as_tensor(np.asarray(embeddings, dtype=np.float32))
        if emb.dim() != 2 or emb.shape[0] != n:
            raise DabError(1, "embeddings must be [len(pks), dim]")
        if emb.shape[1] != self.dim:
            raise DabError(3, f"different vector dimensions {emb.shape[1]} and {self.dim}")
        known = set(self._meta)
        if len(self) and set(meta) != known:
            raise DabError(1, f"filter columns changed: {sorted(meta)} vs {sorted(known)}")
        encoded = {k: self._encode(k, v) for k, v in meta.items()}
        for k, v in encoded.items():
            if v.shape[0] != n:
                raise DabError(1, f"filter column {k!r} has {v.shape[0]} values for {n} rows")
        self.store.append(emb.to(self.device))
        before = len(self)
        self._pks.extend(np.asarray(pks, dtype=np.int64))
        self._doc.extend(np.zeros(n, np.int64) if document_ids is None else np.asarray(document_ids, dtype=np.int64))
        self._order.extend(np.arange(n, dtype=np.int64) if orders is None else np.asarray(orders, dtype=np.int64))
        self._dead.extend(np.zeros(n, np.bool_))
        if texts is not None and self._text is None and any(t is not None for t in texts):
            self._text = [None] * before
        if self._text is not None:
            self._text.extend(texts if texts is not None else [None] * n)
        for k, v in encoded.items():
            self._meta.setdefault(k, _Grow(np.int64)).extend(v)
        self._pk_sorted = None

    def load_queryset(self, qs, field: str = "embedding", columns: Optional[Dict[str, str]] = None, batch: int = 8192,
                      document_field: Optional[str] = "document_id", text_field: Optional[str] = "text",
                      order_field: Optional[str] = "order") -> int:
        """Bulk load from the database column: every row of ``qs`` whose ``field`` is not NULL, streamed in chunks
        (``qs.exclude(<field>__isnull=True).values_list(...).iterator(chunk_size=batch)``) and appended with its
        pk, document id, text, order and the filter columns asked for - ``columns`` maps the name used in
        ``ChunkQuerySet.filter`` to the ORM path it is read from, e.g.
        ``{"document__wiki__bot": "document__wiki__bot_id", "document__processing__status": "document__processing__status"}``
        (the filters of assistant/bot/services/context_service/steps/embeddings.py:26-29).  For
        ``Document.content_embedding`` (assistant/storage/models.py:13) pass ``document_field="pk"``,
        ``text_field=None``, ``order_field=None``.  The store is a cache of that column: this is how it is (re)built."""
        columns = dict(columns or {})
        names = ["pk", field] + [f for f in (document_field, text_field, order_field) if f] + list(columns.values())
        rows = qs.exclude(**{f"{field}__isnull": True}).values_list(*names)
        rows = rows.iterator(chunk_size=batch) if hasattr(rows, "iterator") else iter(rows)
        total = 0
        while True:
            chunk = []
            for r in rows:
                chunk.append(r)
                if len(chunk) == batch:
                    break
            if not chunk:
                break
            cols = list(zip(*chunk))
            pos = 2
            docs = texts = orders = None
            if document_field:
                docs, pos = cols[pos], pos + 1
            if text_field:
                texts, pos = cols[pos], pos + 1
            if order_field:
                orders, pos = cols[pos], pos + 1
            meta = {name: list(cols[pos + i]) for i, name in enumerate(columns)}
            emb = np.asarray([np.asarray(e, dtype=np.float32) for e in cols[1]], dtype=np.float32)
            self.add(list(cols[0]), emb, docs, texts, orders, **meta)
            total += len(chunk)
        return total

    def _rows_of_pks(self, pks: Iterable[int]) -> np.ndarray:
        """Store rows of the given pks (unknown and already deleted pks are skipped) through a sorted index."""
        want = np.unique(np.fromiter((int(p) for p in pks), dtype=np.int64))
        if self._pk_sorted is None:
            order = np.argsort(self._pks.view(), kind="stable")
            self._pk_sorted = (self._pks.view()[order], order)
        spk, rows = self._pk_sorted
        pos = np.searchsorted(spk, want)
        ok = (pos < spk.shape[0])
        ok[ok] &= spk[pos[ok]] == want[ok]
        found = rows[pos[ok]]
        return found[~self._dead.view()[found]]

    def delete(self, pks: Iterable[int]) -> int:
        """Remove rows by pk: what ``Document.objects.filter(...).delete()`` cascades to on finalisation
        (assistant/processing/tasks.py:67-74).  Rows become tombstones on the device (never returned
        again, excluded from every bitmap); returns how many rows were removed."""
        rows = self._rows_of_pks(pks)
        if rows.shape[0] == 0:
            return 0
        self.store.delete_rows(torch.from_numpy(rows.astype(np.int64)).to(self.device))
        self._dead.view()[rows] = True
        self._n_dead += int(rows.shape[0])
        return int(rows.shape[0])

    def compact(self) -> int:
        """Physically drop the tombstones (what VACUUM does for the table after the cascade deletes):
        the device store moves the surviving rows (and its filter columns) down in order and the host columns
        follow with the same survivors, so store row ``i`` is again row ``i`` of every column.  Returns rows removed."""
        if not self._n_dead:
            return 0
        for name in list(self._dev_cols):                # every row's values on the device before the rows move
            self._device_column(name)
        keep = np.flatnonzero(~self._dead.view())
        removed = self.store.compact()
        if removed != self._n_dead or len(self.store) != keep.shape[0]:
            raise DabError(1, f"compaction removed {removed} rows, {self._n_dead} were tombstoned")
        for col in (self._pks, self._doc, self._order, self._dead, *self._meta.values()):
            col.keep(keep)
        if self._text is not None:
            self._text = [self._text[i] for i in keep.tolist()]
        self._n_dead = 0
        self._pk_sorted = None
        self._doc_dev = None
        for name in self._dev_cols:                      # the HBM columns were compacted with the rows
            self._dev_cols[name][1] = len(self)
        return removed

    def live_mask(self) -> Optional[np.ndarray]:
        return ~self._dead.view() if self._n_dead else None

    def save(self, path: str) -> None:
        """Snapshot of the shard: rows in the store's dtype + the host columns, as one ``.npz`` of plain arrays
        (loading never unpickles anything).  The device store is a cache of the DB column (it can always be rebuilt
        from PostgreSQL); the snapshot only avoids re-reading and re-uploading 768 floats per row at start-up."""
        n = len(self)
        blocks = []
        for r0 in range(0, n, 1 << 18):
            rows, _, _ = self.store.export(r0, min(1 << 18, n - r0))
            blocks.append(rows.view(torch.int16).cpu().numpy() if rows.dtype == torch.bfloat16 else rows.cpu().numpy())
        data = np.concatenate(blocks, 0) if blocks else np.zeros((0, self.dim), np.float32)
        extra = {}
        for k, col in self._meta.items():
            extra[f"meta__{k}"] = col.view()
            if k in self._decode:
                values = np.asarray(self._decode[k])
                if values.dtype == object:
                    raise DabError(1, f"filter column {k!r} mixes value types; snapshots hold numbers or strings only")
                extra[f"dict__{k}"] = values
        texts = self._text if self._text is not None else [None] * n
        enc = [None if t is None else t.encode("utf-8") for t in texts]   # one UTF-8 blob + lengths (-1 = NULL)
        np.savez(path, rows=data, bf16=np.asarray(self.store.dtype == torch.bfloat16), dim=np.asarray(self.dim),
                 pks=self._pks.view(), docs=self._doc.view(), orders=self._order.view(),
                 text_blob=np.frombuffer(b"".join(e for e in enc if e is not None), dtype=np.uint8),
                 text_len=np.asarray([-1 if e is None else len(e) for e in enc], dtype=np.int64),
                 deleted=np.flatnonzero(self._dead.view()).astype(np.int64), **extra)

    @classmethod
    def load(cls, path: str, capacity: Optional[int] = None, device=None) -> "DeviceTable":
        z = np.load(path, allow_pickle=False)
        bf16 = bool(z["bf16"])
        n = z["rows"].shape[0]
        t = cls(max(capacity or n, 1), int(z["dim"]), torch.bfloat16 if bf16 else torch.float32, device)
        meta = {}
        for k in z.files:
            if k.startswith("meta__"):
                name = k[len("meta__"):]
                codes = z[k]
                meta[name] = z[f"dict__{name}"][codes] if f"dict__{name}" in z.files else codes
        blob, lens = z["text_blob"].tobytes(), z["text_len"]
        ends = np.cumsum(np.maximum(lens, 0))
        texts = [None if l < 0 else blob[e - l:e].decode("utf-8") for l, e in zip(lens.tolist(), ends.tolist())]
        for r0 in range(0, n, 1 << 18):
            r1 = min(n, r0 + (1 << 18))
            blk = torch.from_numpy(z["rows"][r0:r1])
            emb = blk.view(torch.bfloat16) if bf16 else blk
            t.add(z["pks"][r0:r1], emb, z["docs"][r0:r1], texts[r0:r1], z["orders"][r0:r1],
                  **{k: v[r0:r1].tolist() for k, v in meta.items()})
        dead = z["deleted"]
        if dead.shape[0]:
            t.delete(z["pks"][dead].tolist())
        return t

    def bitmap_from_mask(self, mask: np.ndarray) -> torch.Tensor:
        """Utility for callers that already hold a row mask (not used on the query path)."""
        return ops.pack_allow_bitmap(torch.from_numpy(np.ascontiguousarray(mask, dtype=np.bool_)).to(self.device))

    # -- queryset filters on the device -------------------------------------------------------------
    _ALIASES = {"id": "pk", "document__id": "document_id", "document": "document_id"}

    def _encode_operand(self, name: str, value) -> Optional[int]:
        """The int64 a filter value compares as; None = a value this column has never held."""
        value = getattr(value, "pk", value)              # model instances compare by pk, like the ORM
        if name in self._codes:
            return self._codes[name].get(value)
        return int(value)

    def _device_column(self, name: str) -> int:
        """Index of the HBM column behind ``name``, with every row added so far uploaded (incrementally)."""
        name = self._ALIASES.get(name, name)
        host = {"pk": self._pks, "document_id": self._doc, "order": self._order}.get(name, self._meta.get(name))
        if host is None:
            raise DabError(1, f"unknown filter column {name!r}; known: pk, document_id, order, {sorted(self._meta)}")
        if name not in self._dev_cols:
            if len(self._dev_cols) >= ops._lib.DAB_MAX_COLUMNS:
                raise DabError(1, f"more than {ops._lib.DAB_MAX_COLUMNS} filter columns in use")
            self._dev_cols[name] = [len(self._dev_cols), 0]
        idx, done = self._dev_cols[name]
        n = len(self)
        if done < n:
            self.store.set_column(idx, torch.from_numpy(np.ascontiguousarray(host.view()[done:n])).to(self.device), done)
            self._dev_cols[name][1] = n
        return idx

    _OPS = {"exact": 0, "lt": 2, "lte": 3, "gt": 4, "gte": 5, "in": 6}
    _NEG = {0: 1, 1: 0, 2: 5, 5: 2, 3: 4, 4: 3, 6: 7, 7: 6}

    def predicate(self, name: str, lookup: str, value, negate: bool = False):
        """``(column, op, operand)`` for ``ChunkStore.build_allow`` from a Django-style lookup."""
        name = self._ALIASES.get(name, name)
        col = self._device_column(name)
        op = self._OPS[lookup]
        coded = name in self._codes
        if coded and lookup not in ("exact", "in"):
            raise DabError(1, f"column {name!r} is not numeric: only exact / in lookups apply")
        if lookup == "in":
            vals = [self._encode_operand(name, v) for v in value]
            operand = torch.as_tensor(sorted(v for v in vals if v is not None), dtype=torch.int64).to(self.device)
        else:
            operand = self._encode_operand(name, value)
            if operand is None:                          # a value the column never held: matches nothing
                op, operand = 6, torch.zeros(0, dtype=torch.int64, device=self.device)
        return (col, self._NEG[op] if negate else op, operand)

    def build_bitmap(self, predicates) -> torch.Tensor:
        """AND of the predicates (and of "row is alive"), evaluated on the device; more than 8 predicates are
        built in groups and combined with a bitwise AND."""
        preds = list(predicates)
        step = ops._lib.DAB_MAX_PREDICATES
        bitmap = self.store.build_allow(preds[:step])
        for i in range(step, len(preds), step):
            bitmap = bitmap & self.store.build_allow(preds[i:i + step])
        return bitmap

    def bitmap_from_pks(self, pks: Iterable[int]) -> torch.Tensor:
        """The queryset as a pk list (what a Django QuerySet can give us): the pks are mapped to store rows through
        the sorted pk index (vectorised, O(len(pks) log N) on the host) and scattered into the bitmap on the device
        (``dab_store_allow_from_rows``, O(len(pks))) - nothing touches all N rows."""
        rows = self._rows_of_pks(pks)
        return self.store.allow_from_rows(torch.from_numpy(np.ascontiguousarray(rows, dtype=np.int64)).to(self.device))

    def search_rows(self, query_embeddings, n: int, bitmap: Optional[torch.Tensor] = None):
        """(distances [Q, n'] float64 ascending, store rows [Q, n'] int64); n' <= n."""
        q = np.asarray(query_embeddings, dtype=np.float32)
        if q.ndim == 1:
            q = q[None]
        if q.shape[-1] != self.dim:
            raise DabError(3, f"different vector dimensions {q.shape[-1]} and {self.dim}")
        if len(self) == 0 or n <= 0:
            return np.zeros((q.shape[0], 0)), np.zeros((q.shape[0], 0), dtype=np.int64)
        if int(n) > ops._lib.DAB_MAX_K:
            raise DabError(ops._lib.DAB_E_CAPACITY, f"n = {n} exceeds DAB_MAX_K = {ops._lib.DAB_MAX_K} rows per search "
                           "(the reference's largest n is 250)")
        k = int(n)
        scores, ids = self.store.search_host(q, k, bitmap)      # one C call: H2D, search, D2H, verdict, retries
        return 1.0 - scores.astype(np.float64), ids

    def search_documents(self, query_embedding, n: int, max_scores_n: int, top_n: int,
                         bitmap: Optional[torch.Tensor] = None) -> List[Tuple[int, float]]:
        """Device form of search_service.py:129-150: top-n rows, grouped per document, best ``top_n``
        ``(document_id, score)`` pairs."""
        q = torch.as_tensor(np.asarray(query_embedding, dtype=np.float32))
        if q.shape[-1] != self.dim:
            raise DabError(3, f"different vector dimensions {q.shape[-1]} and {self.dim}")
        if len(self) == 0:
            return []
        if self._doc_dev is None or self._doc_dev.numel() != len(self):
            self._doc_dev = torch.from_numpy(np.ascontiguousarray(self._doc.view())).to(self.device)
        if int(n) > ops._lib.DAB_MAX_K:
            raise DabError(ops._lib.DAB_E_CAPACITY, f"max_scores_n * top_n * 10 = {n} exceeds DAB_MAX_K = {ops._lib.DAB_MAX_K}")
        k = int(n)
        scores, rows = self.store.search(q.to(self.device), k, bitmap)
        docs, dscores = ops.doc_aggregate(scores, rows, self._doc_dev, max_scores_n, top_n)
        docs, dscores = docs[0].cpu().tolist(), dscores[0].cpu().tolist()
        return [(int(d), float(s)) for d, s in zip(docs, dscores) if d >= 0]

    def hits(self, dist_row: np.ndarray, rows_row: np.ndarray) -> List[ChunkRow]:
        out = []
        pks, docs, orders = self._pks.view(), self._doc.view(), self._order.view()
        for d, r in zip(dist_row, rows_row):
            if r < 0:
                break  # fewer matching rows than n, like [:n] on a shorter queryset
            out.append(ChunkRow(int(pks[r]), int(docs[r]), None if self._text is None else self._text[r], int(orders[r]), float(d)))
        return out

    def close(self):
        self.store.close()


_LOOKUPS = ("exact", "in", "lt", "lte", "gt", "gte")


class ChunkQuerySet:
    """A filterable view over a ``DeviceTable`` - the part of Django's QuerySet the reference's
    callers use before handing ``qs`` to the search (``.filter(...)`` with equality, ``__in`` and
    comparison lookups; e.g. ``Question.objects.filter(document__wiki__bot=bot, ...)`` in
    assistant/bot/services/context_service/steps/embeddings.py:26-29 and
    ``filter(document__id__lt=...)`` in assistant/processing/documents/steps/questions.py:124).

    A queryset is a list of predicates; nothing is evaluated until the search needs the allow bitmap,
    and then ONE kernel evaluates them for every row on the device (``dab_store_build_allow``) - no
    per-row work on the host, whatever the size of the table."""

    def __init__(self, table: DeviceTable, predicates: Optional[list] = None):
        self.table = table
        self._preds = list(predicates or [])       # (column name, lookup, value, negate)

    def _apply(self, kwargs, negate: bool) -> "ChunkQuerySet":
        preds = list(self._preds)
        for key, value in kwargs.items():
            name, op = key, "exact"
            if "__" in key and key.rsplit("__", 1)[1] in _LOOKUPS:
                name, op = key.rsplit("__", 1)
            if op == "in":
                value = list(value)
            preds.append((name, op, value, negate))
        return ChunkQuerySet(self.table, preds)

    def filter(self, **kwargs) -> "ChunkQuerySet":
        return self._apply(kwargs, False)

    def exclude(self, **kwargs) -> "ChunkQuerySet":
        return self._apply(kwargs, True)

    def all(self) -> "ChunkQuerySet":
        return self

    def allow_bitmap(self) -> Optional[torch.Tensor]:
        """None = every (live) row; else the device bitmap of the conjunction."""
        if not self._preds:
            return None            # tombstones are skipped by the scan itself
        return self.table.build_bitmap([self.table.predicate(n, op, v, neg) for n, op, v, neg in self._preds])

    def count(self) -> int:
        if not self._preds:
            return len(self.table) - getattr(self.table, "_n_dead", 0)
        bm = self.allow_bitmap()
        n = len(self.table)
        bits = (bm.view(torch.uint8)[:, None] >> torch.arange(8, device=bm.device, dtype=torch.uint8)[None, :]) & 1
        return int(bits.reshape(-1)[:n].sum().item())


# --------------------------------------------------------------------------- registry / configuration
_tables: Dict[Tuple[str, str], DeviceTable] = {}
_embedder = None


def register_table(model_label: str, field: str, table: DeviceTable) -> None:
    """Make ``table`` the device copy of ``<model_label>.<field>`` (e.g. ``("storage.question",
    "embedding")``) for searches that arrive with a Django queryset."""
    _tables[(model_label.lower(), field)] = table


def configure(embedder=None) -> None:
    """Set the embedder ``get_embedding`` uses - the stand-in for
    ``get_ai_embdedder(settings.EMBEDDING_AI_MODEL)`` (search_service.py:156-157)."""
    global _embedder
    _embedder = embedder


def _table_and_bitmap(qs, field: str):
    if isinstance(qs, ChunkQuerySet):
        return qs.table, qs.allow_bitmap(), None
    model = getattr(qs, "model", None)
    if model is None:
        raise DabError(1, "qs must be a ChunkQuerySet or a Django QuerySet")
    label = getattr(getattr(model, "_meta", None), "label_lower", getattr(model, "__name__", "").lower())
    table = _tables.get((label, field))
    if table is None:
        raise DabError(1, f"no device table registered for {label}.{field}")
    pks = list(qs.values_list("pk", flat=True))
    return table, table.bitmap_from_pks(pks), model


def _log_search(what: str, table, n_queries: int, n: int, bitmap, t0: float, t1: float, t2: float) -> None:
    """One DEBUG line per device search: sizes, bytes streamed, and where the wall time went - the counterpart of the
    reference's TimeDebugger around its search (assistant/utils/debug.py:5-30)."""
    if logger.isEnabledFor(logging.DEBUG):
        rows = len(table)
        elem = 2 if table.store.dtype == torch.bfloat16 else 4
        logger.debug("%s: rows=%d dim=%d queries=%d n=%d filtered=%s corpus_bytes=%d filter_ms=%.3f search_ms=%.3f",
                     what, rows, table.dim, n_queries, n, bitmap is not None, rows * (table.dim * elem + 4),
                     (t1 - t0) * 1e3, (t2 - t1) * 1e3)


def _search_sync(query_embedding, qs, n: int, field: str):
    t0 = time.perf_counter()
    table, bitmap, model = _table_and_bitmap(qs, field)
    t1 = time.perf_counter()
    dist, rows = table.search_rows(query_embedding, n, bitmap)
    _log_search("_objects_embedding_search", table, 1, n, bitmap, t0, t1, time.perf_counter())
    hits = table.hits(dist[0], rows[0])
    if model is None:
        return hits
    objs = model.objects.in_bulk([h.pk for h in hits])
    out = []
    for h in hits:
        o = objs.get(h.pk)
        if o is not None:
            o.distance = h.distance
            out.append(o)
    return out


# --------------------------------------------------------------------------- the reference's API
async def _objects_embedding_search(query_embedding: List[float], qs, n: int = 10, field: str = 'embedding') -> List[Any]:
    """search_service.py:185-196: the top ``n`` objects of ``qs`` ordered by cosine distance to
    ``query_embedding``, each with ``.distance`` (ascending).  The blocking device call runs in a
    worker thread (the reference hops threads with ``sync_to_async``, :191)."""
    return await asyncio.to_thread(_search_sync, query_embedding, qs, n, field)


async def embedding_search_documents(query_embedding: List[float], qs, n: int = 10):
    """search_service.py:161-166."""
    return await _objects_embedding_search(query_embedding, qs, n, field='content_embedding')


async def embedding_search_questions(query_embedding: List[float], qs, n: int = 10):
    """search_service.py:169-174."""
    return await _objects_embedding_search(query_embedding, qs, n)


async def embedding_search_sentences(query_embedding: List[float], qs, n: int = 10):
    """search_service.py:177-182."""
    return await _objects_embedding_search(query_embedding, qs, n)


def _search_many_sync(query_embeddings, qs, n: int, field: str):
    t0 = time.perf_counter()
    table, bitmap, model = _table_and_bitmap(qs, field)
    t1 = time.perf_counter()
    dist, rows = table.search_rows(query_embeddings, n, bitmap)
    _log_search("embedding_search_many", table, len(query_embeddings), n, bitmap, t0, t1, time.perf_counter())
    per_query = [table.hits(dist[i], rows[i]) for i in range(dist.shape[0])]
    if model is None:
        return per_query
    objs = model.objects.in_bulk(sorted({h.pk for hits in per_query for h in hits}))
    out = []
    for hits in per_query:
        cur = []
        for h in hits:
            o = objs.get(h.pk)
            if o is not None:
                import copy
                o = copy.copy(o)          # the same row may be a hit of several queries with different distances
                o.distance = h.distance
                cur.append(o)
        out.append(cur)
    return out


async def embedding_search_many(query_embeddings: List[List[float]], qs, n: int = 10, field: str = 'embedding') -> List[List[Any]]:
    """Batched ``_objects_embedding_search``: one corpus pass answers all queries.  This is what the
    ingest-time duplicate check needs - ``MergeQuestionsStep.run``
    (assistant/processing/documents/steps/questions.py:121-126) issues one top-1 search per new
    question against the same filtered queryset; here the whole document's questions go in one call."""
    if not query_embeddings:
        return []
    return await asyncio.to_thread(_search_many_sync, query_embeddings, qs, n, field)


def aggregate_documents(top_objects: Sequence[Any], max_scores_n: int, top_n: int) -> List[Tuple[int, float]]:
    """search_service.py:133-150 without the ORM fetch: group hits by ``document_id`` in rank order,
    keep documents with at least ``max_scores_n`` hits, score ``1 - mean(first max_scores_n
    distances)``, best ``top_n``."""
    docs = defaultdict(list)
    for obj in top_objects:
        docs[obj.document_id].append(obj)
    doc_scores = {
        doc_id: 1 - sum([o.distance for o in v[:max_scores_n]]) / max_scores_n
        for doc_id, v in docs.items() if len(v) >= max_scores_n
    }
    result = sorted(doc_scores.items(), key=lambda x: x[1], reverse=True)
    return result[:top_n]


async def embedding_search(query: str, qs, max_scores_n: int = 10, top_n: int = 10) -> List[Tuple[Any, float]]:
    """search_service.py:111-152: embed the query, take the ``max_scores_n * top_n * 10`` nearest
    rows (:129-131), aggregate per document."""
    logger.info(f'Embedding search for query: {query}')
    query_embedding = await get_embedding(query)
    n = max_scores_n * top_n * 10
    if isinstance(qs, ChunkQuerySet):
        # search + grouping both on the device: only <= top_n (doc id, score) pairs come back to the host
        scored = await asyncio.to_thread(qs.table.search_documents, query_embedding, n, max_scores_n, top_n,
                                         qs.allow_bitmap())
        lookup = qs.table.documents
        return [(lookup.get(doc_id, doc_id), score) for doc_id, score in scored]
    top_objects = await _objects_embedding_search(query_embedding, qs, n=n)
    scored = aggregate_documents(top_objects, max_scores_n, top_n)
    doc_model = qs.model._meta.get_field('document').related_model
    found = {d.id: d for d in await asyncio.to_thread(
        lambda: list(doc_model.objects.filter(id__in=[d for d, _ in scored])))}
    return [(found[doc_id], score) for doc_id, score in scored if doc_id in found]


async def get_embedding(text: str) -> List[float]:
    """search_service.py:155-158: ``get_ai_embdedder(settings.EMBEDDING_AI_MODEL)`` unless an embedder
    was set with ``configure``."""
    embedder = _embedder
    if embedder is None:
        from .embedders import get_ai_embdedder
        embedder = get_ai_embdedder()
    return (await embedder.embeddings([text]))[0]
