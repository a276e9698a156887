"""Oracle for exact cosine top-k retrieval (retrieve half of the hot path).

TEST INFRASTRUCTURE ONLY - see ``oracle/__init__.py``.

What it restates (reference file:line, relative to /root/reference):

* ``cosine_similarity`` / ``embeddings_similarity`` - ``assistant/rag/services/search_service.py:17-21,32-35``
  (NumPy, Python lists -> float64).
* the legacy exact scan - ``search_service.py:76-84`` (per-row cosine, ``np.argsort(sims)[::-1]``).
* ``_objects_embedding_search`` - ``search_service.py:185-196``: rows ordered by
  ``distance = CosineDistance(field, q)`` ascending, first ``n``.  The arithmetic behind
  ``CosineDistance`` is pgvector's ``<=>`` (third party, absent from the reference tree,
  **parity unpinned**); its published algorithm is restated in ``oracle/scan_ref.c`` and,
  for small inputs, in :func:`pgvector_cosine_distance` below.
* doc-level aggregation - ``search_service.py:129-150``.

Ranking contract of the build (BASELINE.json north_star): score descending, row index
ascending on ties; rows whose distance is NaN (zero vectors) sort last, like PostgreSQL's
ORDER BY ... ASC puts NaN after every number.
"""
from __future__ import annotations

from collections import defaultdict
from typing import Dict, List, Sequence, Tuple

import numpy as np

# ----------------------------------------------------------------------------- bf16 helpers


def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even fp32 -> bf16, returned as uint16 bit patterns."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    rounding = ((u >> 16) & 1) + np.uint32(0x7FFF)
    out = ((u + rounding) >> 16).astype(np.uint16)
    nan = np.isnan(x)
    if nan.any():
        out = np.where(nan, np.uint16(0x7FC0), out)
    return out


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    """Exact upcast of bf16 bit patterns to fp32."""
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


# ----------------------------------------------------------------------------- reference NumPy statement


def cosine_similarity(a, b):
    """search_service.py:17-21 restated: dot / (|a| * |b|) in the dtype NumPy picks
    (float64 when fed Python lists, as ``embeddings_similarity`` :32-35 does)."""
    a = np.asarray(a)
    b = np.asarray(b)
    return np.dot(a, b) / (np.linalg.norm(a) * np.linalg.norm(b))


def legacy_scan_rank(query: Sequence[float], rows: Sequence[Sequence[float]]):
    """search_service.py:76-84 restated: one cosine per stored row in a Python loop,
    then indices by descending similarity (``np.argsort(sims)[::-1]``, not stable)."""
    q = np.array(query)
    sims = [cosine_similarity(q, np.array(r)) for r in rows]
    order = np.argsort(sims)[::-1]
    return order, np.asarray(sims)


# ----------------------------------------------------------------------------- pgvector <=> restatement


def pgvector_cosine_distance(a: np.ndarray, b: np.ndarray) -> float:
    """Cosine distance as pgvector's ``<=>`` publishes it: three fp32 running sums
    (a.b, a.a, b.b), ``similarity = (double)dot / sqrt((double)na * (double)nb)``,
    clamped to [-1, 1], distance ``1 - similarity``; NaN when either vector is zero.
    Pure-Python loop - small inputs only (``oracle/scan_ref.c`` is the fast twin)."""
    a = np.asarray(a, dtype=np.float32)
    b = np.asarray(b, dtype=np.float32)
    if a.shape != b.shape:
        raise ValueError("different vector dimensions %d and %d" % (a.size, b.size))
    dot = np.float32(0)
    na = np.float32(0)
    nb = np.float32(0)
    for x, y in zip(a, b):
        dot = np.float32(dot + np.float32(x * y))
        na = np.float32(na + np.float32(x * x))
        nb = np.float32(nb + np.float32(y * y))
    den = np.sqrt(np.float64(na) * np.float64(nb))
    with np.errstate(invalid="ignore", divide="ignore"):
        sim = np.float64(dot) / den
    if np.isnan(sim):
        return float("nan")
    sim = min(1.0, max(-1.0, float(sim)))
    return 1.0 - sim


# ----------------------------------------------------------------------------- the build's ranking oracle


def exact_scores(rows_f32: np.ndarray, queries_f32: np.ndarray) -> np.ndarray:
    """Cosine of every stored row with every query, accumulated in float64 from the
    *stored* values (callers upcast bf16 rows exactly first).  Shape [Q, N].
    Zero rows / zero queries give NaN (pgvector semantics)."""
    r = np.asarray(rows_f32, dtype=np.float64)
    q = np.atleast_2d(np.asarray(queries_f32, dtype=np.float64))
    rn = np.sqrt((r * r).sum(axis=1))
    qn = np.sqrt((q * q).sum(axis=1))
    with np.errstate(invalid="ignore", divide="ignore"):
        s = (q @ r.T) / (qn[:, None] * rn[None, :])
    return np.clip(s, -1.0, 1.0)


def rank_topk(scores_1d: np.ndarray, k: int, allow: np.ndarray | None = None):
    """Indices of the top-k scores under (score desc, index asc); NaN rows last and
    never returned ahead of a number.  ``allow`` is an optional boolean row mask (the
    queryset filter).  Returns fewer than k entries when fewer rows qualify - exactly
    like ``[:n]`` on a shorter queryset (search_service.py:194)."""
    s = np.asarray(scores_1d, dtype=np.float64)
    idx = np.arange(s.shape[0])
    keep = ~np.isnan(s)
    if allow is not None:
        keep &= np.asarray(allow, dtype=bool)
    idx = idx[keep]
    order = np.lexsort((idx, -s[idx]))
    top = idx[order[:k]]
    return top.astype(np.int64), s[top]


def topk_oracle(rows_f32, queries_f32, k, allow=None):
    """[Q] lists of (ids int64, scores float64) - the ground truth of the parity tests."""
    S = exact_scores(rows_f32, queries_f32)
    return [rank_topk(S[i], k, allow) for i in range(S.shape[0])], S


def check_topk_parity(S_row: np.ndarray, got_ids: np.ndarray, got_scores: np.ndarray, k: int,
                      allow: np.ndarray | None = None, gap: float = 1e-6, rtol: float = 1e-5,
                      floor: float = 2e-7) -> None:
    """Assert the north_star parity bar for one query.

    * ``got_ids`` must equal the oracle's ids position by position, except where the
      oracle scores of the two ids differ by < ``gap`` (1e-6: adjacent scores closer than
      fp32 accumulation noise may swap);
    * no id may repeat, every id must be allowed, and each reported score must be within
      ``rtol`` RELATIVE (1e-5 * |score|, north_star's bar) of the float64 oracle score of the
      id it is reported for; ``floor`` (2e-7 absolute) only covers scores so close to zero
      that fp32 arithmetic cannot meet a relative bound.
    """
    want_ids, want_scores = rank_topk(S_row, k, allow)
    n = len(want_ids)
    got_ids = np.asarray(got_ids)[:n]
    got_scores = np.asarray(got_scores, dtype=np.float64)[:n]
    assert len(got_ids) == n, f"expected {n} results, got {len(got_ids)}"
    assert len(set(got_ids.tolist())) == n, "duplicate ids in result"
    if allow is not None:
        assert np.asarray(allow, dtype=bool)[got_ids].all(), "filtered-out row returned"
    for pos in range(n):
        g, w = int(got_ids[pos]), int(want_ids[pos])
        if g != w:
            assert abs(S_row[g] - S_row[w]) < gap, (
                f"rank {pos}: got id {g} (oracle score {S_row[g]!r}) but oracle has id {w} "
                f"({S_row[w]!r}); gap {abs(S_row[g] - S_row[w]):.3e} >= {gap}")
        assert abs(got_scores[pos] - S_row[g]) <= max(rtol * abs(S_row[g]), floor), (
            f"rank {pos}: score {got_scores[pos]!r} vs oracle {S_row[g]!r}")


def merge_topk_lists(gathered_scores: np.ndarray, gathered_ids: np.ndarray, k: int):
    """Checker for the cross-shard merge: ``[R,Q,k]`` per-shard lists -> ``[Q,k]`` under
    (score desc, id asc); id -1 marks an empty slot.  The single-table ORDER BY ... LIMIT n of
    search_service.py:192-194 applied to the union of the shard results."""
    gs = np.asarray(gathered_scores, dtype=np.float64)
    gi = np.asarray(gathered_ids, dtype=np.int64)
    R, Q, _ = gs.shape
    out_s = np.full((Q, k), -np.inf)
    out_i = np.full((Q, k), -1, dtype=np.int64)
    for q in range(Q):
        s, i = gs[:, q].ravel(), gi[:, q].ravel()
        keep = i >= 0
        s, i = s[keep], i[keep]
        order = np.lexsort((i, -s))[:k]
        out_s[q, :len(order)] = s[order]
        out_i[q, :len(order)] = i[order]
    return out_s, out_i


# ----------------------------------------------------------------------------- doc-level aggregation


def doc_level_scores(hit_doc_ids: Sequence[int], hit_distances: Sequence[float],
                     max_scores_n: int, top_n: int) -> List[Tuple[int, float]]:
    """search_service.py:133-150 restated.  ``hit_*`` are the top-n hits in rank order
    (ascending distance).  Hits are grouped by document in that order; a document with
    fewer than ``max_scores_n`` hits is dropped; its score is
    ``1 - sum(first max_scores_n distances) / max_scores_n``; documents are sorted by
    score descending (Python's stable sort; the reference's pre-sort order comes from an
    unordered ``filter(id__in=...)`` query, so ties are resolved here by first
    appearance in the hit list) and cut to ``top_n``."""
    groups: Dict[int, List[float]] = defaultdict(list)
    for d, dist in zip(hit_doc_ids, hit_distances):
        groups[int(d)].append(float(dist))
    scored = [(d, 1 - sum(v[:max_scores_n]) / max_scores_n)
              for d, v in groups.items() if len(v) >= max_scores_n]
    scored.sort(key=lambda x: x[1], reverse=True)
    return scored[:top_n]
