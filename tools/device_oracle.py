"""Device oracle: the second level of the two-level oracle of SURVEY.md section 8(c).

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  Nothing under ``django-assistant-bot_b200/`` imports this file; it
is used by ``tests/`` (parity at BASELINE sizes, where the NumPy oracle of ``oracle/search_ref.py`` would
take hours and the rows do not fit host memory) and by ``bench.py``'s ``verified`` guard.  It is plain
PyTorch - no kernel of this repository takes part - and is itself pinned against ``oracle.search_ref`` on a
<= 1M-row subsample in ``tests/test_gpu_at_scale.py``.

What it computes is the ranking ``_objects_embedding_search`` asks PostgreSQL for
(assistant/rag/services/search_service.py:192-194: ORDER BY cosine distance ASC LIMIT n) under the build's
tie-break (score desc, row id asc):

1. candidate pass - block by block, an fp32 ``torch.matmul`` (TF32 off) of the unit-norm queries with the
   upcast stored rows, divided by the rows' fp32 norms, ``topk`` of ``k + pad`` per block, merged; the
   candidates' stored rows travel with them;
2. exact pass - the surviving ``k + pad`` candidates of every query are re-scored in FLOAT64 from their stored
   values (bf16 upcast exactly), like ``oracle.search_ref.exact_scores``, and sorted lexicographically.

The pad absorbs fp32 noise of pass 1 at the cut (a row whose float64 score belongs in the top-k always
survives unless more than ``pad`` rows lie within ~1e-6 of the k-th score).
"""
from __future__ import annotations

from typing import Callable, Iterable, Optional, Tuple

import torch


def _no_tf32():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        torch.set_float32_matmul_precision("highest")
    except Exception:  # pragma: no cover
        pass


def oracle_topk(blocks: Iterable[Tuple[int, torch.Tensor]], queries: torch.Tensor, k: int, pad: int = 64,
                allow: Optional[Callable[[int, int], Optional[torch.Tensor]]] = None,
                q_chunk: int = 256) -> Tuple[torch.Tensor, torch.Tensor]:
    """``blocks`` yields ``(first_global_row, rows [n, D])`` with rows in the STORE dtype (bf16 or fp32) on the
    device; ``queries`` [Q, D] fp32 (any norm).  ``allow(row0, n)`` may return a bool [n] row filter for a block.
    Returns ``(ids [Q, k+pad] int64, scores [Q, k+pad] float64)`` sorted (score desc, id asc); slots beyond the
    rankable rows hold id -1 / score -inf.  Zero rows never rank (pgvector gives them NaN distance)."""
    _no_tf32()
    dev = queries.device
    Q, D = queries.shape
    kk = k + pad
    q32 = torch.nn.functional.normalize(queries.float(), dim=1)
    best_s = torch.full((Q, kk), float("-inf"), device=dev)
    best_i = torch.full((Q, kk), -1, dtype=torch.int64, device=dev)
    best_r = None
    for row0, rows in blocks:
        n = rows.shape[0]
        if n == 0:
            continue
        if best_r is None:
            best_r = torch.zeros((Q, kk, D), dtype=rows.dtype, device=dev)
        part = rows.float()
        nrm = part.norm(dim=1)
        dead = ~(nrm > 0) | ~torch.isfinite(nrm)
        if allow is not None:
            a = allow(row0, n)
            if a is not None:
                dead = dead | ~a.to(dev)
        kb = min(kk, n)
        for q0 in range(0, Q, q_chunk):
            q1 = min(Q, q0 + q_chunk)
            sc = (q32[q0:q1] @ part.T) / nrm[None, :]
            sc[:, dead] = float("-inf")
            top = torch.topk(sc, kb, dim=1)
            del sc
            ids = top.indices + row0
            ids = torch.where(torch.isfinite(top.values), ids, torch.full_like(ids, -1))
            cs = torch.cat([best_s[q0:q1], top.values], 1)
            ci = torch.cat([best_i[q0:q1], ids], 1)
            cr = torch.cat([best_r[q0:q1], rows[top.indices.reshape(-1)].view(q1 - q0, kb, D)], 1)
            keep = torch.topk(cs, kk, dim=1)
            best_s[q0:q1] = keep.values
            best_i[q0:q1] = torch.gather(ci, 1, keep.indices)
            best_r[q0:q1] = torch.gather(cr, 1, keep.indices[:, :, None].expand(-1, -1, D))
            del cs, ci, cr, keep, top
        del part, nrm
    if best_r is None:
        return best_i, best_s.double()
    # exact pass in float64 from the stored values
    q64 = queries.double()
    q64 = q64 / q64.norm(dim=1, keepdim=True)
    s64 = torch.empty((Q, kk), dtype=torch.float64, device=dev)
    for q0 in range(0, Q, 32):
        q1 = min(Q, q0 + 32)
        r64 = best_r[q0:q1].double()
        s = torch.einsum("qkd,qd->qk", r64, q64[q0:q1]) / r64.norm(dim=2)
        s64[q0:q1] = s.clamp(-1.0, 1.0)
    s64 = torch.where(best_i >= 0, s64, torch.full_like(s64, float("-inf")))
    # lexicographic (score desc, id asc): stable sort by id, then stable sort by score
    big = torch.where(best_i >= 0, best_i, torch.full_like(best_i, 2 ** 62))
    o1 = torch.sort(big, dim=1, stable=True).indices
    s1, i1 = torch.gather(s64, 1, o1), torch.gather(best_i, 1, o1)
    o2 = torch.sort(s1, dim=1, descending=True, stable=True).indices
    return torch.gather(i1, 1, o2), torch.gather(s1, 1, o2)


def merge_rank_lists(ids_list, scores_list, kk: int):
    """Merge per-shard oracle lists ([Q, kk] each, disjoint id ranges) into the global one (same ordering)."""
    ci, cs = torch.cat(ids_list, 1), torch.cat(scores_list, 1)
    big = torch.where(ci >= 0, ci, torch.full_like(ci, 2 ** 62))
    o1 = torch.sort(big, dim=1, stable=True).indices
    s1, i1 = torch.gather(cs, 1, o1), torch.gather(ci, 1, o1)
    o2 = torch.sort(s1, dim=1, descending=True, stable=True).indices
    return torch.gather(i1, 1, o2)[:, :kk], torch.gather(s1, 1, o2)[:, :kk]


def parity_report(got_scores: torch.Tensor, got_ids: torch.Tensor, want_ids: torch.Tensor, want_scores: torch.Tensor,
                  k: int, gap: float = 1e-6, rtol: float = 1e-5, floor: float = 2e-7) -> dict:
    """The north_star parity bar for every query at once (same rules as ``oracle.search_ref.check_topk_parity``):

    * position by position the returned id equals the oracle's id, except where the float64 oracle scores of the
      two ids differ by < ``gap`` (near-ties may swap);
    * no id repeats; unused slots are (-1, -inf) exactly where the oracle runs out of rankable rows;
    * every returned score is within ``max(rtol * |s|, floor)`` of the float64 oracle score of the id it is
      reported for (1e-5 RELATIVE; the floor only covers scores so close to zero that fp32 cannot do better).

    Returns counts of violations and the worst relative score error; ``ok`` is True when all counts are zero."""
    Q = got_ids.shape[0]
    got_ids = got_ids[:, :k]
    got_scores = got_scores[:, :k].double()
    wi, ws = want_ids, want_scores
    match = got_ids[:, :, None] == wi[:, None, :]                       # [Q, k, kk]
    valid_got = got_ids >= 0
    found = match.any(-1)
    pos = match.to(torch.uint8).argmax(-1)
    s_of_got = torch.gather(ws, 1, pos)                                  # oracle score of the id we returned
    want_valid = wi[:, :k] >= 0
    # slots
    slot_bad = (valid_got != want_valid)
    empty_bad = (~valid_got) & ~(torch.isinf(got_scores) & (got_scores < 0))
    unknown = valid_got & ~found
    differs = valid_got & want_valid & found & (got_ids != wi[:, :k])
    order_bad = differs & ((s_of_got - ws[:, :k]).abs() >= gap)
    err = (got_scores - s_of_got).abs()
    tol = torch.clamp(rtol * s_of_got.abs(), min=floor)
    score_bad = valid_got & found & ~(err <= tol)
    rel = torch.where(valid_got & found, err / s_of_got.abs().clamp(min=1e-30), torch.zeros_like(err))
    srt = torch.sort(torch.where(valid_got, got_ids, -1 - torch.arange(got_ids.shape[1], device=got_ids.device)[None, :]), dim=1).values
    dup = (srt[:, 1:] == srt[:, :-1]).any() if got_ids.shape[1] > 1 else torch.tensor(False)
    out = dict(queries=int(Q), k=int(k), slot_mismatch=int(slot_bad.sum()), bad_empty_slot=int(empty_bad.sum()),
               id_not_in_oracle_list=int(unknown.sum()), order_violations=int(order_bad.sum()),
               swapped_near_ties=int((differs & ~order_bad).sum()), score_violations=int(score_bad.sum()),
               duplicate_ids=int(bool(dup)), max_rel_score_err=float(rel.max()) if rel.numel() else 0.0,
               max_abs_score_err=float(torch.where(valid_got & found, err, torch.zeros_like(err)).max()) if err.numel() else 0.0)
    out["ok"] = all(out[x] == 0 for x in ("slot_mismatch", "bad_empty_slot", "id_not_in_oracle_list", "order_violations",
                                           "score_violations", "duplicate_ids"))
    return out


def doc_level_scores(hit_doc_ids, hit_distances, max_scores_n: int, top_n: int):
    """The grouping of ``embedding_search`` (assistant/rag/services/search_service.py:133-150) restated for the
    checker: hits in rank order are grouped per document, documents with fewer than ``max_scores_n`` hits are
    dropped, score = 1 - mean of the first ``max_scores_n`` distances, best ``top_n`` (ties: first appearance)."""
    groups, order = {}, []
    for d, dist in zip(hit_doc_ids, hit_distances):
        d = int(d)
        if d not in groups:
            groups[d] = []
            order.append(d)
        groups[d].append(float(dist))
    scored = [(d, 1 - sum(groups[d][:max_scores_n]) / max_scores_n) for d in order if len(groups[d]) >= max_scores_n]
    scored.sort(key=lambda x: x[1], reverse=True)
    return scored[:top_n]
