"""Queryset filters evaluated on the device (dab_store_set_column / dab_store_build_allow) against the NumPy
statement of the same WHERE clauses - the filters every reference call carries
(assistant/bot/services/context_service/steps/embeddings.py:26-29, assistant/processing/documents/steps/questions.py:124)."""
import asyncio

import numpy as np
import pytest
import torch

from oracle import search_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _unpack(bitmap, n):
    b = bitmap.view(torch.uint8).cpu().numpy()
    return np.unpackbits(b, bitorder="little")[:n].astype(bool)


def _table(n=5000, dim=64, seed=3):
    from dab_b200 import search_service as ss
    rng = np.random.default_rng(seed)
    emb = rng.standard_normal((n, dim)).astype(np.float32)
    cols = dict(pks=(np.arange(n) * 3 + 7).tolist(), docs=(np.arange(n) // 10).tolist(), bot=(rng.integers(0, 4, n)).tolist(),
                status=[["completed", "processing", "failed"][i] for i in rng.integers(0, 3, n)])
    t = ss.DeviceTable(n + 100, dim, torch.float32, DEV)
    half = n // 2                                           # two appends: columns are uploaded incrementally
    for a, b in ((0, half), (half, n)):
        t.add(cols["pks"][a:b], emb[a:b], cols["docs"][a:b], document__wiki__bot=cols["bot"][a:b],
              document__processing__status=cols["status"][a:b])
    return ss, t, emb, {k: np.asarray(v) for k, v in cols.items()}


def test_every_lookup_matches_numpy():
    ss, t, emb, c = _table()
    n = len(t)
    qs = ss.ChunkQuerySet(t)
    cases = [
        (qs.filter(document__wiki__bot=2), c["bot"] == 2),
        (qs.filter(document__wiki__bot=2, document__processing__status="completed"), (c["bot"] == 2) & (c["status"] == "completed")),
        (qs.filter(document__id__lt=123), c["docs"] < 123),
        (qs.filter(document_id__lte=123).filter(document_id__gt=17), (c["docs"] <= 123) & (c["docs"] > 17)),
        (qs.filter(document__id__gte=400), c["docs"] >= 400),
        (qs.filter(pk__in=[7, 10, 13, 999999, 7 + 3 * 4999]), np.isin(c["pks"], [7, 10, 13, 7 + 3 * 4999])),
        (qs.exclude(document__processing__status="failed"), c["status"] != "failed"),
        (qs.exclude(document__wiki__bot__in=[0, 3]), ~np.isin(c["bot"], [0, 3])),
        (qs.filter(document__processing__status__in=["completed", "never-seen"]), c["status"] == "completed"),
        (qs.filter(document__processing__status="never-seen"), np.zeros(n, bool)),
        (qs.exclude(document__processing__status="never-seen"), np.ones(n, bool)),
    ]
    many = qs
    want_many = np.ones(n, bool)
    for j in range(10):                                     # more predicates than one kernel call takes
        many = many.exclude(document__id=j * 7)
        want_many &= c["docs"] != j * 7
    cases.append((many, want_many))
    for got_qs, want in cases:
        np.testing.assert_array_equal(_unpack(got_qs.allow_bitmap(), n), want)
        assert got_qs.count() == int(want.sum())
    with pytest.raises(Exception):
        qs.filter(document__processing__status__lt="x").allow_bitmap()   # no order on dictionary-encoded columns
    t.close()


def test_filtered_search_tombstones_and_compaction():
    ss, t, emb, c = _table()
    n = len(t)
    q = emb[17] + 0.01
    want_mask = (c["bot"] == 1) & (c["docs"] < 300)
    qs = ss.ChunkQuerySet(t).filter(document__wiki__bot=1, document__id__lt=300)
    hits = asyncio.run(ss._objects_embedding_search(q.tolist(), qs, n=20))
    S = search_ref.exact_scores(emb, q[None])[0]
    ids, sc = search_ref.rank_topk(S, 20, want_mask)
    assert [h.pk for h in hits] == c["pks"][ids].tolist()
    np.testing.assert_allclose([h.distance for h in hits], 1 - sc, atol=2e-6)
    # delete some rows: the bitmap itself excludes tombstones
    gone = c["pks"][ids[:5]].tolist()
    t.delete(gone)
    alive = ~np.isin(c["pks"], gone)
    np.testing.assert_array_equal(_unpack(qs.allow_bitmap(), n), want_mask & alive)
    np.testing.assert_array_equal(_unpack(t.build_bitmap([]), n), alive)
    hits2 = asyncio.run(ss._objects_embedding_search(q.tolist(), qs, n=15))
    assert [h.pk for h in hits2] == c["pks"][ids[5:20]].tolist()
    # compaction moves the HBM columns with the rows
    t.compact()
    keep = np.flatnonzero(alive)
    np.testing.assert_array_equal(_unpack(qs.allow_bitmap(), len(t)), want_mask[keep])
    hits3 = asyncio.run(ss._objects_embedding_search(q.tolist(), qs, n=15))
    assert [h.pk for h in hits3] == [h.pk for h in hits2]
    # the queryset as a pk list (what a Django QuerySet gives): pk IN (...) on the device
    pk_list = c["pks"][keep][::7].tolist() + [123456789]
    np.testing.assert_array_equal(_unpack(t.bitmap_from_pks(pk_list), len(t)), np.isin(c["pks"][keep], pk_list))
    t.close()


def test_raw_c_abi_predicates_and_errors():
    from dab_b200 import _lib, ops
    N, D = 1000, 32
    store = ops.ChunkStore(N, D, torch.float32, DEV, id_base=500)
    store.append(torch.randn(N, D, device=DEV))
    col = torch.arange(N, device=DEV, dtype=torch.int64) % 10
    store.set_column(3, col)
    bm = store.build_allow([(3, _lib.DAB_OP_EQ, 4), (_lib.DAB_COL_ID, _lib.DAB_OP_GE, 1000)])
    want = (np.arange(N) % 10 == 4) & (np.arange(N) + 500 >= 1000)
    np.testing.assert_array_equal(_unpack(bm, N), want)
    # a queryset as a list of rows: device scatter, tombstones and out-of-range rows skipped, duplicates harmless
    store.delete_rows(torch.tensor([7], device=DEV))
    listed = torch.tensor([3, 7, 64, 999, 3, 1000, -1], device=DEV)
    want_rows = np.zeros(N, bool)
    want_rows[[3, 64, 999]] = True
    np.testing.assert_array_equal(_unpack(store.allow_from_rows(listed), N), want_rows)
    assert not _unpack(store.allow_from_rows(torch.zeros(0, dtype=torch.int64, device=DEV)), N).any()
    with pytest.raises(ops.DabError):
        store.build_allow([(5, _lib.DAB_OP_EQ, 1)])          # never set
    with pytest.raises(ops.DabError):
        store.set_column(8, col)
    with pytest.raises(ops.DabError):
        store.set_column(0, col, row0=1)                     # beyond capacity
    # a filtered search through the raw bitmap
    q = torch.randn(2, D, device=DEV)
    s, i = store.search(q, 5, bm)
    assert want[(i.cpu().numpy() - 500).ravel()].all()
    store.close()


def test_stateful_table_fuzz_slice():
    """A fixed-seed slice of tools/fuzz_table.py: random add / delete / compact / snapshot sequences on a DeviceTable,
    every step followed by a filtered search checked against a NumPy model of the table."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_table
    rng = np.random.default_rng(424242)
    for s in range(40):
        fuzz_table.one_sequence(rng, s)
