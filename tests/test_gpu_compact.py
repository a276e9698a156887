"""Physical compaction of tombstones (dab_store_compact): the surviving rows keep their order, ids
travel with explicit-id stores, implicit ids are renumbered, genuine zero vectors stay, and the
searches before and after agree with the oracle over the survivors."""
import asyncio

import numpy as np
import pytest
import torch

from oracle import search_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _store(n, d, dtype, explicit, seed):
    from dab_b200 import ops
    g = torch.Generator().manual_seed(seed)
    rows = torch.randn(n, d, generator=g)
    rows[5] = 0                                   # a genuine zero vector: never ranks, but is not a tombstone
    ids = (torch.randperm(10 * n, generator=g)[:n] + 1000).to(torch.int64) if explicit else None
    st = ops.ChunkStore(n + 16, d, dtype, DEV)
    st.append(rows.to(DEV), ids.to(DEV) if explicit else None)
    return st, rows.to(dtype).float().numpy(), (ids.numpy() if explicit else np.arange(n))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("explicit", [False, True])
@pytest.mark.parametrize("n,d", [(50_000, 768), (1000, 104), (300_000, 64), (777, 8)])
def test_compact_drops_tombstones_keeps_order(dtype, explicit, n, d):
    st, stored, ids = _store(n, d, dtype, explicit, seed=n + d)
    rng = np.random.default_rng(n)
    dead = np.unique(np.concatenate([rng.choice(n, size=n // 3, replace=False), [0, n - 1, 7]]))
    dead = dead[dead != 5]
    st.delete_rows(torch.from_numpy(dead).to(DEV))
    alive = np.ones(n, dtype=bool)
    alive[dead] = False
    q = torch.randn(4, d, generator=torch.Generator().manual_seed(1))
    k = 10
    s_before, i_before = st.search(q.to(DEV), k)
    assert st.compact() == len(dead) and len(st) == alive.sum()
    assert st.compact() == 0                                       # idempotent
    rows_out, inv_out, ids_out = st.export(0, len(st))
    np.testing.assert_array_equal(rows_out.float().cpu().numpy(), stored[alive])
    want_ids = ids[alive] if explicit else np.arange(alive.sum())
    np.testing.assert_array_equal(ids_out.cpu().numpy(), want_ids)
    inv = inv_out.cpu().numpy()
    zero_at = int(alive[:5].sum())                                 # where the zero vector went
    assert inv[zero_at] == 0 and not np.signbit(inv[zero_at]) and (np.delete(inv, zero_at) > 0).all()
    # same answers as before the compaction (ids for explicit stores, survivors' positions otherwise)
    s_after, i_after = st.search(q.to(DEV), k)
    np.testing.assert_allclose(s_after.cpu().numpy(), s_before.cpu().numpy(), rtol=0, atol=0)
    new_row_of_old = np.cumsum(alive) - 1
    want = i_before.cpu().numpy() if explicit else new_row_of_old[i_before.cpu().numpy()]
    np.testing.assert_array_equal(i_after.cpu().numpy(), want)
    # and the oracle over the survivors
    S = search_ref.exact_scores(stored[alive], q.numpy())
    ia = i_after.cpu().numpy()
    if explicit:
        back = {int(v): j for j, v in enumerate(want_ids)}
        ia = np.vectorize(back.get)(ia)
    for qi in range(q.shape[0]):
        search_ref.check_topk_parity(S[qi], ia[qi], s_after[qi].cpu().numpy(), k)
    # the store keeps working: append after compaction lands behind the survivors
    extra = torch.randn(8, d, generator=torch.Generator().manual_seed(2))
    st.append(extra.to(DEV), (torch.arange(8) + 10_000_000).to(DEV) if explicit else None)
    _, _, ids2 = st.export(len(st) - 8, 8)
    np.testing.assert_array_equal(ids2.cpu().numpy(), np.arange(8) + (10_000_000 if explicit else alive.sum()))
    st.close()


def test_compact_everything_and_nothing():
    from dab_b200 import ops
    st = ops.ChunkStore(100, 64, torch.float32, DEV)
    assert st.compact() == 0                                       # empty store
    st.append(torch.randn(100, 64).to(DEV))
    assert st.compact() == 0                                       # nothing deleted
    st.delete_rows(torch.arange(100).to(DEV))
    assert st.compact() == 100 and len(st) == 0
    s, i = st.search(torch.randn(2, 64).to(DEV), 5)
    assert (i.cpu().numpy() == -1).all()
    st.append(torch.randn(3, 64).to(DEV))
    s, i = st.search(torch.randn(1, 64).to(DEV), 5)
    assert sorted(i.cpu().numpy()[0][:3].tolist()) == [0, 1, 2]
    st.close()


def test_device_table_compact_keeps_pks_and_filters():
    """DeviceTable.compact: host columns follow the device rows; every queryset answers as before."""
    from dab_b200 import search_service as ss
    rng = np.random.default_rng(11)
    n, d = 6000, 768
    emb = rng.standard_normal((n, d)).astype(np.float32)
    pks = np.arange(100, 100 + n)
    docs = rng.integers(1, 60, size=n)
    t = ss.DeviceTable(n, d, torch.float32, DEV)
    t.add(pks, emb, document_ids=docs, texts=[f"s{i}" for i in range(n)], document__wiki__bot=docs % 3)
    gone = pks[(docs % 5 == 0)]
    assert t.delete(gone.tolist()) == len(gone)
    q = rng.standard_normal(d).astype(np.float32)
    before = asyncio.run(ss._objects_embedding_search(q.tolist(), ss.ChunkQuerySet(t).filter(document__wiki__bot=1), n=25))
    assert t.compact() == len(gone) and len(t) == n - len(gone) and t.compact() == 0
    after = asyncio.run(ss._objects_embedding_search(q.tolist(), ss.ChunkQuerySet(t).filter(document__wiki__bot=1), n=25))
    assert [h.id for h in after] == [h.id for h in before]
    assert [h.text for h in after] == [h.text for h in before]
    np.testing.assert_allclose([h.distance for h in after], [h.distance for h in before], atol=0)
    assert ss.ChunkQuerySet(t).count() == n - len(gone)
    alive = ~np.isin(pks, gone)
    S = search_ref.exact_scores(emb[alive], q)[0]
    want, _ = search_ref.rank_topk(S, 10)
    hits = asyncio.run(ss._objects_embedding_search(q.tolist(), ss.ChunkQuerySet(t), n=10))
    assert [h.id for h in hits] == list(pks[alive][want])
    # deleting and adding keep working on the compacted table
    assert t.delete([hits[0].id]) == 1
    t.add([999_999], emb[:1], document_ids=[1], texts=["new"], document__wiki__bot=[1])
    hits2 = asyncio.run(ss._objects_embedding_search(emb[0].tolist(), ss.ChunkQuerySet(t), n=3))
    assert hits[0].id not in [h.id for h in hits2] and hits2[0].id in (999_999, int(pks[0]))
    t.close()
