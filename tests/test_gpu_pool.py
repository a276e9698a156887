"""Parity of the fused pool+normalise CUDA kernel with the oracle (through the C-ABI)."""
import os

import numpy as np
import pytest
import torch

from oracle import pool_ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _run(hidden, mask, **kw):
    from dab_b200 import ops
    dev = torch.device("cuda:0")
    return ops.pool_l2norm(hidden.to(dev), None if mask is None else mask.to(dev), **kw)


def test_pool_matches_reference_golden(golden_dir):
    """The CUDA kernel on the padded batch == the reference's per-text output, normalised
    (assistant/ai/embedders/transformers.py:25), 1e-5 relative."""
    g = np.load(os.path.join(golden_dir, "pool_ref.npz"))
    out, inv = _run(torch.from_numpy(g["hidden"]), torch.from_numpy(g["mask"]), want_inv_norm=True)
    ref = g["ref_embeddings"]
    ref_unit = ref / np.linalg.norm(ref, axis=1, keepdims=True)
    np.testing.assert_allclose(out.cpu().numpy(), ref_unit, rtol=1e-5, atol=2e-6)
    # out / inv is the reference's raw (un-normalised) mean
    np.testing.assert_allclose(out.cpu().numpy() / inv.cpu().numpy()[:, None], ref, rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("B,T,D", [(32, 256, 384), (3, 7, 768), (5, 512, 1024), (2, 33, 2048), (1, 1, 8), (4, 19, 4096),
                                   (300, 16, 64)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("mask_dtype", [torch.int64, torch.uint8, None])
def test_pool_parity(B, T, D, dtype, mask_dtype):
    g = torch.Generator().manual_seed(B * 1000 + T + D)
    hidden = torch.randn(B, T, D, generator=g).to(dtype)
    mask = None
    if mask_dtype is not None:
        lengths = torch.randint(1, T + 1, (B,), generator=g)
        mask = (torch.arange(T)[None, :] < lengths[:, None]).to(mask_dtype)
        if B > 1:
            mask[1] = torch.randint(0, 2, (T,), generator=g).to(mask_dtype)  # ragged, non-prefix mask
    out, ob, inv = _run(hidden, mask, want_bf16=True, want_inv_norm=True)
    m = np.ones((B, T), dtype=np.int64) if mask is None else mask.numpy()
    want, winv = pool_ref.masked_mean_pool(hidden.float().numpy(), m)
    np.testing.assert_allclose(out.cpu().numpy(), want, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(inv.cpu().numpy(), winv, rtol=1e-5)
    np.testing.assert_array_equal(ob.cpu().view(torch.int16).numpy(), out.cpu().to(torch.bfloat16).view(torch.int16).numpy())


def test_pool_all_masked_row_and_empty_batch():
    hidden = torch.randn(3, 9, 16)
    mask = torch.ones(3, 9, dtype=torch.int64)
    mask[1] = 0
    out = _run(hidden, mask).cpu().numpy()
    assert np.all(out[1] == 0) and np.isfinite(out).all()
    assert _run(torch.zeros(0, 4, 16), None).shape == (0, 16)


def test_pool_argument_errors():
    from dab_b200 import DabError
    with pytest.raises(DabError, match="multiple of 4"):
        _run(torch.zeros(1, 2, 6), None)
    with pytest.raises(DabError, match="mask shape"):
        _run(torch.zeros(1, 2, 8), torch.ones(1, 3, dtype=torch.int64))


def test_ingest_pools_and_appends_into_the_owning_shard():
    """The write path as a path (assistant/processing/documents/steps/embeddings.py:27-39): hidden states in,
    rows in the shard out - pool, normalise, append, on the device; then the rows are searchable."""
    from dab_b200 import sharded
    g = torch.Generator().manual_seed(11)
    B, T, D = 48, 40, 256
    hidden = torch.randn(B, T, D, generator=g)
    lengths = torch.randint(1, T + 1, (B,), generator=g)
    mask = (torch.arange(T)[None] < lengths[:, None]).to(torch.int64)
    want, _ = pool_ref.masked_mean_pool(hidden.numpy(), mask.numpy())
    for dtype, tol in ((torch.float32, 2e-6), (torch.bfloat16, 2 ** -8)):
        sh = sharded.ShardedStore(1000, D, dtype, DEV)            # world size 1: this rank owns rows [0, 1000)
        assert sh.ingest(hidden[:20].to(DEV), mask[:20].to(DEV)) == 20
        assert sh.ingest(hidden[20:].to(DEV), mask[20:].to(DEV)) == B - 20 and len(sh.local) == B
        rows, inv, ids = sh.local.export(0, B)
        np.testing.assert_allclose(rows.float().cpu().numpy(), want, rtol=0, atol=tol)
        assert ids.cpu().tolist() == list(range(B))
        s, i = sh.search(torch.from_numpy(want[5:6]).to(DEV), 1)
        assert i[0, 0].item() == 5 and abs(s[0, 0].item() - 1.0) < 1e-3
        sh.close()


@pytest.mark.parametrize("store_dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("hidden_dtype", [torch.float32, torch.bfloat16])
def test_store_ingest_equals_pool_then_append(store_dtype, hidden_dtype):
    """dab_store_ingest (the pool kernel writing straight into the store) leaves exactly what dab_pool_l2norm followed
    by dab_store_append leaves: the same stored rows, the same 1/|row| (taken from the stored values), the same ids."""
    from dab_b200 import ops
    g = torch.Generator().manual_seed(23)
    B, T, D = 37, 33, 768
    hidden = torch.randn(B, T, D, generator=g).to(hidden_dtype).to(DEV)
    lengths = torch.randint(0, T + 1, (B,), generator=g)          # a zero-length text too: a zero row, never ranked
    lengths[3] = 0
    mask = (torch.arange(T)[None] < lengths[:, None]).to(torch.int64).to(DEV)
    a = ops.ChunkStore(100, D, store_dtype, DEV, id_base=40)
    b = ops.ChunkStore(100, D, store_dtype, DEV, id_base=40)
    first = torch.randn(5, D, generator=g).to(DEV)
    for st in (a, b):
        st.append(first)                                           # rows already there: the ingest continues after them
    a_rows_before = a.export(0, 5)[0]
    assert a.ingest(hidden, mask) == B and len(a) == 5 + B
    if store_dtype == torch.bfloat16:
        _, rows = ops.pool_l2norm(hidden, mask, want_bf16=True)
    else:
        rows = ops.pool_l2norm(hidden, mask)
    b.append(rows)
    ra, ia, da = a.export(0, 5 + B)
    rb, ib, db = b.export(0, 5 + B)
    assert torch.equal(ra[5:], rb[5:]) and torch.equal(da, db) and torch.equal(ra[:5], a_rows_before)
    np.testing.assert_allclose(ia.cpu().numpy(), ib.cpu().numpy(), rtol=2e-7, atol=0)
    assert ia[5 + 3].item() == 0.0 and da[-1].item() == 40 + 5 + B - 1
    q = ra[10:11].float()
    sa, ida = a.search(q, 3)
    sb, idb = b.search(q, 3)
    assert torch.equal(ida, idb) and ida[0, 0].item() == 40 + 10
    with pytest.raises(ops.DabError):
        a.ingest(torch.randn(100, T, D, device=DEV), None)          # beyond capacity
    a.close(); b.close()
