"""Pins the oracle against the reference's own outputs (tests/golden, made by oracle/make_golden.py)
and against known-answer vectors.  CPU only."""
import os

import numpy as np
import pytest

from oracle import c_oracle, pool_ref, search_ref


@pytest.fixture(scope="module")
def pool_golden(golden_dir):
    return np.load(os.path.join(golden_dir, "pool_ref.npz"))


@pytest.fixture(scope="module")
def search_golden(golden_dir):
    return np.load(os.path.join(golden_dir, "search_ref.npz"))


# ---------------------------------------------------------------- pool half (pinned)

def test_reference_mean_matches_reference_output(pool_golden):
    """oracle.reference_mean on the per-text hidden states == TransformersEmbedder.embeddings output
    (assistant/ai/embedders/transformers.py:25), 1e-5 relative."""
    g = pool_golden
    off = 0
    for i, L in enumerate(g["lengths"]):
        h = g["per_text_hidden"][off:off + L][None]
        off += L
        np.testing.assert_allclose(pool_ref.reference_mean(h), g["ref_embeddings"][i], rtol=1e-5, atol=1e-6)


def test_masked_pool_of_padded_batch_matches_reference(pool_golden):
    """Known-answer (5): masked pool of the padded batch == per-text reference output."""
    g = pool_golden
    raw, _ = pool_ref.masked_mean_pool(g["hidden"], g["mask"], normalize=False)
    np.testing.assert_allclose(raw, g["ref_embeddings"], rtol=1e-5, atol=2e-6)
    unit, inv = pool_ref.masked_mean_pool(g["hidden"], g["mask"], normalize=True)
    ref_unit = g["ref_embeddings"] / np.linalg.norm(g["ref_embeddings"], axis=1, keepdims=True)
    np.testing.assert_allclose(unit, ref_unit, rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(np.linalg.norm(unit.astype(np.float64), axis=1), 1.0, rtol=1e-6)


def test_pool_c_oracle_matches_numpy(pool_golden):
    g = pool_golden
    want, _ = pool_ref.masked_mean_pool(g["hidden"], g["mask"])
    for strict in (True, False):
        got = c_oracle.pool(g["hidden"], g["mask"], strict=strict, nthreads=2)
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)


def test_pool_edge_cases():
    h = np.random.default_rng(0).standard_normal((3, 7, 16)).astype(np.float32)
    m = np.ones((3, 7), dtype=np.int64)
    m[1] = 0                      # all-masked row -> zero vector, not NaN
    m[2, 3:] = 0
    out, inv = pool_ref.masked_mean_pool(h, m)
    assert np.all(out[1] == 0) and np.isfinite(out).all()
    np.testing.assert_allclose(out[2], (h[2, :3].mean(0) / np.linalg.norm(h[2, :3].mean(0))), rtol=1e-5)
    np.testing.assert_allclose(c_oracle.pool(h, m), out, rtol=1e-5, atol=1e-6)
    import torch
    t = pool_ref.masked_mean_pool_torch_cpu(torch.from_numpy(h), torch.from_numpy(m)).numpy()
    np.testing.assert_allclose(t, out, rtol=1e-5, atol=1e-6)


# ---------------------------------------------------------------- retrieve half

def test_cosine_matches_reference_numpy(search_golden):
    """oracle cosine == reference embeddings_similarity (search_service.py:17-21,32-35)."""
    g = search_golden
    S = search_ref.exact_scores(g["rows"], g["queries"])
    np.testing.assert_allclose(S, g["sims"], rtol=1e-12, atol=1e-12)
    for qi in range(g["queries"].shape[0]):
        for r in (0, 7, 63):
            got = search_ref.cosine_similarity(g["queries"][qi].tolist(), g["rows"][r].tolist())
            assert got == pytest.approx(g["sims"][qi, r], rel=1e-14)


def test_known_answers(search_golden):
    ka = search_golden["known_answers"]
    assert ka[0] == 0.0 and ka[1] == pytest.approx(1.0) and ka[2] == pytest.approx(-1.0)
    assert ka[3] == pytest.approx(ka[4], abs=1e-8)           # scale invariance (fp32 rounding of 3*x only)
    e = np.eye(8, dtype=np.float32)
    assert search_ref.pgvector_cosine_distance(e[0], e[1]) == 1.0
    assert search_ref.pgvector_cosine_distance(e[0], e[0]) == 0.0
    assert search_ref.pgvector_cosine_distance(e[0], -e[0]) == 2.0
    assert np.isnan(search_ref.pgvector_cosine_distance(e[0], np.zeros(8, np.float32)))
    with pytest.raises(ValueError, match="different vector dimensions"):
        search_ref.pgvector_cosine_distance(e[0], np.zeros(4, np.float32))


def test_legacy_ranking_agrees_where_gaps_allow(search_golden):
    """search_service.py:79-82 ranking vs the oracle's (score desc, index asc)."""
    g = search_golden
    S = g["sims"]
    for qi in range(S.shape[0]):
        ids, sc = search_ref.rank_topk(S[qi], 10)
        legacy = g["legacy_order"][qi][:10]
        for pos in range(10):
            if ids[pos] != legacy[pos]:
                assert abs(S[qi, ids[pos]] - S[qi, legacy[pos]]) < 1e-6
        assert np.all(np.diff(sc) <= 0)
    # identical / opposite query rows planted by make_golden
    assert search_ref.rank_topk(S[1], 1)[0][0] == 20 and S[1, 20] == pytest.approx(1.0)
    assert np.argmin(S[2]) == 21 and S[2, 21] == pytest.approx(-1.0)


def test_tie_break_and_duplicates(search_golden):
    g = search_golden
    S = search_ref.exact_scores(g["rows"], g["rows"][3:4])[0]
    ids, _ = search_ref.rank_topk(S, 3)
    assert ids[0] == 3 and ids[1] == 10            # duplicate rows: lower index first
    S7 = search_ref.exact_scores(g["rows"], g["rows"][7:8])[0]
    assert abs(S7[7] - S7[11]) < 1e-12             # scaled copy has the same cosine


def test_c_scan_matches_python_pgvector_statement(search_golden):
    g = search_golden
    rows, q = g["rows"][:12], g["queries"][0]
    want = np.array([search_ref.pgvector_cosine_distance(r, q) for r in rows])
    got = c_oracle.distances(rows, q, strict=True)
    np.testing.assert_array_equal(got, want)       # same sequential fp32 sums -> bit identical
    vec = c_oracle.distances(rows, q, strict=False)
    np.testing.assert_allclose(vec, want, rtol=0, atol=2e-6)


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_c_search_matches_oracle_ranking(dtype):
    rng = np.random.default_rng(7)
    N, D, Q, k = 5000, 96, 5, 17
    rows = rng.standard_normal((N, D)).astype(np.float32)
    rows[100] = 0                                  # zero row -> NaN distance -> never returned
    rows[4000] = rows[12]
    qs = rng.standard_normal((Q, D)).astype(np.float32)
    allow = rng.random(N) < 0.7
    if dtype == "bf16":
        bits = search_ref.f32_to_bf16_bits(rows)
        stored = search_ref.bf16_bits_to_f32(bits)
        crow = bits
    else:
        stored = rows
        crow = rows
    S = search_ref.exact_scores(stored, qs)
    for a in (None, allow):
        ids, dist, cnt = c_oracle.search(crow, qs, k, allow=a, nthreads=3)
        for qi in range(Q):
            assert cnt[qi] == k
            search_ref.check_topk_parity(S[qi], ids[qi], 1.0 - dist[qi], k, allow=a)
            assert 100 not in ids[qi]


def test_search_fewer_rows_than_k():
    rng = np.random.default_rng(1)
    rows = rng.standard_normal((5, 8)).astype(np.float32)
    rows[2] = 0
    ids, dist, cnt = c_oracle.search(rows, rows[0], 10)
    assert cnt[0] == 4 and ids[0, 0] == 0 and (ids[0, 4:] == -1).all()
    top, sc = search_ref.rank_topk(search_ref.exact_scores(rows, rows[0])[0], 10)
    assert list(top) == list(ids[0, :4])


def test_bf16_roundtrip():
    x = np.array([1.0, -2.5, 3.14159, 1e-8, 65504.0, 0.0], dtype=np.float32)
    b = search_ref.f32_to_bf16_bits(x)
    y = search_ref.bf16_bits_to_f32(b)
    assert np.all(np.abs(y - x) <= np.abs(x) * 2.0 ** -8)
    import torch
    tb = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    np.testing.assert_array_equal(b, tb)


def test_doc_level_scores():
    """search_service.py:133-150: docs with < max_scores_n hits are dropped; score 1 - mean(first n)."""
    docs = [1, 2, 1, 3, 2, 1, 2]
    dist = [0.1, 0.2, 0.3, 0.35, 0.4, 0.5, 0.6]
    got = search_ref.doc_level_scores(docs, dist, max_scores_n=2, top_n=5)
    assert [d for d, _ in got] == [1, 2]
    assert got[0][1] == pytest.approx(1 - (0.1 + 0.3) / 2)
    assert got[1][1] == pytest.approx(1 - (0.2 + 0.4) / 2)


from hypothesis import given, settings, strategies as st


@settings(max_examples=25, deadline=None)
@given(st.integers(0, 2**31 - 1))
def test_ranking_invariants(seed):
    """Permutation of rows permutes ids; positive scaling of rows/query leaves the ranking alone."""
    rng = np.random.default_rng(seed)
    N, D, k = 200, 24, 9
    rows = rng.standard_normal((N, D)).astype(np.float32)
    q = rng.standard_normal(D).astype(np.float32)
    S = search_ref.exact_scores(rows, q)[0]
    ids, _ = search_ref.rank_topk(S, k)
    perm = rng.permutation(N)
    S2 = search_ref.exact_scores(rows[perm], q)[0]
    ids2, _ = search_ref.rank_topk(S2, k)
    assert list(perm[ids2]) == list(ids)
    scale = rng.uniform(0.1, 10, size=(N, 1)).astype(np.float32)
    S3 = search_ref.exact_scores(rows * scale, 3.0 * q)[0]
    search_ref.check_topk_parity(S, search_ref.rank_topk(S3, k)[0], S3[search_ref.rank_topk(S3, k)[0]], k)


def test_bf16_golden_rows_regenerate_and_oracle_matches_reference(golden_dir):
    """search_ref_bf16.npz stores only the reference's outputs; the bf16-valued rows come back from the seed
    (checksum pinned) and the oracle's float64 cosine must reproduce the reference's numbers on them."""
    from oracle import make_golden
    g = np.load(os.path.join(golden_dir, "search_ref_bf16.npz"))
    rows = make_golden.golden_bf16_rows()
    assert float(rows.astype(np.float64).sum()) == float(g["row_checksum"][0])
    S = search_ref.exact_scores(rows, g["queries"])
    sims = g["sims"]
    assert np.isnan(sims[:, 200]).all() and np.isnan(S[:, 200]).all()          # the zero row
    ok = ~np.isnan(sims)
    np.testing.assert_allclose(S[ok], sims[ok], rtol=1e-12, atol=1e-12)
    ids, _ = search_ref.rank_topk(sims[0], 50)
    assert ids[:3].tolist() == [9, 150, 151]                                    # triplicate: index order
