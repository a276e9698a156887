"""Parity at the BASELINE.json sizes, inside ``pytest -m gpu`` ("two-level oracle").

The NumPy oracle (``oracle/search_ref.py``) cannot rank 10^7..10^8 rows, so these tests use the device oracle
(``tools/device_oracle.py``: plain torch, fp32 candidate pass with TF32 off + float64 exact pass +
lexicographic tie-break) - and the first test pins that device oracle to the NumPy oracle on a subsample, so
the chain is  reference semantics -> oracle/search_ref.py -> device oracle -> CUDA path at full size.

Shapes: configs[1] (1M x 768 fp32, Q=1, k=10: 3 sampling levels), configs[2] (10M x 1024 bf16, Q=64, k=100: 5
levels) and a 50M x 768 bf16 shard with Q=256, k=100 (the same 6 levels as the 100M-row configs[3]); EVERY
query is checked.  Each corpus carries the tie-break set of SURVEY 8(d): 1024 rows duplicated at random higher
indices (seed 99), with the first queries aimed at duplicated rows so the ties land inside the top-k.
"""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

from oracle import search_ref  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
BLOCK = 1_000_000


class Corpus:
    """Synthetic corpus generated block by block (never materialised on the host): rows ~ N(0,1)^D normalised,
    seed 1234 + block; 1024 source rows of block 0 are copied over random higher rows (seed 99)."""

    def __init__(self, n, dim, dtype, n_dup=1024):
        self.n, self.dim, self.dtype = n, dim, dtype
        rng = np.random.default_rng(99)
        n_dup = min(n_dup, n // 4)
        self.src = np.sort(rng.choice(min(n, BLOCK) // 2, size=n_dup, replace=False))
        lo = min(n, BLOCK) // 2
        dst = np.unique(rng.integers(lo, n, size=2 * n_dup))
        self.dst = np.sort(rng.permutation(dst)[:n_dup])
        self.src = self.src[:len(self.dst)]
        self.src_rows = None

    def _ensure_src(self):
        if self.src_rows is None:
            self.src_rows = self._raw(0)[torch.as_tensor(self.src, device=DEV)].clone()

    def _raw(self, b):
        g = torch.Generator(device=DEV).manual_seed(1234 + b)
        rows = min(BLOCK, self.n - b * BLOCK)
        x = torch.randn(rows, self.dim, generator=g, device=DEV)
        return torch.nn.functional.normalize(x, dim=1)

    def blocks(self):
        """(first row, rows [n, D] in the store dtype) - exactly what was appended to the store."""
        self._ensure_src()
        for b in range((self.n + BLOCK - 1) // BLOCK):
            x = self._raw(b)
            r0 = b * BLOCK
            sel = (self.dst >= r0) & (self.dst < r0 + x.shape[0])
            if sel.any():
                x[torch.as_tensor(self.dst[sel] - r0, device=DEV)] = self.src_rows[torch.as_tensor(np.nonzero(sel)[0], device=DEV)]
            yield r0, x.to(self.dtype)
            del x

    def queries(self, Q):
        g = torch.Generator(device=DEV).manual_seed(4321)
        q = torch.nn.functional.normalize(torch.randn(Q, self.dim, generator=g, device=DEV), dim=1)
        # aim the first queries at duplicated rows: both copies tie at the top of the list
        self._ensure_src()
        m = min(Q, 8, len(self.src))
        noise = 0.05 * torch.randn(m, self.dim, generator=g, device=DEV) / self.dim ** 0.5
        q[:m] = torch.nn.functional.normalize(self.src_rows[:m] + noise, dim=1)
        return q.contiguous()


def _fill(store, corpus):
    for _, rows in corpus.blocks():
        store.append(rows)
    torch.cuda.synchronize()


def _check_against_device_oracle(corpus, queries, k, scores, ids, min_tied_queries):
    import device_oracle
    wi, ws = device_oracle.oracle_topk(corpus.blocks(), queries, k, pad=64)
    rep = device_oracle.parity_report(scores, ids, wi, ws, k)
    assert rep["ok"], rep
    # exact ties (the planted duplicates produce bit-identical scores) must come out in id order
    s, i = scores.cpu().numpy(), ids.cpu().numpy()
    tied = (s[:, 1:] == s[:, :-1]) & (i[:, 1:] >= 0)
    assert (i[:, 1:][tied] > i[:, :-1][tied]).all(), "equal scores not in ascending id order"
    assert int(tied.any(axis=1).sum()) >= min_tied_queries, "the planted duplicates did not reach the top-k"
    return rep


def test_device_oracle_is_pinned_to_the_numpy_oracle():
    """Level 2 against level 1 on a subsample the NumPy oracle finishes in seconds: identical ids for every query
    (ties included) and float64-equal scores, for a bf16 and an fp32 corpus."""
    import device_oracle
    for dtype, n, dim in ((torch.bfloat16, 300_000, 768), (torch.float32, 200_000, 768)):
        corpus = Corpus(n, dim, dtype)
        q = corpus.queries(12)
        k = 100
        wi, ws = device_oracle.oracle_topk(((r0, x) for r0, x in corpus.blocks()), q, k, pad=64)
        stored = torch.cat([x for _, x in corpus.blocks()]).float().cpu().numpy()
        S = search_ref.exact_scores(stored, q.cpu().numpy())
        wi, ws = wi.cpu().numpy(), ws.cpu().numpy()
        for qi in range(q.shape[0]):
            ids, sc = search_ref.rank_topk(S[qi], k)
            np.testing.assert_array_equal(wi[qi, :k], ids)
            np.testing.assert_allclose(ws[qi, :k], sc, rtol=0, atol=1e-12)


@pytest.mark.parametrize("name,n,dim,dtype,Q,k,levels", [
    ("c2", 1_000_000, 768, torch.float32, 1, 10, 3),
    ("c3", 10_000_000, 1024, torch.bfloat16, 64, 100, 5),
    ("c4-shard", 50_000_000, 768, torch.bfloat16, 256, 100, 6),
])
def test_baseline_config_every_query_matches_the_device_oracle(name, n, dim, dtype, Q, k, levels):
    from dab_b200 import ops
    corpus = Corpus(n, dim, dtype)
    store = ops.ChunkStore(n, dim, dtype, DEV)
    _fill(store, corpus)
    assert store.describe_plan(Q, k)["num_levels"] == levels
    q = corpus.queries(Q)
    scores, ids = store.search(q, k)
    _check_against_device_oracle(corpus, q, k, scores, ids, min_tied_queries=min(Q, 8))
    if name == "c4-shard":
        # the same store at the north_star query counts (single-CTA tensor kernel / pair kernel shapes)
        for q_small in (64, 1):
            s2, i2 = store.search(q[:q_small].contiguous(), k)
            np.testing.assert_array_equal(i2.cpu().numpy(), ids[:q_small].cpu().numpy())
            np.testing.assert_array_equal(s2.cpu().numpy(), scores[:q_small].cpu().numpy())
        # and through the host-buffer entry point
        s3, i3 = store.search_host(q.cpu().numpy(), k)
        np.testing.assert_array_equal(i3, ids.cpu().numpy())
    store.close()
    torch.cuda.empty_cache()


def test_fuzz_slice_of_200_random_configurations():
    """A fixed-seed slice of tools/fuzz_search.py (random store type, row length, N, Q, k, filters, duplicates,
    zero rows; float64 scores from the stored values) so the driver's GPU run exercises it."""
    import fuzz_search
    rng = np.random.default_rng(20261107)
    for c in range(200):
        fuzz_search.one_case(rng, c)
