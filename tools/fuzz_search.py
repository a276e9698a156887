"""Randomised parity sweep of the search on one GPU: random store types, row lengths, sizes, query counts, k and
row filters, each checked query by query with the oracle's parity rule against float64 scores computed by torch on
the GPU from the stored values.  A tool for the GPU box (the committed tests cover fixed shapes); prints the first
failing configuration with its seed.  usage: python tools/fuzz_search.py [cases] [seed]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "django-assistant-bot_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dab_b200 import ops  # noqa: E402
from oracle import search_ref  # noqa: E402

DEV = "cuda:0"


def one_case(rng: np.random.Generator, case: int) -> str:
    dtype = torch.bfloat16 if rng.random() < 0.5 else torch.float32
    step = 8 if dtype == torch.bfloat16 else 4
    D = int(rng.choice([step * int(rng.integers(1, 200)), 64, 128, 384, 768, 1024]))
    Q = int(rng.choice([1, 2, 3, 4, 5, 16, 33, 64, 65, 96, 100, 128, 129, 192, 200, 256, int(rng.integers(1, 257))]))
    N = int(rng.choice([1, 7, 300, 4097, int(rng.integers(1, 20_000)), int(rng.integers(20_000, 400_000))]))
    while N * Q > 30_000_000 or N * D > 200_000_000:
        N = max(1, N // 2)
    k = int(rng.choice([1, 5, 10, 100, 250, int(rng.integers(1, 400))]))
    use_filter = rng.random() < 0.25
    desc = f"case {case}: dtype={str(dtype).split('.')[-1]} N={N} D={D} Q={Q} k={k} filter={use_filter}"
    g = torch.Generator(device=DEV).manual_seed(int(rng.integers(1 << 30)))
    rows = torch.randn(N, D, generator=g, device=DEV)
    if N > 10 and rng.random() < 0.5:                        # duplicates, a zero row, a scaled copy
        rows[N // 2] = rows[1]
        rows[N // 3] = 0
        rows[N - 1] = 3.0 * rows[2]
    queries = torch.randn(Q, D, generator=g, device=DEV)
    if N > 3 and rng.random() < 0.5:
        queries[0] = rows[1]
    store = ops.ChunkStore(N, D, dtype, DEV)
    store.append(rows)
    allow_np, bitmap = None, None
    if use_filter:
        allow = torch.rand(N, generator=g, device=DEV) < float(rng.choice([0.02, 0.5, 0.9]))
        allow_np = allow.cpu().numpy()
        bitmap = ops.pack_allow_bitmap(allow)
    scores, ids = store.search(queries, k, bitmap)
    plan = store.describe_plan(Q, k)["path"]
    stored = rows.to(dtype).double()
    qn = queries.double()
    qn = qn / qn.norm(dim=1, keepdim=True)
    nrm = stored.norm(dim=1)
    S = (qn @ stored.T) / torch.where(nrm > 0, nrm, torch.ones_like(nrm))[None, :]
    S[:, nrm == 0] = float("nan")
    S = S.cpu().numpy()
    ids, scores = ids.cpu().numpy(), scores.cpu().numpy()
    for qi in range(Q):
        try:
            search_ref.check_topk_parity(S[qi], ids[qi], scores[qi], k, allow=allow_np)
        except AssertionError as e:
            raise AssertionError(f"{desc} path={plan} query {qi}: {e}") from None
    store.close()
    return f"{desc} path={plan} ok"


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rng = np.random.default_rng(seed)
    t0 = time.time()
    for c in range(cases):
        msg = one_case(rng, c)
        if c % 10 == 0:
            print(msg, flush=True)
    print(f"{cases} cases passed in {time.time() - t0:.1f} s (seed {seed})")


if __name__ == "__main__":
    main()
