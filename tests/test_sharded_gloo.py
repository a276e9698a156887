"""Host logic of the N>1 path on CPU: world_size 2 over gloo (no GPU needed).
Each rank ranks its row shard with the oracle, the lists are exchanged with the same
gather_topk the GPU path uses, and the merged result must equal the single-table ranking."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import search_ref


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, N, D, Q, k, ret):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "django-assistant-bot_b200"))
    sys.path.insert(0, root)
    from dab_b200 import sharded
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(123)
    rows = rng.standard_normal((N, D)).astype(np.float32)
    rows[N // 2 + 3] = rows[5]                       # a tie that straddles the shard boundary
    queries = rng.standard_normal((Q, D)).astype(np.float32)
    queries[0] = rows[5]
    lo, hi = sharded.shard_bounds(N, world, rank)
    S_local = search_ref.exact_scores(rows[lo:hi], queries)
    ls = np.full((Q, k), -np.inf, dtype=np.float32)
    li = np.full((Q, k), -1, dtype=np.int64)
    for q in range(Q):
        ids, sc = search_ref.rank_topk(S_local[q], k)
        ls[q, :len(ids)] = sc
        li[q, :len(ids)] = ids + lo                  # ids are global row indices
    gs, gi = sharded.gather_topk(torch.from_numpy(ls), torch.from_numpy(li))
    assert gs.shape == (world, Q, k) and gi.shape == (world, Q, k)
    assert torch.equal(gi[rank], torch.from_numpy(li))           # rank-major layout
    ms, mi = search_ref.merge_topk_lists(gs.numpy(), gi.numpy(), k)
    S = search_ref.exact_scores(rows, queries)
    for q in range(Q):
        want_ids, want_sc = search_ref.rank_topk(S[q], k)
        assert list(mi[q]) == list(want_ids), (rank, q)
        np.testing.assert_allclose(ms[q], want_sc, rtol=1e-6)
    assert mi[0, 0] == 5 and mi[0, 1] == N // 2 + 3               # cross-shard tie: lower id first
    dist.barrier()
    dist.destroy_process_group()
    ret.put(rank)


@pytest.mark.timeout(180)
def test_two_rank_gather_and_merge_matches_single_table():
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 4001, 48, 3, 12, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(150)
        assert p.exitcode == 0, f"rank exited with {p.exitcode}"
    assert sorted(ret.get(timeout=5) for _ in range(world)) == [0, 1]


def test_shard_bounds_cover_everything_once():
    from dab_b200 import sharded
    for n in (0, 1, 7, 100, 100_000_001):
        for w in (1, 2, 3, 4, 8):
            b = [sharded.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
    for row in (0, 49, 50, 99):
        g = sharded.owner_of_row(row, 100, 4)
        lo, hi = sharded.shard_bounds(100, 4, g)
        assert lo <= row < hi
