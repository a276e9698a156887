"""The single-process sharded store (dab_sharded_*): G shards behind one call, no launcher, no torch.distributed.
On a one-GPU box the shards share cuda:0 (two or three "ranks" on one device: the publish / wait / pull / merge of
the exchange kernel all run, only the wire is missing); with >= 2 GPUs the same test also runs one shard per device."""
import numpy as np
import pytest
import torch

from oracle import search_ref

pytestmark = pytest.mark.gpu


def _layouts():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    out = [[0], [0, 0], [0, 0, 0]]
    if n >= 2:
        out.append(list(range(min(n, 8))))
    return out


@pytest.mark.parametrize("devices", _layouts(), ids=lambda d: "dev" + "".join(map(str, d)))
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_single_process_sharded_search_matches_oracle(devices, dtype):
    from dab_b200 import sharded
    N, D = 90_001, 256
    g = torch.Generator().manual_seed(17)
    rows = torch.randn(N, D, generator=g)
    rows[N // 2 + 5] = rows[3]                       # duplicate across shards: the lower global id comes first
    sh = sharded.LocalShardedStore(N, D, dtype, devices, max_q=64, max_k=128)
    assert [b[0] for b in sh.bounds] == [i * N // len(devices) for i in range(len(devices))]
    for i, (lo, hi) in enumerate(sh.bounds):
        sh.append(i, rows[lo:hi].to(sh.shards[i].device))
    for s in sh.shards:
        torch.cuda.synchronize(s.device)
    stored = rows.to(dtype).float().numpy()
    for Q, k in ((1, 10), (64, 100), (5, 128)):
        q = torch.randn(Q, D, generator=g)
        q[0] = rows[3]
        scores, ids = sh.search_host(q.numpy(), k)
        S = search_ref.exact_scores(stored, q.numpy())
        for qi in range(Q):
            search_ref.check_topk_parity(S[qi], ids[qi], scores[qi], k)
        assert ids[0, 0] == 3 and ids[0, 1] == N // 2 + 5
    with pytest.raises(Exception):
        sh.search_host(torch.randn(65, D).numpy(), 10)          # beyond max_q
    sh.close()


def test_single_process_ingest_lands_in_the_owning_shard():
    from dab_b200 import sharded
    from oracle import pool_ref
    B, T, D = 32, 24, 128
    g = torch.Generator().manual_seed(4)
    sh = sharded.LocalShardedStore(2 * B, D, torch.float32, [0, 0], max_q=8, max_k=16)
    wants = []
    for i in range(2):
        hidden = torch.randn(B, T, D, generator=g)
        mask = (torch.arange(T)[None] < torch.randint(1, T + 1, (B, 1), generator=g)).to(torch.int64)
        assert sh.ingest(i, hidden.to("cuda:0"), mask.to("cuda:0")) == B
        wants.append(pool_ref.masked_mean_pool(hidden.numpy(), mask.numpy())[0])
    torch.cuda.synchronize()
    want = np.concatenate(wants)
    scores, ids = sh.search_host(want[[5, B + 7]], 1)
    assert ids[:, 0].tolist() == [5, B + 7] and np.allclose(scores[:, 0], 1.0, atol=1e-5)
    sh.close()
