"""dab_exchange_merge_topk on one GPU (world size 1: publish, wait and gather all hit the local block)."""
import ctypes
import time

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_exchange_single_rank_roundtrip_and_epochs():
    from dab_b200 import _lib
    L = _lib.lib()
    R, max_q, max_k = 1, 16, 32
    nbytes = L.dab_exchange_bytes(R, max_q, max_k)
    assert nbytes > 0
    block = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
    bases = torch.tensor([block.data_ptr()], dtype=torch.int64, device=DEV)
    status_in = torch.tensor([0, 1], dtype=torch.int32, device=DEV)
    status_out = torch.zeros(2, dtype=torch.int32, device=DEV)
    g = torch.Generator().manual_seed(0)
    for epoch in (1, 2, 3):
        Q, k = 5, 20
        sc = torch.rand(Q, k, generator=g).sort(dim=1, descending=True).values
        ids = torch.randperm(Q * k, generator=g).view(Q, k)
        ids[2, 15:] = -1
        sc[2, 15:] = float("-inf")
        ps, pi = ctypes.c_void_p(), ctypes.c_void_p()
        assert L.dab_exchange_slot(block.data_ptr(), R, max_q, max_k, epoch, ctypes.byref(ps), ctypes.byref(pi)) == 0
        off_s, off_i = ps.value - block.data_ptr(), pi.value - block.data_ptr()
        block[off_s: off_s + Q * k * 4].view(torch.float32).view(Q, k).copy_(sc)
        block[off_i: off_i + Q * k * 8].view(torch.int64).view(Q, k).copy_(ids)
        out_s = torch.empty(Q, k, device=DEV)
        out_i = torch.empty(Q, k, dtype=torch.int64, device=DEV)
        rc = L.dab_exchange_merge_topk(bases.data_ptr(), 0, R, max_q, max_k, Q, k, epoch, status_in.data_ptr(),
                                       out_s.data_ptr(), out_i.data_ptr(), status_out.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream)
        assert rc == 0, L.dab_last_error()
        np.testing.assert_array_equal(out_i.cpu().numpy(), ids.numpy())
        np.testing.assert_array_equal(out_s.cpu().numpy(), sc.numpy())
        assert status_out.cpu().tolist() == [0, 1]
    assert L.dab_exchange_merge_topk(bases.data_ptr(), 0, R, max_q, max_k, 17, 5, 4, None, out_s.data_ptr(), out_i.data_ptr(),
                                     None, None) == _lib.DAB_E_CAPACITY
    assert L.dab_exchange_merge_topk(bases.data_ptr(), 0, R, max_q, max_k, 5, 5, 0, None, out_s.data_ptr(), out_i.data_ptr(),
                                     None, None) == _lib.DAB_E_INVALID


def test_exchange_two_ranks_on_one_gpu_publish_wait_pull():
    """The R > 1 path on a single GPU: two half-stores and two staging blocks on cuda:0 play ranks 0 and 1, each
    on its own stream; both launch dab_exchange_merge_topk, so publish (release stores into the peer's block),
    wait (acquire loads until the peer's epoch arrives) and pull (loads from the peer's block) all run for real -
    only the wire is missing.  Checked against the NumPy oracle over the whole corpus, for three epochs (both
    staging parities) and with rank 1 enqueued long after rank 0 is already spinning."""
    from dab_b200 import _lib, ops
    from oracle import search_ref
    L = _lib.lib()
    R, max_q, max_k = 2, 32, 64
    N, D = 120_000, 256
    g = torch.Generator().manual_seed(5)
    rows = torch.randn(N, D, generator=g)
    rows[N // 2 + 7] = rows[11]                      # a duplicate across the shard boundary: id order decides
    half = N // 2
    stores = [ops.ChunkStore(half, D, torch.bfloat16, DEV, id_base=r * half) for r in range(R)]
    for r in range(R):
        stores[r].append(rows[r * half:(r + 1) * half].to(DEV))
    stored = rows.to(torch.bfloat16).float().numpy()
    nbytes = L.dab_exchange_bytes(R, max_q, max_k)
    blocks = [torch.zeros(nbytes, dtype=torch.uint8, device=DEV) for _ in range(R)]
    bases = torch.tensor([b.data_ptr() for b in blocks], dtype=torch.int64, device=DEV)
    streams = [torch.cuda.Stream(device=DEV) for _ in range(R)]
    torch.cuda.synchronize()
    for epoch, (Q, k) in enumerate([(5, 20), (32, 64), (1, 10)], start=1):
        queries = torch.randn(Q, D, generator=g)
        queries[0] = rows[11]
        qd = queries.to(DEV)
        outs = []
        torch.cuda.synchronize()
        for r in range(R):
            with torch.cuda.stream(streams[r]):
                ps, pi = ctypes.c_void_p(), ctypes.c_void_p()
                assert L.dab_exchange_slot(blocks[r].data_ptr(), R, max_q, max_k, epoch, ctypes.byref(ps), ctypes.byref(pi)) == 0
                off_s, off_i = ps.value - blocks[r].data_ptr(), pi.value - blocks[r].data_ptr()
                slot = (blocks[r][off_s: off_s + Q * k * 4].view(torch.float32).view(Q, k),
                        blocks[r][off_i: off_i + Q * k * 8].view(torch.int64).view(Q, k))
                if r == 1:
                    time.sleep(0.05)                 # rank 0's kernel is already on the GPU, waiting for this rank's flag
                stores[r].search_async(qd, k, out=slot)
                out_s = torch.empty(Q, k, device=DEV)
                out_i = torch.empty(Q, k, dtype=torch.int64, device=DEV)
                st = torch.zeros(2, dtype=torch.int32, device=DEV)
                rc = L.dab_exchange_merge_topk(bases.data_ptr(), r, R, max_q, max_k, Q, k, epoch,
                                               L.dab_store_status_dev(stores[r]._h), out_s.data_ptr(), out_i.data_ptr(),
                                               st.data_ptr(), streams[r].cuda_stream)
                assert rc == 0, L.dab_last_error()
                outs.append((out_s, out_i, st))
        torch.cuda.synchronize()
        S = search_ref.exact_scores(stored, queries.numpy())
        for r in range(R):
            out_s, out_i, st = outs[r]
            assert st.cpu().tolist() == [0, 0]
            for qi in range(Q):
                search_ref.check_topk_parity(S[qi], out_i[qi].cpu().numpy(), out_s[qi].cpu().numpy(), k)
        assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][0], outs[1][0])
        assert outs[0][1][0, 0].item() == 11 and outs[0][1][0, 1].item() == N // 2 + 7
    for s in stores:
        s.close()
