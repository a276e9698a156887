"""dab_exchange_merge_topk on one GPU (world size 1: publish, wait and gather all hit the local block)."""
import ctypes

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
