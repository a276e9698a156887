"""Small end-to-end run of every kernel, meant to be executed under `compute-sanitizer --tool memcheck`
(after the same command has passed without it).  Sizes are tiny so the instrumented run is quick."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "django-assistant-bot_b200"))
import torch
from dab_b200 import ops

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
# pool
h = torch.randn(5, 37, 384, generator=g)
m = (torch.arange(37)[None] < torch.tensor([1, 37, 20, 5, 11])[:, None]).long()
for dt in (torch.float32, torch.bfloat16):
    ops.pool_l2norm(h.to(dt).to(dev), m.to(dev), want_bf16=True, want_inv_norm=True)
# searches: CUDA-core, tcgen05, tcgen05 cluster kernel; tail tiles, filters
for dtype, path, N, D, Q, k in ((torch.float32, "simt", 5003, 768, 3, 10), (torch.bfloat16, "simt", 4999, 392, 2, 5),
                                (torch.bfloat16, "tcgen05", 5001, 768, 40, 10), (torch.bfloat16, "tcgen05", 5001, 768, 130, 10),
                                (torch.bfloat16, "tcgen05", 700, 1024, 256, 20)):
    rows = torch.randn(N, D, generator=g)
    q = torch.randn(Q, D, generator=g)
    st = ops.ChunkStore(N, D, dtype, dev)
    st.set_path(path)
    st.append(rows.to(dev))
    allow = ops.pack_allow_bitmap((torch.rand(N, generator=g) < 0.5).to(dev))
    st.search(q.to(dev), k)
    st.search(q.to(dev), k, allow)
    st.close()
# merge
ops.merge_topk(torch.rand(4, 3, 50).sort(dim=2, descending=True).values.to(dev), torch.randint(0, 10**6, (4, 3, 50)).to(dev))
torch.cuda.synchronize()
print("sanitize target ok")
