"""The C5 ingest batch (256 x 512 x 1024 fp32, lengths U{16..512}, seed 555) through the pool kernel: time per call."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "django-assistant-bot_b200"))
import torch  # noqa: E402
from dab_b200 import _lib  # noqa: E402
if os.environ.get("DAB_EXP_LIB"):
    _lib.LIB_PATH = os.path.abspath(os.environ["DAB_EXP_LIB"])
from dab_b200 import ops  # noqa: E402
dev = "cuda:0"
B, T, D = 256, 512, 1024
hs = [torch.randn(B, T, D, device=dev) for _ in range(4)]
lengths = torch.randint(16, T + 1, (B,), generator=torch.Generator().manual_seed(555))
mask = (torch.arange(T)[None] < lengths[:, None]).to(torch.int64).to(dev)
for i in range(8):
    ops.pool_l2norm(hs[i % 4], mask, want_bf16=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 200
e0.record()
for i in range(reps):
    ops.pool_l2norm(hs[i % 4], mask, want_bf16=True)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / reps * 1e3
gb = (int(mask.sum()) * D * 4 + B * T * 8 + B * D * 6) / 1e9
print(f"lib={os.path.basename(_lib.LIB_PATH)} splits={os.environ.get('DAB_POOL_SPLITS', 'auto')}: {us:.1f} us  {gb / us * 1e6:.0f} GB/s")
