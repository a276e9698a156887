"""Pool kernel bandwidth against batch size and mask density (CUDA events around 50 back-to-back calls)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "django-assistant-bot_b200"))
import torch  # noqa: E402

from dab_b200 import ops  # noqa: E402

dev = "cuda:0"
T, D = 512, 1024
for B, dense in [(256, False), (256, True), (1024, False), (1024, True), (2048, False)]:
    h = torch.randn(B, T, D, device=dev)
    lens = torch.full((B,), T, device=dev) if dense else torch.randint(1, T + 1, (B,), device=dev)
    mask = (torch.arange(T, device=dev)[None, :] < lens[:, None]).to(torch.int64)
    live = int(mask.sum())
    for _ in range(5):
        ops.pool_l2norm(h, mask)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 50
    e0.record()
    for _ in range(reps):
        ops.pool_l2norm(h, mask)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    gb = (live * D * 4 + B * T * 8 + B * D * 4) / 1e9
    print(f"B={B} dense={dense}: {ms * 1e3:.1f} us  {gb / (ms * 1e-3):.0f} GB/s over {gb:.3f} GB", flush=True)
    del h, mask
