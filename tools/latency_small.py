"""Latency of one checked host-buffer search (dab_store_search_host) on small stores - the reference's
operating point (one query, a few thousand chunks).  Prints the median wall-clock microseconds."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "django-assistant-bot_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dab_b200 import ops  # noqa: E402

for n, d, dt in [(10_000, 384, torch.float32), (10_000, 768, torch.float32), (100_000, 768, torch.float32), (100_000, 768, torch.bfloat16)]:
    st = ops.ChunkStore(n, d, dt, "cuda:0")
    st.append(torch.randn(n, d, device="cuda:0"))
    for Q in (1, 8):
        q = np.random.default_rng(0).standard_normal((Q, d)).astype(np.float32)
        for _ in range(20):
            st.search_host(q, 10)
        ts = []
        for _ in range(300):
            t0 = time.perf_counter()
            st.search_host(q, 10)
            ts.append(time.perf_counter() - t0)
        ts = np.sort(np.asarray(ts)) * 1e6
        print(f"n={n} d={d} {str(dt).split('.')[-1]} Q={Q}: median {ts[150]:.1f} us  p10 {ts[30]:.1f}  p90 {ts[270]:.1f}  plan={st.describe_plan(Q, 10)['path']} levels={st.describe_plan(Q, 10)['num_levels']}", flush=True)
    st.close()
