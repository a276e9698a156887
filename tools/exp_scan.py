"""Kernel experiments: time the scan launches of one search shape with an alternative build of the
library (DAB_EXP_LIB=path/to/lib.so), e.g. one compiled with experiment macros.  Results of
experimental builds are timing-only (they may rank wrongly); nothing here is part of the product."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "django-assistant-bot_b200"))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from dab_b200 import _lib  # noqa: E402

if os.environ.get("DAB_EXP_LIB"):
    _lib.LIB_PATH = os.path.abspath(os.environ["DAB_EXP_LIB"])

from dab_b200 import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=12_500_000)
    ap.add_argument("--dim", type=int, default=768)
    ap.add_argument("--q", type=int, default=256)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--path", default="auto")
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    dt = torch.bfloat16 if a.dtype == "bf16" else torch.float32
    dev = "cuda:0"
    st = ops.ChunkStore(a.rows, a.dim, dt, dev)
    g = torch.Generator(device=dev).manual_seed(1)
    step = 1_000_000
    for r0 in range(0, a.rows, step):
        n = min(step, a.rows - r0)
        st.append(torch.randn(n, a.dim, generator=g, device=dev, dtype=torch.float32).to(dt))
    st.set_path(a.path)
    q = torch.randn(a.q, a.dim, generator=g, device=dev)
    for _ in range(3):
        st.search_async(q, a.k)
    torch.cuda.synchronize()
    from bench import ClockSampler
    cs = ClockSampler(0)
    st.set_profile(True)
    st.get_profile()
    cs.start()
    for _ in range(a.reps):
        st.search_async(q, a.k)
    torch.cuda.synchronize()
    clocks = cs.stop()
    p = st.get_profile()
    plan = st.describe_plan(a.q, a.k)
    print(f"lib={os.path.basename(_lib.LIB_PATH)} path={plan.get('path')} scan_ms/search={p['scan_ms'] / a.reps:.3f} "
          f"search_ms={p['search_ms'] / a.reps:.3f} launches/search={p['scan_launches'] / a.reps:.1f} sm_mhz={clocks.get('sm_mhz')} "
          f"reasons={clocks.get('reasons')}", flush=True)


if __name__ == "__main__":
    main()
