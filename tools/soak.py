"""Soak test: several host threads hammer two stores (bf16 and fp32) with searches of changing shape on their own
streams for a fixed time; every answer is compared with the answer the same store gave for the same query batch in
a quiet, single-threaded pass (itself checked against the float64 oracle).  Catches rare ordering faults (per-call
verdict slots, scratch hand-over between streams, programmatic dependent launch) that the unit tests would only hit
by luck.  usage: python tools/soak.py [seconds] [threads]"""
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "django-assistant-bot_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from dab_b200 import ops  # noqa: E402
from oracle import search_ref  # noqa: E402

DEV = "cuda:0"


def main():
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    nthreads = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    g = torch.Generator().manual_seed(99)
    stores, cases = [], []
    for dtype, N, D in ((torch.bfloat16, 300_000, 768), (torch.float32, 120_000, 384)):
        rows = torch.randn(N, D, generator=g)
        st = ops.ChunkStore(N, D, dtype, DEV)
        st.append(rows.to(DEV))
        stored = rows.to(dtype).float().numpy()
        for Q, k in ((1, 5), (3, 10), (17, 100), (64, 100), (130, 10), (256, 50)):
            q = torch.randn(Q, D, generator=g)
            s, i = st.search(q.to(DEV), k)
            S = search_ref.exact_scores(stored, q.numpy())
            for qi in range(0, Q, max(1, Q // 8)):
                search_ref.check_topk_parity(S[qi], i[qi].cpu().numpy(), s[qi].cpu().numpy(), k)
            cases.append((st, q.to(DEV), k, s.clone(), i.clone()))
        stores.append(st)
    torch.cuda.synchronize()
    stop = time.time() + seconds
    counts = [0] * nthreads
    errors = []

    def worker(t):
        rng = np.random.default_rng(t)
        stream = torch.cuda.Stream(device=DEV)
        with torch.cuda.stream(stream):
            while time.time() < stop and not errors:
                st, q, k, want_s, want_i = cases[int(rng.integers(len(cases)))]
                try:
                    if rng.random() < 0.3:
                        hs, hi = st.search_host(q.cpu().numpy(), k)
                        ok = np.array_equal(hi, want_i.cpu().numpy()) and np.array_equal(hs, want_s.cpu().numpy())
                    else:
                        s, i = st.search(q, k)
                        ok = torch.equal(i, want_i) and torch.equal(s, want_s)
                    if not ok:
                        errors.append(f"thread {t}: answer differs (Q={q.shape[0]}, k={k}, dtype={st.dtype})")
                except Exception as e:  # noqa: BLE001
                    errors.append(f"thread {t}: {e!r}")
                counts[t] += 1

    ths = [threading.Thread(target=worker, args=(t,)) for t in range(nthreads)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    print(f"soak: {sum(counts)} searches from {nthreads} threads in {seconds:.0f} s, errors: {errors[:3] if errors else 'none'}")
    sys.exit(1 if errors else 0)


if __name__ == "__main__":
    main()
