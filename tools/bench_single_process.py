"""configs[3] through the SINGLE-PROCESS sharded store (dab_sharded_*): one Python process, no torchrun, all visible
GPUs.  Times LocalShardedStore.search_host (host queries in, merged host results out) for Q = 256 / 64 / 1 and checks
every query against the device oracle.  Prints one JSON line per query count.  Builder-run yardstick for the
torchrun numbers of bench.py (same corpus: seed 1234 + global block index).
usage: python tools/bench_single_process.py [--rows 100000000] [--steps 20] [--devices 0,1,...]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "django-assistant-bot_b200"), os.path.join(ROOT, "tools")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import device_oracle  # noqa: E402
from bench import BLOCK_ROWS, gen_block, gen_queries, load_peaks  # noqa: E402
from dab_b200 import sharded  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--dim", type=int, default=768)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--devices", default="")
    ap.add_argument("--no-verify", action="store_true")
    a = ap.parse_args()
    devices = [int(x) for x in a.devices.split(",")] if a.devices else list(range(torch.cuda.device_count()))
    G, N, D, k = len(devices), a.rows, a.dim, a.k
    peaks = load_peaks()
    sh = sharded.LocalShardedStore(N, D, torch.bfloat16, devices, max_q=256, max_k=k)

    def blocks(g):
        lo, hi = sh.bounds[g]
        dev = sh.shards[g].device
        r = lo
        while r < hi:
            b = r // BLOCK_ROWS
            b0 = b * BLOCK_ROWS
            with torch.cuda.device(dev):
                blk = gen_block(torch, b, min(BLOCK_ROWS, N - b0), D, dev)
                part = blk[r - b0: min(hi, b0 + BLOCK_ROWS) - b0]
            yield r, part
            r += part.shape[0]

    t0 = time.time()
    for g in range(G):
        with torch.cuda.device(sh.shards[g].device):
            for _, part in blocks(g):
                sh.append(g, part)
    for g in range(G):
        torch.cuda.synchronize(sh.shards[g].device)
    build_s = time.time() - t0
    queries = gen_queries(torch, 256, D)
    oracle = None
    for Q in (256, 64, 1):
        q = queries[:Q].numpy()
        for _ in range(5):
            sh.search_host(q, k)
        t0 = time.perf_counter()
        for _ in range(a.steps):
            scores, ids = sh.search_host(q, k)
        ms = (time.perf_counter() - t0) / a.steps * 1e3
        verified = None
        if not a.no_verify:
            if oracle is None:
                li, ls = [], []
                for g in range(G):
                    dev = sh.shards[g].device
                    with torch.cuda.device(dev):
                        wi, ws = device_oracle.oracle_topk(((r, p.to(torch.bfloat16)) for r, p in blocks(g)), queries.to(dev), k, pad=64)
                    li.append(wi.to("cuda:%d" % devices[0]))
                    ls.append(ws.to("cuda:%d" % devices[0]))
                oracle = device_oracle.merge_rank_lists(li, ls, k + 64)
            d0 = "cuda:%d" % devices[0]
            rep = device_oracle.parity_report(torch.from_numpy(scores).to(d0), torch.from_numpy(ids).to(d0), oracle[0][:Q], oracle[1][:Q], k)
            assert rep["ok"], rep
            verified = f"all {Q} queries match the device oracle (worst relative score error {rep['max_rel_score_err']:.1e})"
        per_gpu_bytes = (N // G) * (D * 2 + 4) + Q * D * 4 + Q * k * 12
        print(json.dumps(dict(mode="single process, dab_sharded_search_host", n_gpus=G, devices=devices, rows=N, dim=D, Q=Q, k=k,
                              e2e_ms_per_step=ms, e2e_queries_per_s=Q / ms * 1e3, hbm_frac_e2e=per_gpu_bytes / (ms * 1e-3) / 1e9 / peaks["hbm"],
                              build_s=build_s, verified=verified)), flush=True)
    sh.close()


if __name__ == "__main__":
    main()
