#!/usr/bin/env python
"""bench.py - the headline measurement: exact cosine top-k queries/sec over the device store.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c4q64|c5] [--impl reference]

One "step" = one pass of the hot path over one batch of synthetic input: Q queries scored against
the whole (sharded) corpus, local top-k, all-gather + k-way merge when N > 1.  Rank 0 prints ONE
JSON line.  For N > 1 the driver launches this under torch.distributed.run (one rank per GPU, NCCL).

  value      whole-job queries/sec with queries already resident in HBM (CUDA events, max over ranks)
  e2e        the same through the public API with HOST buffers: pinned queries -> H2D -> checked
             search (+ merge) -> D2H of scores/ids, every step
  roofline   the scan kernel: algorithmic bytes (or flops) / its own event-timed duration, against
             MEASURED_PEAKS.json
  cpu_baseline  the oracle (oracle/scan_ref.c, pgvector `<=>` restatement, OpenMP over host cores) on a
             bounded sample of the same workload, extrapolated linearly in rows - a reported
             baseline, not the target
`--impl reference` times that CPU restatement as the reference arm (the reference delegates this
path to PostgreSQL+pgvector, unavailable on this box; kind "port").
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "django-assistant-bot_b200"), os.path.join(ROOT, "tools")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

# BASELINE.json configs (index 1..4); SURVEY.md section 8(d) shapes
WORKLOADS = {
    "c2": dict(kind="search", rows=1_000_000, dim=768, dtype="f32", Q=1, k=10,
               desc="1M x 768 fp32 store, Q=1, top-10 (BASELINE.json configs[1])"),
    "c3": dict(kind="search", rows=10_000_000, dim=1024, dtype="bf16", Q=64, k=100,
               desc="10M x 1024 bf16 store, Q=64, top-100, exact fp32 rescoring (configs[2])"),
    "c4": dict(kind="search", rows=100_000_000, dim=768, dtype="bf16", Q=256, k=100,
               desc="100M x 768 bf16 store row-sharded over the GPUs, Q=256, top-100 (configs[3])"),
    "c4q64": dict(kind="search", rows=100_000_000, dim=768, dtype="bf16", Q=64, k=100,
                  desc="100M x 768 bf16 store row-sharded over the GPUs, Q=64, top-100 (north_star target)"),
    "f32q32": dict(kind="search", rows=4_000_000, dim=768, dtype="f32", Q=32, k=10,
                   desc="4M x 768 fp32 store (what pgvector holds), Q=32, top-10: kind::tf32 tensor pass + fp32 rescoring"),
    "c5": dict(kind="pool", B=256, T=512, dim=1024, dtype="f32",
               desc="ingest: masked mean-pool + L2 norm, 256 chunks x 512 tok x 1024-d fp32 per step per GPU (configs[4])"),
}
# The metric is quoted "at 1/2/4/8 B200": that is configs[3] (100M x 768 bf16 sharded over the GPUs, Q=256,
# top-100), which fits one 180 GB B200 (153.6 GB of rows), so it is the N=1 workload too (strong scaling).
DEFAULT_WORKLOAD = "c4"
BLOCK_ROWS = 1_000_000


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), tf=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    tf_burst=float(d["bf16_tflops"]), source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf=1400.0, tf_burst=1590.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "200"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def gen_block(torch, block_idx: int, rows: int, dim: int, device):
    """Rows ~ N(0,1)^dim, L2-normalised in fp32; seed depends only on the GLOBAL block index, so
    the corpus is the same however many GPUs it is sharded over."""
    g = torch.Generator(device=device).manual_seed(1234 + block_idx)
    x = torch.randn(rows, dim, generator=g, device=device, dtype=torch.float32)
    return torch.nn.functional.normalize(x, dim=1)


def gen_queries(torch, Q: int, dim: int):
    g = torch.Generator().manual_seed(4321)
    return torch.nn.functional.normalize(torch.randn(Q, dim, generator=g), dim=1)


def cpu_search_sample(w, sample_rows_np, queries_np, budget_s: float, threads: int):
    """Times oracle/scan_ref.c (vectorised build, OpenMP) on `sample_rows_np`; returns
    (queries/sec extrapolated to the full row count, description)."""
    from oracle import c_oracle
    n_s = sample_rows_np.shape[0]
    t0 = time.perf_counter()
    c_oracle.search(sample_rows_np, queries_np[:1], w["k"], nthreads=threads)
    t1 = time.perf_counter() - t0
    nq = int(max(1, min(queries_np.shape[0], budget_s / max(t1, 1e-6))))
    reps = 1 if nq > 1 else int(max(1, min(200, budget_s / max(t1, 1e-6))))
    t0 = time.perf_counter()
    for _ in range(reps):
        c_oracle.search(sample_rows_np, queries_np[:nq], w["k"], nthreads=threads)
    dt = (time.perf_counter() - t0) / (reps * nq)          # seconds per query on the sample
    per_query_full = dt * (w["rows"] / n_s)
    desc = (f"{nq} quer{'y' if nq == 1 else 'ies'} x {reps} rep(s) over the first {n_s} rows "
            f"({per_query_full * 1e3:.1f} ms/query after linear extrapolation to {w['rows']} rows)")
    return 1.0 / per_query_full, desc


def run_reference_arm(args, w, rank, world):
    """--impl reference: the CPU restatement of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    import numpy as np
    import torch
    from oracle import c_oracle, search_ref
    c_oracle.build()
    threads = os.cpu_count() or 1
    if w["kind"] == "pool":
        from oracle import pool_ref
        torch.manual_seed(777)
        Bs = 32
        hidden = torch.randn(Bs, w["T"], w["dim"])
        lengths = torch.randint(16, w["T"] + 1, (Bs,), generator=torch.Generator().manual_seed(555))
        mask = (torch.arange(w["T"])[None] < lengths[:, None]).long()
        times = []
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            pool_ref.masked_mean_pool_torch_cpu(hidden, mask)
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
        val = Bs / statistics.mean(times)
        line = dict(metric="ingest pool+L2norm chunks/sec", value=val, unit="chunks/s", n_gpus=args.gpus, steps=args.steps,
                    warmup=args.warmup, ms_per_step=1e3 * statistics.mean(times) * w["B"] / Bs, higher_is_better=True,
                    scaling="weak", vs_baseline=None, dtype="f32", data="synthetic", impl="reference",
                    config={"workload": args.workload, "desc": w["desc"]},
                    cpu_baseline=dict(value=val, unit="chunks/s", cores=torch.get_num_threads(), kind="port",
                                      sample=f"{Bs} chunks per step, CPU torch masked mean + F.normalize"),
                    e2e=dict(value=val, unit="chunks/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
        print(json.dumps(line))
        return
    n_s = min(w["rows"], 250_000)
    rng = np.random.default_rng(1234)
    rows = rng.standard_normal((n_s, w["dim"]), dtype=np.float32)
    rows /= np.linalg.norm(rows, axis=1, keepdims=True)
    if w["dtype"] == "bf16":
        rows = search_ref.f32_to_bf16_bits(rows)
    q = gen_queries(torch, w["Q"], w["dim"]).numpy()
    # one step = a bounded sample: up to 4 queries over n_s rows
    nq = min(4, w["Q"])
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        c_oracle.search(rows, q[:nq], w["k"], nthreads=threads)
        if i >= args.warmup:
            times.append(time.perf_counter() - t0)
    step_s = statistics.mean(times)
    per_query_full = step_s / nq * (w["rows"] / n_s)
    val = 1.0 / per_query_full
    sample = (f"per step {nq} queries over a {n_s}-row sample ({step_s * 1e3:.1f} ms measured per step); per-query time scaled "
              f"linearly to {w['rows']} rows = {per_query_full * 1e3:.0f} ms/query, i.e. {per_query_full * w['Q']:.1f} s for the "
              f"{w['Q']}-query batch (pgvector scans one query at a time: Q queries = Q passes)")
    # ms_per_step is what one step of THIS run took (the bounded sample); `value` is the extrapolation
    line = dict(metric="exact cosine top-k queries/sec", value=val, unit="queries/s", n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=step_s * 1e3, higher_is_better=True, scaling="strong",
                vs_baseline=None, dtype="f32", data="synthetic", impl="reference",
                config=search_config(args, w, w["rows"] // max(1, args.gpus)),
                engine=dict(path="cpu", kernel_path="oracle/scan_ref.c (OpenMP)", exchange="none",
                            arithmetic="fp32 running sums, double division (pgvector <=> restatement)"),
                extrapolated=dict(rows_sampled=n_s, queries_sampled=nq, rows_full=w["rows"], ms_per_full_step=per_query_full * w["Q"] * 1e3),
                cpu_baseline=dict(value=val, unit="queries/s", cores=threads, kind="port", sample=sample),
                e2e=dict(value=val, unit="queries/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


def search_config(args, w, n_local):
    """The `config` object of a search line: the workload and nothing else, identical for our arm and the
    reference arm (how each arm computes it is in the separate `engine` object)."""
    elem = 2 if w["dtype"] == "bf16" else 4
    return {"workload": args.workload, "desc": w["desc"], "rows": w["rows"], "rows_per_gpu": n_local, "dim": w["dim"],
            "store_dtype": w["dtype"], "Q": w["Q"], "k": w["k"],
            "l2": "inputs larger than L2: every step streams the whole shard "
                  f"({n_local * w['dim'] * elem / 1e9:.2f} GB >> 126 MB L2)"}


class SearchRun:
    """One (sharded) store of a search workload plus the measurements taken on it.  Every figure of the
    JSON line - headline and `targets` alike - comes from `measure`, i.e. from identical timing code."""

    def __init__(self, env, w, exchange="auto", path="auto"):
        self.env, self.w = env, w
        torch, dev = env.torch, env.dev
        from dab_b200 import sharded
        self.N, self.D, self.k = w["rows"], w["dim"], w["k"]
        self.sdtype = torch.bfloat16 if w["dtype"] == "bf16" else torch.float32
        self.elem = 2 if w["dtype"] == "bf16" else 4
        self.lo, self.hi = env.rank * self.N // env.world, (env.rank + 1) * self.N // env.world
        self.sh = sharded.ShardedStore(self.N, self.D, self.sdtype, dev, exchange=exchange, max_q=max(w["Q"], 1),
                                       max_k=max(self.k, 1))
        assert (self.sh.lo, self.sh.hi) == (self.lo, self.hi)
        self.store = self.sh.local
        self.store.set_path(path)
        self.exchange = "none" if env.world == 1 else ("p2p" if self.sh.peer is not None else "nccl")
        self.sample_host = None
        self.queries_host = gen_queries(torch, w["Q"], self.D).pin_memory()
        self.queries_dev = self.queries_host.to(dev)
        self._oracle = None

    def blocks(self):
        """(first global row, rows in the store dtype) of this rank's shard, regenerated block by block."""
        torch = self.env.torch
        r = self.lo
        while r < self.hi:
            b = r // BLOCK_ROWS
            b0 = b * BLOCK_ROWS
            blk = gen_block(torch, b, min(BLOCK_ROWS, self.N - b0), self.D, self.env.dev)
            part = blk[r - b0: min(self.hi, b0 + BLOCK_ROWS) - b0]
            yield r, part
            r += part.shape[0]
            del blk, part

    def build(self, keep_cpu_sample=False):
        import numpy as np
        torch = self.env.torch
        for r, part in self.blocks():
            self.store.append(part)
            if keep_cpu_sample and self.sample_host is None:
                s = part[:min(part.shape[0], 250_000)]
                self.sample_host = (s.to(torch.bfloat16).view(torch.int16).cpu().numpy().view(np.uint16)
                                    if self.sdtype == torch.bfloat16 else s.cpu().numpy())
        torch.cuda.synchronize()

    def close(self):
        self.sh.close()
        self._oracle = None
        self.env.torch.cuda.empty_cache()

    # ------------------------------------------------------------------ ground truth (not timed)
    def oracle(self):
        """Device oracle (tools/device_oracle.py: fp32 candidate pass + float64 exact pass, plain torch) for ALL
        queries of the workload over the whole corpus; computed once, shared by every query count measured."""
        if self._oracle is None:
            env, torch = self.env, self.env.torch
            import device_oracle
            pad = 64
            stored = ((r, part.to(self.sdtype)) for r, part in self.blocks())
            wi, ws = device_oracle.oracle_topk(stored, self.queries_dev, self.k, pad=pad)
            if env.world > 1:
                gi = [torch.empty_like(wi) for _ in range(env.world)]
                gs = [torch.empty_like(ws) for _ in range(env.world)]
                env.dist.all_gather(gi, wi.contiguous())
                env.dist.all_gather(gs, ws.contiguous())
                wi, ws = device_oracle.merge_rank_lists(gi, gs, self.k + pad)
            self._oracle = (wi, ws)
            torch.cuda.empty_cache()
        return self._oracle

    def verify(self, res_s, res_i, Q):
        import device_oracle
        wi, ws = self.oracle()
        rep = device_oracle.parity_report(res_s[:Q], res_i[:Q], wi[:Q], ws[:Q], self.k)
        assert rep["ok"], f"bench result disagrees with the device oracle: {rep}"
        return (f"top-{self.k} of all {Q} quer{'y' if Q == 1 else 'ies'} match the device oracle (fp32 candidate pass + float64 "
                f"rescoring, plain torch) over all {self.N} rows: ids position by position (near-ties < 1e-6 swapped: "
                f"{rep['swapped_near_ties']}), scores within 1e-5 relative (worst {rep['max_rel_score_err']:.1e})")

    # ------------------------------------------------------------------ timing
    def measure(self, Q, steps, warmup, verify=True, count_launches=False):
        env, torch, dist, L = self.env, self.env.torch, self.env.dist, self.env.L
        from dab_b200 import ops
        store, sh, k, D, dev, world = self.store, self.sh, self.k, self.D, self.env.dev, self.env.world
        q_dev = self.queries_dev[:Q].contiguous()
        q_host = self.queries_host[:Q]
        q_np = q_host.numpy()
        out_host_s = torch.empty(Q, k, dtype=torch.float32).pin_memory()
        out_host_i = torch.empty(Q, k, dtype=torch.int64).pin_memory()
        gs = torch.empty(world, Q, k, dtype=torch.float32, device=dev) if world > 1 else None
        gi = torch.empty(world, Q, k, dtype=torch.int64, device=dev) if world > 1 else None
        def step_device():
            if sh.peer is not None:                         # one fused publish/wait/gather/merge kernel over NVLink
                slot = sh.peer.next_slot(Q, k)
                store.search_async(q_dev, k, out=slot)
                # every search has its own verdict slot: ask for it after the search was enqueued
                return sh.peer.merge(Q, k, L.dab_store_status_dev(store._h))
            s, i = store.search_async(q_dev, k)
            if world > 1:
                dist.all_gather_into_tensor(gs, s)
                dist.all_gather_into_tensor(gi, i)
                s, i = ops.merge_topk(gs, gi)
            return s, i

        def step_e2e():
            if world == 1:
                # the reference-facing call: host buffers in and out through one C-ABI entry point
                # (H2D of the queries, search, D2H of scores/ids, verdict check inside)
                return store.search_host(q_np, k)
            return sh.search_host(q_host, k, out=(out_host_s, out_host_i))   # pinned in, pinned out, checked

        def timed(fn, n):
            env.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                res = fn()
            e1.record()
            env.barrier()
            return env.max_over_ranks(e0.elapsed_time(e1)), res

        for _ in range(warmup):
            step_device()
        torch.cuda.synchronize()
        status = L.dab_store_search_status(store._h, torch.cuda.current_stream().cuda_stream)
        assert status == 0, f"warm-up search incomplete: {L.dab_last_error()}"
        launches0 = L.dab_launch_count()
        ms_total, (res_s, res_i) = timed(step_device, steps)
        launches = L.dab_launch_count() - launches0
        status = L.dab_store_search_status(store._h, torch.cuda.current_stream().cuda_stream)
        assert status == 0, f"timed search incomplete: {L.dab_last_error()}"
        for _ in range(warmup):          # the host-buffer path has its own first-call costs (pinned copies, lazy loads)
            step_e2e()
        ms_e2e, _ = timed(step_e2e, steps)

        # roofline pass: the scan kernels bracketed by events on their own stream
        store.set_profile(True)
        env.barrier()
        for _ in range(steps):
            step_device()
        prof = store.get_profile()
        store.set_profile(False)
        scan_ms = env.max_over_ranks(prof["scan_ms"]) / steps            # per search, slowest rank
        search_ms = env.max_over_ranks(prof["search_ms"]) / steps
        n_local = self.hi - self.lo
        alg_bytes = n_local * (D * self.elem + 4) + Q * D * 4 + Q * k * 12       # SURVEY 8(d), per GPU per query batch
        alg_flops = 2.0 * Q * n_local * D
        peaks = env.peaks
        out = dict(Q=Q, k=k, value=Q * steps / (ms_total * 1e-3), ms_per_step=ms_total / steps,
                   e2e=dict(value=Q * steps / (ms_e2e * 1e-3), unit="queries/s", h2d_bytes_per_step=Q * D * 4,
                            d2h_bytes_per_step=Q * k * 12, ms_per_step=ms_e2e / steps),
                   scan_ms=scan_ms, search_ms=search_ms, scan_launches=prof["scan_launches"] / max(1, steps),
                   alg_bytes=alg_bytes, alg_flops=alg_flops, launches=int(launches),
                   kernel_path=store.describe_plan(Q, k)["path"],
                   hbm_frac=alg_bytes / (scan_ms * 1e-3) / 1e9 / peaks["hbm"],
                   hbm_frac_step=alg_bytes / (ms_total / steps * 1e-3) / 1e9 / peaks["hbm"],
                   hbm_frac_e2e=alg_bytes / (ms_e2e / steps * 1e-3) / 1e9 / peaks["hbm"],
                   tensor_frac=alg_flops / (scan_ms * 1e-3) / 1e12 / peaks["tf"])
        out["verified"] = self.verify(res_s, res_i, Q) if verify else None
        out["_res"] = (res_s, res_i)          # the last timed search's answer (dropped before printing)
        return out

    def measure_filtered(self, steps, warmup):
        """A single-query search over a FILTERED queryset, the reference's everyday call
        (assistant/bot/services/context_service/steps/embeddings.py:26-30: bot + status filter, n = 5;
        assistant/processing/documents/steps/questions.py:124: document__id__lt).  Two int64 columns live in HBM
        beside the rows; every step builds the allow bitmap on the device (dab_store_build_allow) and searches
        with it - host query in, host result out.  Checked against the device oracle with the same row mask."""
        env, torch, L = self.env, self.env.torch, self.env.L
        from dab_b200 import _lib
        import device_oracle
        assert env.world == 1
        store, dev, n, k = self.store, env.dev, self.hi - self.lo, 5
        rows = torch.arange(n, device=dev, dtype=torch.int64)
        store.set_column(0, rows % 4)                      # "bot id"
        store.set_column(1, rows // 10)                    # "document id" (10 chunks per document)
        doc_limit = int(0.8 * n) // 10
        preds = [(0, _lib.DAB_OP_EQ, 1), (1, _lib.DAB_OP_LT, doc_limit)]
        q_np = self.queries_host[:1].numpy()

        def step():
            return store.search_host(q_np, k, store.build_allow(preds))

        def timed(fn, reps):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                r = fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps, r
        for _ in range(warmup):
            step()
        ms, (got_s, got_i) = timed(step, steps)
        ms_bitmap, _ = timed(lambda: store.build_allow(preds), steps)
        allow = lambda r0, m: ((torch.arange(r0, r0 + m, device=dev) % 4 == 1) & (torch.arange(r0, r0 + m, device=dev) // 10 < doc_limit))
        wi, ws = device_oracle.oracle_topk(((r, p.to(self.sdtype)) for r, p in self.blocks()), self.queries_dev[:1], k, pad=32, allow=allow)
        rep = device_oracle.parity_report(torch.from_numpy(got_s).to(dev), torch.from_numpy(got_i).to(dev), wi, ws, k)
        assert rep["ok"], f"filtered search disagrees with the device oracle: {rep}"
        bitmap_bytes = n * (2 * 8 + 4) + n // 8
        return dict(desc=f"{self.w['desc']} - ONE query, top-{k}, queryset filtered on two HBM columns (bot == 1 AND document_id < "
                         f"{doc_limit}: {0.25 * 0.8:.0%} of the rows pass), bitmap built on the device every step, host buffers in and out",
                    Q=1, k=k, rows=n, e2e=1e3 / ms, unit="queries/s", e2e_ms_per_step=ms, build_allow_ms=ms_bitmap,
                    build_allow_hbm_frac=bitmap_bytes / (ms_bitmap * 1e-3) / 1e9 / env.peaks["hbm"],
                    verified=f"top-{k} identical to the device oracle under the same row mask (worst relative score error {rep['max_rel_score_err']:.1e})")

    def measure_docsearch(self, steps, warmup):
        """The reference's document-level query (`embedding_search`, assistant/rag/services/search_service.py:111-152,
        called with max_scores_n = 5, top_n = 5 from steps/embeddings.py:47): the n = 250 nearest chunks of ONE query,
        grouped per document ON THE DEVICE (dab_doc_aggregate), <= 5 (document, score) pairs back to the host."""
        env, torch = self.env, self.env.torch
        from dab_b200 import ops
        import device_oracle
        assert env.world == 1
        store, dev, n = self.store, env.dev, self.hi - self.lo
        max_scores_n, top_n = 5, 5
        k = max_scores_n * top_n * 10                      # search_service.py:129-131
        # synthetic random rows have no document structure: 40 interleaved "documents" (row r belongs to document
        # r % 40) give the grouping something to do - ~6 of the 250 hits per document, several documents qualify
        n_docs = 40
        doc_of_row = torch.arange(n, device=dev, dtype=torch.int64) % n_docs
        q_host = self.queries_host[:1].clone().pin_memory()
        qd = q_host.to(dev)
        out_host = torch.empty(2, top_n, dtype=torch.float64).pin_memory()

        def step():
            q = q_host.to(dev, non_blocking=True)
            scores, rows = store.search(q, k)              # checked search (verdict read = the only sync besides the result)
            docs, dscores = ops.doc_aggregate(scores, rows, doc_of_row, max_scores_n, top_n)
            out_host[0].copy_(docs[0], non_blocking=True)
            out_host[1].copy_(dscores[0], non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return scores, rows, docs, dscores

        for _ in range(warmup):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            scores, rows, docs, dscores = step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        wi, ws = device_oracle.oracle_topk(((r, p.to(self.sdtype)) for r, p in self.blocks()), qd, k, pad=64)
        rep = device_oracle.parity_report(scores, rows, wi, ws, k)
        assert rep["ok"], f"n=250 search disagrees with the device oracle: {rep}"
        want = device_oracle.doc_level_scores((rows[0] % n_docs).tolist(), (1.0 - scores[0].double()).tolist(), max_scores_n, top_n)
        got = [(int(d), float(s)) for d, s in zip(docs[0].tolist(), dscores[0].tolist()) if d >= 0]
        assert [d for d, _ in got] == [d for d, _ in want] and all(abs(a[1] - b[1]) < 1e-5 for a, b in zip(got, want)), (got, want)
        return dict(desc=f"{self.w['desc']} - the reference's document-level query: ONE query, n = {k} nearest chunks, grouped per "
                         f"document on the device (max_scores_n = {max_scores_n}, top_n = {top_n}), host query in, <= {top_n} documents out",
                    Q=1, k=k, rows=n, e2e=1e3 / ms, unit="queries/s", e2e_ms_per_step=ms, documents_returned=len(got),
                    verified=f"top-{k} chunks identical to the device oracle; document ranking identical to the restated grouping "
                             f"of search_service.py:133-150 ({len(got)} documents)")

    def roofline(self, m):
        """The `roofline` object of the line from one measurement."""
        peaks, w = self.env.peaks, self.w
        t_hbm, t_tf = m["alg_bytes"] / (peaks["hbm"] * 1e9), m["alg_flops"] / (peaks["tf"] * 1e12)
        if t_tf > t_hbm and w["dtype"] == "bf16":
            roof = dict(bound="tensor", achieved=m["alg_flops"] / (m["scan_ms"] * 1e-3) / 1e12, peak=peaks["tf"], unit="TFLOP/s")
            roof["hbm_frac"] = m["hbm_frac"]
        else:
            roof = dict(bound="hbm", achieved=m["alg_bytes"] / (m["scan_ms"] * 1e-3) / 1e9, peak=peaks["hbm"], unit="GB/s")
        roof["frac"] = roof["achieved"] / roof["peak"]
        roof["kernel"] = "scan (all level launches of one search)"
        roof["scan_ms_per_search"] = m["scan_ms"]
        roof["search_ms_per_search"] = m["search_ms"]
        roof["scan_launches_per_search"] = m["scan_launches"]
        roof["peak_source"] = peaks["source"]
        if roof["bound"] == "hbm":
            roof["peak_note"] = "the HBM peak is the driver's COPY measurement (read + write); a read-only scan can run a few per cent above it"
        else:
            roof["peak_note"] = ("the tensor peak is cuBLAS bf16 under the sustained power cap; this kernel also streams the whole shard "
                                 "from HBM and runs power-capped (see clocks.reasons)")
        roof["traffic"] = None
        return roof

    def target(self, m):
        """Compact form of a measurement for the `targets` object."""
        return dict(Q=m["Q"], k=m["k"], rows=self.N, rows_per_gpu=self.hi - self.lo, dim=self.D, store_dtype=self.w["dtype"],
                    value=m["value"], unit="queries/s", ms_per_step=m["ms_per_step"], e2e=m["e2e"]["value"],
                    e2e_ms_per_step=m["e2e"]["ms_per_step"], scan_ms=m["scan_ms"], hbm_frac=m["hbm_frac"],
                    hbm_frac_step=m["hbm_frac_step"], hbm_frac_e2e=m["hbm_frac_e2e"], tensor_frac=m["tensor_frac"],
                    kernel_path=m["kernel_path"], exchange=self.exchange, verified=m["verified"])


class Env:
    pass


def vendor_gemm_yardstick(env, rows, dim, Q, reps=100):
    """What cuBLAS reaches on the scan's OWN shape - [rows x dim] bf16 streamed from HBM times [dim x Q] bf16, about
    half a second back to back so the power cap settles - against the 8192^3 L2-resident GEMM MEASURED_PEAKS.json
    quotes as the tensor roof.  It writes the [rows x Q] score matrix, which the scan never does; it selects nothing.
    A yardstick for roofline.frac of the Q=256 scan, not part of the product path (no library GEMM is)."""
    torch, dev = env.torch, env.dev
    a = torch.randn(rows, dim, device=dev, dtype=torch.bfloat16)
    b = torch.randn(Q, dim, device=dev, dtype=torch.bfloat16)
    out = torch.empty(rows, Q, device=dev, dtype=torch.bfloat16)
    for _ in range(5):
        torch.matmul(a, b.T, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        torch.matmul(a, b.T, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    tf = 2.0 * rows * dim * Q / (ms * 1e-3) / 1e12
    del a, b, out
    torch.cuda.empty_cache()
    return dict(desc=f"cuBLAS (torch.matmul) bf16 [{rows} x {dim}] x [{dim} x {Q}] -> bf16, {reps} calls back to back: the per-GPU "
                     "shard shape of the 8-GPU run, as a library GEMM", ms=ms, tflops=tf, frac_of_tensor_peak=tf / env.peaks["tf"],
                note="yardstick only: the product path calls no library GEMM")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("DAB_BENCH_WORKLOAD", DEFAULT_WORKLOAD), choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override the corpus row count (debugging)")
    ap.add_argument("--path", default="auto", choices=["auto", "simt", "tcgen05"])
    ap.add_argument("--exchange", default="auto", choices=["auto", "p2p", "nccl"], help="N>1: how local top-k lists are exchanged")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-targets", action="store_true", help="headline only: skip the `targets` sub-measurements")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    w = dict(WORKLOADS[args.workload])
    if args.rows:
        w["rows"] = args.rows
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, w, rank, world)
        return

    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from dab_b200 import _lib
    env = Env()
    env.torch, env.dist, env.dev, env.rank, env.world = torch, dist, dev, rank, world
    env.peaks = load_peaks()
    env.L = _lib.lib()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    env.barrier, env.max_over_ranks = barrier, max_over_ranks

    if w["kind"] == "pool":
        return bench_pool(args, w, env)

    verify = not args.no_verify
    targets = {}
    with_targets = args.workload == DEFAULT_WORKLOAD and not args.no_targets and not args.rows
    # ------------------------------------------------------------------ N=1: the single-GPU configs first
    # configs[1] and configs[2] do not fit beside the 154 GB store of configs[3]: measured and freed before it is built
    if with_targets:
        # configs[4] (ingest): pool + append into the owning shard on every rank, no collective
        m5 = measure_ingest(env, dict(WORKLOADS["c5"]), args.steps, args.warmup)
        targets["c5"] = dict(desc=WORKLOADS["c5"]["desc"], value=m5["value"], unit="chunks/s", ms_per_step=m5["ms_per_step"],
                             pool_ms=m5["pool_ms"], e2e=m5["e2e"]["value"], hbm_frac=m5["hbm_frac"], hbm_frac_step=m5["hbm_frac_step"],
                             verified=m5["verified"], config=m5["config"])
    if with_targets and world == 1:
        for name in ("c2", "c3"):
            sw = dict(WORKLOADS[name])
            run = SearchRun(env, sw, path=args.path)
            run.build()
            m = run.measure(sw["Q"], args.steps, args.warmup, verify=verify)
            m.pop("_res")
            targets[name] = dict(run.target(m), desc=sw["desc"])
            if name == "c3":
                targets["c3_filtered_q1"] = run.measure_filtered(args.steps, args.warmup)
                targets["c3_docsearch_q1"] = run.measure_docsearch(args.steps, args.warmup)
            run.close()
            del run

        targets["vendor_gemm_shard"] = vendor_gemm_yardstick(env, 12_500_000, w["dim"], w["Q"])

    # ------------------------------------------------------------------ the headline store
    run = SearchRun(env, w, exchange=args.exchange, path=args.path)
    run.build(keep_cpu_sample=(rank == 0 and world == 1 and not args.no_cpu_baseline))
    Q, k = w["Q"], w["k"]
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    head = run.measure(Q, args.steps, args.warmup, verify=False)
    clocks = sampler.stop() if rank == 0 else None
    roof = run.roofline(head)
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        roof["traffic"] = json.load(open(tpath)).get(args.workload)
    if with_targets:
        # north_star: "top-100 ... at >= 80 % of aggregate HBM roofline for Q <= 64" - the same store, the first 64
        # queries and the first query alone (the reference's own traffic is single-query: steps/embeddings.py:30,47)
        for name, q in (("q64", 64), ("q1", 1)):
            if q < Q:
                m = run.measure(q, args.steps, args.warmup, verify=False)
                m["verified"] = None
                targets[name] = m
    # correctness guard (not timed): one oracle pass over the corpus serves every query count measured on it
    verified = None
    if verify:
        # EVERY query of the last timed search of each query count is checked
        verified = run.verify(*head["_res"], Q)
        for name in ("q64", "q1"):
            if name in targets:
                targets[name]["verified"] = run.verify(*targets[name]["_res"], targets[name]["Q"])
    for name in ("q64", "q1"):
        if name in targets:
            targets[name] = dict(run.target(targets[name]), desc=f"{w['desc']} - first {targets[name]['Q']} quer"
                                 f"{'y' if targets[name]['Q'] == 1 else 'ies'} of the batch (north_star target shape)")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    tensor = head["kernel_path"] != "simt"
    arithmetic = ("bf16 x bf16 -> fp32 tcgen05 scan, then fp32 CUDA-core rescoring of the kept candidates"
                  if w["dtype"] == "bf16" and tensor else
                  "tf32 x tf32 -> fp32 tcgen05 scan, then fp32 CUDA-core rescoring of the kept candidates"
                  if tensor else "fp32 FMA dot products")
    line = dict(metric="exact cosine top-k queries/sec", value=head["value"], unit="queries/s", n_gpus=world, steps=args.steps,
                warmup=args.warmup, ms_per_step=head["ms_per_step"], higher_is_better=True, scaling="strong",
                vs_baseline=None, dtype=("bf16" if w["dtype"] == "bf16" else "tf32") if tensor else "f32", data="synthetic",
                config=search_config(args, w, w["rows"] // world),
                engine=dict(path=args.path, kernel_path=head["kernel_path"], exchange=run.exchange, arithmetic=arithmetic),
                e2e=head["e2e"], gpu_launches=head["launches"], clocks=clocks, roofline=roof, verified=verified)
    if targets:
        line["targets"] = targets
    if world == 1 and not args.no_cpu_baseline and run.sample_host is not None:
        from oracle import c_oracle
        c_oracle.build()
        threads = os.cpu_count() or 1
        v, desc = cpu_search_sample(w, run.sample_host, run.queries_host.numpy(), 12.0, threads)
        line["cpu_baseline"] = dict(value=v, unit="queries/s", cores=threads, kind="port", sample=desc)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def measure_ingest(env, w, steps, warmup, cpu_baseline=False):
    """configs[4] as a PATH, not a kernel: every rank pools + normalises its own batch of encoder hidden states
    and appends the rows to its own shard of a row-sharded store (chunk -> GPU = row -> shard: no collective).
    One step = B chunks per GPU; `value` = chunks/s over all ranks (weak scaling)."""
    torch, dev, rank, world, L = env.torch, env.dev, env.rank, env.world, env.L
    from dab_b200 import ops, sharded
    B, T, D = w["B"], w["T"], w["dim"]
    total = w.get("rows", 1_000_000)
    sh = sharded.ShardedStore(total, D, torch.bfloat16, dev, exchange="nccl")
    g = torch.Generator(device=dev).manual_seed(777 + rank)
    nb = 4                                   # resident batches rotated so that no step re-reads L2-resident data
    hidden = [torch.randn(B, T, D, generator=g, device=dev) for _ in range(nb)]
    lengths = torch.randint(16, T + 1, (B,), generator=torch.Generator().manual_seed(555))
    mask = (torch.arange(T)[None] < lengths[:, None]).to(torch.int64)
    mask_dev = mask.to(dev)
    hidden_host = torch.randn(B, T, D).pin_memory()
    mask_host = mask.pin_memory()
    ack_host = torch.empty(B, dtype=torch.float32).pin_memory()
    it = [0]
    cap = sh.hi - sh.lo

    def room():
        if len(sh.local) + B > cap:          # the shard is full: start over (the timed work per step is unchanged)
            sh.local.reset()

    def step_device():
        it[0] += 1
        room()
        return sh.ingest(hidden[it[0] % nb], mask_dev)

    def step_e2e():
        # host buffers in (the encoder's output arriving from elsewhere), the rows' inverse norms back as the receipt
        h = hidden_host.to(dev, non_blocking=True)
        m = mask_host.to(dev, non_blocking=True)
        room()
        n0 = len(sh.local)
        sh.ingest(h, m)
        _, inv, _ = sh.local.export(n0, B)
        ack_host.copy_(inv, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def timed(fn, n):
        env.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        env.barrier()
        return env.max_over_ranks(e0.elapsed_time(e1))

    for _ in range(warmup):
        step_device()
    l0 = L.dab_launch_count()
    ms = timed(step_device, steps)
    launches = L.dab_launch_count() - l0
    # the pool kernel alone, for the roofline (the append adds a pass over B x D bf16 = 0.2 % of the bytes)
    def pool_only():
        it[0] += 1
        return ops.pool_l2norm(hidden[it[0] % nb], mask_dev, want_bf16=True)
    for _ in range(3):
        pool_only()
    ms_pool = timed(pool_only, steps)
    for _ in range(3):
        step_e2e()
    e2e_steps = max(3, steps // 4)
    ms_e2e = timed(step_e2e, e2e_steps)
    # parity of the path: what landed in the shard is the pooled batch (checked against a torch fp32 statement)
    sh.local.reset()
    sh.ingest(hidden[0], mask_dev)
    rows, inv, ids = sh.local.export(0, B)
    m = mask_dev.float()
    want = torch.nn.functional.normalize((hidden[0] * m[:, :, None]).sum(1) / m.sum(1, keepdim=True).clamp(min=1), dim=1)
    err = float((rows.float() - want).abs().max())
    assert err < 2 ** -8, f"ingested rows differ from the pooled batch by {err}"      # bf16 rounding of unit-norm rows
    assert int(ids[0]) == sh.lo and int(ids[-1]) == sh.lo + B - 1
    # the kernel skips masked-out token rows, so the bytes it must move are the unmasked ones
    tokens = int(mask.sum())
    alg_bytes = tokens * D * 4 + B * T * 8 + B * D * (4 + 2)
    nominal_bytes = B * (T * D * 4 + T * 8 + D * 4)
    ach = alg_bytes / (ms_pool / steps * 1e-3) / 1e9
    out = dict(value=world * B * steps / (ms * 1e-3), unit="chunks/s", ms_per_step=ms / steps, pool_ms=ms_pool / steps,
               e2e=dict(value=world * B * e2e_steps / (ms_e2e * 1e-3), unit="chunks/s",
                        h2d_bytes_per_step=B * T * D * 4 + B * T * 8, d2h_bytes_per_step=B * 4, ms_per_step=ms_e2e / e2e_steps),
               launches=int(launches), hbm_frac=ach / env.peaks["hbm"],
               hbm_frac_step=alg_bytes / (ms / steps * 1e-3) / 1e9 / env.peaks["hbm"],
               roofline=dict(bound="hbm", achieved=ach, peak=env.peaks["hbm"], unit="GB/s", frac=ach / env.peaks["hbm"],
                             traffic=None, kernel="pool_l2norm_kernel", peak_source=env.peaks["source"],
                             note=f"algorithmic bytes count only unmasked token rows ({tokens} of {B * T}); with every row "
                                  f"counted: {nominal_bytes / (ms_pool / steps * 1e-3) / 1e9:.0f} GB/s"),
               verified=f"rows landed in shard rows [{sh.lo}, {sh.lo + B}) equal the pooled batch (max abs diff {err:.1e} after bf16 rounding)",
               config={"B": B, "T": T, "dim": D, "store": f"{total} x {D} bf16 row-sharded over {world} GPU(s)",
                       "l2": f"{nb} resident batches of {B * T * D * 4 / 1e6:.0f} MB rotated (> 126 MB L2)"})
    if cpu_baseline and rank == 0:
        from oracle import pool_ref
        hs, msk = hidden_host[:32], mask[:32]
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 10.0:
            pool_ref.masked_mean_pool_torch_cpu(hs, msk)
            reps += 1
        dt = (time.perf_counter() - t0) / reps
        out["cpu_baseline"] = dict(value=32 / dt, unit="chunks/s", cores=torch.get_num_threads(), kind="port",
                                   sample=f"{reps} reps of 32 chunks, CPU torch masked mean + F.normalize")
    sh.close()
    del hidden
    torch.cuda.empty_cache()
    return out


def bench_pool(args, w, env):
    import torch.distributed as dist
    sampler = ClockSampler(env.dev.index)
    if env.rank == 0:
        sampler.start()
    m = measure_ingest(env, w, args.steps, args.warmup, cpu_baseline=(env.world == 1 and not args.no_cpu_baseline))
    clocks = sampler.stop() if env.rank == 0 else None
    if env.rank != 0:
        if env.world > 1:
            dist.destroy_process_group()
        return
    line = dict(metric="ingest pool+L2norm chunks/sec", value=m["value"], unit="chunks/s", n_gpus=env.world, steps=args.steps,
                warmup=args.warmup, ms_per_step=m["ms_per_step"], higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic", config=dict(m["config"], workload=args.workload, desc=w["desc"]),
                e2e=m["e2e"], gpu_launches=m["launches"], clocks=clocks, roofline=m["roofline"], verified=m["verified"],
                pool_ms=m["pool_ms"])
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        line["roofline"]["traffic"] = json.load(open(tpath)).get(args.workload)
    if "cpu_baseline" in m:
        line["cpu_baseline"] = m["cpu_baseline"]
    print(json.dumps(line))
    if env.world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
