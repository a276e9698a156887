import sys, time, os
sys.path.insert(0, "django-assistant-bot_b200"); sys.path.insert(0, ".")
import torch
from dab_b200 import ops, _lib
dev = torch.device("cuda:0")
N, D, Q, k = 2_000_000, 1024, 64, 100
store = ops.ChunkStore(N, D, torch.bfloat16, dev)
for b in range(2):
    store.append(torch.nn.functional.normalize(torch.randn(1_000_000, D, device=dev), dim=1))
qh = torch.nn.functional.normalize(torch.randn(Q, D), dim=1).pin_memory()
L = _lib.lib()
def t(fn, n=20):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
qd = qh.to(dev)
print("async", t(lambda: store.search_async(qd, k)))
def f_status():
    store.search_async(qd, k)
    L.dab_store_search_status(store._h, torch.cuda.current_stream().cuda_stream)
print("async+status", t(f_status))
def f_isnan():
    s, i = store.search_async(qd, k)
    L.dab_store_search_status(store._h, torch.cuda.current_stream().cuda_stream)
    bool(torch.isnan(s).any())
print("async+status+isnan", t(f_isnan))
print("checked", t(lambda: store.search(qd, k)))
def f_e2e():
    q = qh.to(dev, non_blocking=True)
    s, i = store.search(q, k)
    s.cpu(); i.cpu()
print("e2e", t(f_e2e))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(20): f_e2e()
pr.disable(); pstats.Stats(pr).sort_stats("cumtime").print_stats(12)
