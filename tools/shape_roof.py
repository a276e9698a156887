"""What the vendor GEMM reaches on the scan's own shape: [rows x dim] bf16 corpus times [dim x Q] bf16 queries,
streamed from HBM, sustained for about a second, with the SM clock sampled meanwhile.  A yardstick for the
wide-query tensor kernel (which additionally thresholds every score and never writes the [rows x Q] matrix).
usage: python tools/shape_roof.py [--rows 12500000] [--dim 768] [--q 256]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import ClockSampler  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=12_500_000)
    ap.add_argument("--dim", type=int, default=768)
    ap.add_argument("--q", type=int, default=256)
    ap.add_argument("--reps", type=int, default=200)
    a = ap.parse_args()
    dev = "cuda:0"
    rows = torch.randn(a.rows, a.dim, device=dev, dtype=torch.bfloat16)
    q = torch.randn(a.q, a.dim, device=dev, dtype=torch.bfloat16)
    out = torch.empty(a.rows, a.q, device=dev, dtype=torch.bfloat16)
    for _ in range(5):
        torch.matmul(rows, q.T, out=out)
    torch.cuda.synchronize()
    cs = ClockSampler(0)
    cs.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        torch.matmul(rows, q.T, out=out)
    e1.record()
    torch.cuda.synchronize()
    clocks = cs.stop()
    ms = e0.elapsed_time(e1) / a.reps
    flops = 2.0 * a.rows * a.dim * a.q
    print(f"cuBLAS bf16 [{a.rows}x{a.dim}] x [{a.dim}x{a.q}] -> bf16: {ms:.3f} ms = {flops / ms / 1e9:.0f} TFLOP/s, "
          f"reads {a.rows * a.dim * 2 / ms / 1e6:.0f} GB/s + writes {a.rows * a.q * 2 / ms / 1e6:.0f} GB/s; clocks {clocks}")


if __name__ == "__main__":
    main()
