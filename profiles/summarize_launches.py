"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time share per kernel.
usage: python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches.md"""
import collections
import csv
import sys


def main(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(row["Metric Unit"], 1.0)
        k = agg[row["Kernel Name"].split("(")[0][:90]]
        k[0] += 1
        k[1] += v
    tot = sum(v[1] for v in agg.values())
    ours = sum(v[1] for n, v in agg.items() if "dab::" in n)
    print(f"| kernel | launches | total us | avg us | share of all | share of dab:: kernels |")
    print("|---|---:|---:|---:|---:|---:|")
    for n, v in sorted(agg.items(), key=lambda x: -x[1][1])[:20]:
        own = f"{100 * v[1] / ours:.1f}%" if "dab::" in n else ""
        print(f"| `{n}` | {v[0]} | {v[1] / 1e3:.1f} | {v[1] / 1e3 / v[0]:.2f} | {100 * v[1] / tot:.1f}% | {own} |")


if __name__ == "__main__":
    main(sys.argv[1])
