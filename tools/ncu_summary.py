"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) into per-kernel totals / shares.
usage: python tools/ncu_summary.py gpurun_out/launches.csv [frames] > profiles/rNN_launches.md"""
import csv
import re
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            unit = r.get("Metric Unit", "ns")
            v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
            name = re.sub(r"\(.*", "", r["Kernel Name"])
            name = name.split("::")[-1]
            rows.append((name, v, r.get("Grid Size", ""), r.get("Block Size", "")))
    tot = defaultdict(lambda: [0, 0.0])
    for n, v, *_ in rows:
        tot[n][0] += 1
        tot[n][1] += v
    total = sum(v for _, v in tot.values())
    print(f"# ncu launch list summary: {len(rows)} launches, {total:.1f} us total ({frames} frames)\n")
    print("| kernel | launches | total us | us/launch | share |")
    print("|---|---|---|---|---|")
    for n, (c, v) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print(f"| {n} | {c} | {v:.1f} | {v / c:.2f} | {100 * v / total:.1f}% |")


if __name__ == "__main__":
    main()
