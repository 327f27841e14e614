#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share.
Usage: python tools/summarize_ncu.py gpurun_out/launches.csv [skip_first_n_launches] > profiles/launches_rNN.md"""
import csv
import re
import sys
from collections import OrderedDict

path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = []
with open(path) as f:
    lines = [ln for ln in f if ln.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        rows.append((r["Kernel Name"], float(r["Metric Value"]), r["Grid Size"], r["Block Size"]))
rows = rows[skip:]
agg = OrderedDict()
for name, ns, grid, block in rows:
    short = name.split(">(")[0] + ">" if ">(" in name else name.split("(")[0]
    a = agg.setdefault(short, [0, 0.0, grid, block])
    a[0] += 1
    a[1] += ns
total = sum(a[1] for a in agg.values())
print(f"# ncu launch list summary: {path} (first {skip} launches skipped), {len(rows)} launches, {total / 1e6:.2f} ms of kernel time")
print("# per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes\n")
print("| kernel | launches | total ms | mean ms | share |")
print("|---|---:|---:|---:|---:|")
for k, (n, ns, grid, block) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{k}` | {n} | {ns / 1e6:.3f} | {ns / n / 1e6:.4f} | {100 * ns / total:.1f}% |")
