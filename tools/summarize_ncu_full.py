#!/usr/bin/env python3
"""Markdown summary of every launch in an `ncu --set full` report: time, DRAM bytes (traffic), DRAM throughput, the busiest
pipes, occupancy, top stall reasons.   python tools/summarize_ncu_full.py <rep> [algorithmic-bytes-json]
The optional JSON list gives the algorithmic bytes of every launch, in launch order (traffic / algorithmic is printed)."""
import csv
import json
import subprocess
import sys

rep = sys.argv[1]
algo = json.loads(sys.argv[2]) if len(sys.argv) > 2 else []   # algorithmic bytes per launch, in launch order
if rep.endswith(".csv"):   # a raw page exported on the GPU box (`ncu -i rep --page raw --csv`): reports are too big to bring back
    out = open(rep).read()
else:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}


def val(r, name, default=float("nan")):
    try:
        return float(r[col[name]].replace(",", ""))
    except (KeyError, ValueError):
        return default


def unit(name):
    return units[col[name]] if name in col else ""


def to_bytes(r, name):
    v, u = val(r, name), unit(name)
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Tbyte": 1e12}.get(u, 1.0)


def to_ms(r, name):
    v, u = val(r, name), unit(name)
    return v * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3, "msecond": 1.0, "usecond": 1e-3, "nsecond": 1e-6, "second": 1e3}.get(u, 1.0)


print(f"# ncu --set full summary of `{rep}` (times under ncu are cold-cache and serialised)\n")
print("| # | kernel | grid x block | ms | DRAM read + write = traffic (GB) | traffic / algorithmic | DRAM % of peak | fp64 pipe % | LSU pipe % | issue active % | warps active % | regs | top stalls (per issue) |")
print("|---|---|---|---:|---|---:|---:|---:|---:|---:|---:|---:|---|")
stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
for i, r in enumerate(data):
    name = r[col["Kernel Name"]]
    short = name.split("(")[0].replace("void qb::", "").replace("void ", "")
    rd, wr = to_bytes(r, "dram__bytes_read.sum"), to_bytes(r, "dram__bytes_write.sum")
    a = algo[i] if i < len(algo) else None
    top = sorted(((val(r, h, 0.0), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for h in stalls), reverse=True)[:3]
    print(f"| {i} | `{short}` | {r[col['Grid Size']]} x {r[col['Block Size']]} | {to_ms(r, 'gpu__time_duration.sum'):.3f} | "
          f"{rd / 1e9:.2f} + {wr / 1e9:.2f} = {(rd + wr) / 1e9:.2f} | {((rd + wr) / a):.3f} |" if a else
          f"| {i} | `{short}` | {r[col['Grid Size']]} x {r[col['Block Size']]} | {to_ms(r, 'gpu__time_duration.sum'):.3f} | "
          f"{rd / 1e9:.2f} + {wr / 1e9:.2f} = {(rd + wr) / 1e9:.2f} | - |", end="")
    print(f" {val(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | "
          f"{val(r, 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{val(r, 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{val(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{val(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} | {val(r, 'launch__registers_per_thread'):.0f} | "
          + ", ".join(f"{n} {v:.2f}" for v, n in top) + " |")
