#!/usr/bin/env python3
"""Per-opcode executed-instruction counts and the hot code segments of the first kernel in an .ncu-rep that holds
SourceCounters: python tools/ncu_hot.py <rep | source-page.csv[:k]> [amplitudes-per-launch-log2=30] [ctas=524288]
(a .csv is `ncu -i rep --page source --csv` exported on the GPU box; `:k` selects the k-th launch in it, default 0)"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
amps = 2 ** int(sys.argv[2]) if len(sys.argv) > 2 else 2 ** 30
ctas = int(sys.argv[3]) if len(sys.argv) > 3 else 524288
which = 0
if ".csv" in rep:
    rep, _, k = rep.partition(".csv")
    rep, which = rep + ".csv", int(k[1:] or 0)
    out = open(rep).read()
else:
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
heads = [k for k, r in enumerate(rows) if "Address" in r]
hdr = rows[heads[which]]
data = [r for r in rows[heads[which] + 1:(heads[which + 1] if which + 1 < len(heads) else len(rows))] if len(r) == len(hdr)]
iS, iE, iN = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
tot = sum(int(r[iE]) for r in data)
samples = sum(int(r[iN]) for r in data)
print(f"{tot} warp instructions = {tot * 32 / amps:.1f} thread instructions per amplitude")
c = collections.Counter()
for r in data:
    t = r[iS].strip().split()
    op = t[1] if t[0].startswith("@") else t[0]
    c[op.split(".")[0] + (".MOV" if ".MOV" in op else "")] += int(r[iE])
print(", ".join(f"{op} {n * 32 / amps:.1f}" for op, n in c.most_common(26)))
stall = {h: sum(int(r[i] or 0) for r in data) for i, h in enumerate(hdr) if h.startswith("stall_") and h.endswith("(Not Issued)")}
allst = sum(stall.values())
print("stall samples (warp not issued): " + ", ".join(f"{h.split()[0]} {n / allst * 100:.1f}%" for h, n in sorted(stall.items(), key=lambda kv: -kv[1])[:8]))
segs, start, prev = [], 0, None
for k, r in enumerate(data):
    e = int(r[iE])
    if prev is not None and e != prev:
        segs.append((start, prev, k - start))
        start = k
    prev = e
segs.append((start, prev, len(data) - start))
for st, e, n in segs:
    share = e * n / tot * 100
    smp = sum(int(data[i][iN]) for i in range(st, st + n)) / max(samples, 1) * 100
    if share > 1.0 or smp > 1.5:
        print(f"@{st:5d} len {n:4d} exec/CTA {e / ctas:7.1f} instr-share {share:5.1f}% stall-sample-share {smp:5.1f}%  {data[st][iS].strip()[:56]}")
