#!/usr/bin/env python3
"""Static instruction mix of one kernel in a built library: python tools/sass_mix.py <lib.so> <mangled-name-substring>."""
import collections
import re
import subprocess
import sys

lib, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
on, mix, n = False, collections.Counter(), 0
for line in out.splitlines():
    if "Function :" in line:
        on = pat in line
        continue
    if not on:
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
    if m:
        op = m.group(1) + (".MOV" if ".MOV" in m.group(2) else "")
        mix[op] += 1
        n += 1
print(f"{n} instructions = {n * 16 / 1024:.1f} KiB")
print(", ".join(f"{k} {v}" for k, v in mix.most_common(22)))
