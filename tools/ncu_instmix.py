#!/usr/bin/env python
"""Executed-instruction mix + hottest source lines of one kernel in an ncu report (needs --import-source on):
    python tools/ncu_instmix.py rep.ncu-rep <kernel-name-regex> [launch-index]"""
import csv
import io
import re
import subprocess
import sys
from collections import Counter

rep, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", "regex:" + pat],
                     capture_output=True, text=True).stdout
blocks = out.split('"Kernel Name"')
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
blk = '"Kernel Name"' + blocks[1 + which]
rows = list(csv.reader(io.StringIO(blk)))
print(rows[0][1][:100])
hdr = rows[1]
ia, isrc, ismp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
ops, smp, tot = Counter(), Counter(), 0
lines = []
for r in rows[2:]:
    if len(r) <= ia or not r[ia].isdigit():
        continue
    n = int(r[ia])
    tot += n
    toks = r[isrc].split()
    op = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
    ops[op] += n
    smp[op] += int(r[ismp])
    lines.append((int(r[ismp]), n, r[isrc].strip()))
print("total warp instructions", tot)
for k, v in ops.most_common(22):
    print(f"{k:12s} {v:12d} {v / tot * 100:5.1f}%  samples {smp[k]}")
print("-- hottest SASS lines by samples")
for s, n, src in sorted(lines, reverse=True)[:25]:
    print(f"{s:6d} {n:10d}  {src[:110]}")
