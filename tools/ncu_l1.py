#!/usr/bin/env python
"""L1 traffic per memory opcode of one kernel in an ncu report (--import-source on): executed instructions, global
tag requests, shared wavefronts and L2 sectors, optionally divided by `norm`:
    python tools/ncu_l1.py rep.ncu-rep <kernel-regex> [norm]"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

rep, pat = sys.argv[1], sys.argv[2]
N = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", "regex:" + pat],
                     capture_output=True, text=True).stdout
blk = '"Kernel Name"' + out.split('"Kernel Name"')[1]
rows = list(csv.reader(io.StringIO(blk)))
hdr = rows[1]
cols = ["Instructions Executed", "L1 Tag Requests Global", "L1 Wavefronts Shared", "L2 Theoretical Sectors Global"]
idx = [hdr.index(c) for c in cols]
isrc = hdr.index("Source")
agg = defaultdict(lambda: [0, 0, 0, 0])
for r in rows[2:]:
    if len(r) <= max(idx) or not r[idx[0]].isdigit():
        continue
    t = r[isrc].split()
    op = t[1] if t[0].startswith("@") else t[0]
    if re.match(r"LDG|STG|LDS|STS|ATOM|RED|SHFL|LDSM|STSM", op):
        for j, i in enumerate(idx):
            agg[op][j] += int(r[i]) if r[i].isdigit() else 0
print(f"{'opcode':28s} {'instr':>10s} {'tag req':>10s} {'smem wavefr':>12s} {'L2 sectors':>11s}")
tot = [0, 0, 0, 0]
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1] - kv[1][2]):
    print(f"{k:28s} " + " ".join(f"{x / N:11.1f}" for x in v))
    tot = [a + b for a, b in zip(tot, v)]
print(f"{'total':28s} " + " ".join(f"{x / N:11.1f}" for x in tot))
