#!/usr/bin/env python
"""Executed warp instructions and stall samples per CUDA source line of one kernel, from an ncu report
(--import-source on) and the line table of the cubin it ran (nvdisasm -g):
    python tools/ncu_lines.py rep.ncu-rep <kernel-regex> <file.cubin> <mangled-name-substring> [norm]
`norm` divides the counts (e.g. the trip count of the hot loop) so the figures read "per pixel"."""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

rep, pat, cubin, sub = sys.argv[1:5]
norm = float(sys.argv[5]) if len(sys.argv) > 5 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", "regex:" + pat],
                     capture_output=True, text=True).stdout
blk = '"Kernel Name"' + out.split('"Kernel Name"')[1]
rows = list(csv.reader(io.StringIO(blk)))
hdr = rows[1]
ia, ismp, iad = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Address")
sass = [(int(r[iad], 16), int(r[ia]), int(r[ismp])) for r in rows[2:] if len(r) > ia and r[ia].isdigit()]
base = sass[0][0]

dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
line_of, cur, infn, inl = {}, None, False, ""
for ln in dis.splitlines():
    if ln.startswith("//--------------------- .text."):
        infn = sub in ln
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m and cur:
        line_of[int(m.group(1), 16)] = cur

agg = defaultdict(lambda: [0, 0])
for a, n, s in sass:
    k = line_of.get(a - base, ("?", 0))
    agg[k][0] += n
    agg[k][1] += s
tot = sum(v[0] for v in agg.values())
print("total warp instructions", tot, " normalised", round(tot / norm, 1))
src = {}
for (f, l), (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:60]:
    if f not in src:
        try:
            src[f] = open(subprocess.run(["bash", "-c", f"ls /root/repo/video-frame*/csrc/{f}"], capture_output=True,
                                         text=True).stdout.strip()).read().splitlines()
        except Exception:
            src[f] = []
    text = src[f][l - 1].strip()[:100] if 0 < l <= len(src[f]) else ""
    print(f"{n / norm:9.1f} {s:7d}  {f}:{l:<5d} {text}")
