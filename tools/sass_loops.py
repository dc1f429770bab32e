#!/usr/bin/env python
"""Static SASS size of a kernel's loops (no GPU needed):  python tools/sass_loops.py <lib.so> <kernel-name-substring>
Prints every backward branch with its body length and the opcode histogram of the largest loop."""
import re
import subprocess
import sys
from collections import Counter

lib, sub = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur, fn = None, {}
for ln in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = m.group(1)
        fn[cur] = []
    elif cur and re.match(r"\s*/\*[0-9a-f]{4}\*/", ln):
        fn[cur].append(ln)
for name, lines in fn.items():
    if sub not in name:
        continue
    addr = lambda l: int(re.match(r"\s*/\*([0-9a-f]{4})\*/", l).group(1), 16)
    print(name[:90], "instructions:", len(lines))
    best = None
    for l in lines:
        if "BRA" in l:
            t = re.search(r"0x([0-9a-f]+)", l.split("BRA")[1])
            if t and int(t.group(1), 16) < addr(l):
                lo, hi = int(t.group(1), 16), addr(l)
                n = (hi - lo) // 16 + 1
                print(f"  loop {lo:#x}..{hi:#x}: {n} instructions")
                if not best or n > best[2]:
                    best = (lo, hi, n)
    if best:
        c = Counter()
        for l in lines:
            if best[0] <= addr(l) <= best[1]:
                t = l.split("*/", 1)[1].strip().split()
                c[(t[1] if t[0].startswith("@") else t[0]).split(".")[0]] += 1
        print("  largest loop:", c.most_common())
