#!/usr/bin/env python
"""Per-kernel census of the Blackwell-native SASS mnemonics in libvfi_b200.so (cuobjdump -sass):
    python tools/sass_census.py > profiles/r02_sass_census.txt
UTCHMMA = tcgen05.mma (bf16), LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA tensor load / store, UTCBAR = tcgen05.commit,
SYNCS = mbarrier ops, HMMA = legacy mma.sync, IDP = dp4a, DADD / DFMA / DMUL = fp64 pipe."""
import glob
import os
import re
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = glob.glob(os.path.join(ROOT, "video-frame*", "libvfi_b200.so"))[0]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
WANT = ["UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "HMMA", "IDP", "DFMA", "SHFL", "LDGSTS"]
kern, cur = OrderedDict(), None
for ln in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = kern.setdefault(re.sub(r"\((?!anonymous).*", "", name)[:78], dict.fromkeys(WANT, 0) | {"instr": 0})
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
    if m and cur is not None:
        cur["instr"] += 1
        op = m.group(1)
        for w in WANT:
            if op.startswith(w):
                cur[w] += 1
print(f"# {os.path.basename(lib)}: SASS mnemonic counts per kernel (sm_100a); kernels without any listed mnemonic are summed at the end")
print(f"{'kernel':80s}" + "".join(f"{w:>9s}" for w in WANT) + f"{'instr':>9s}")
tot = dict.fromkeys(WANT, 0) | {"instr": 0}
rest = 0
for k, v in kern.items():
    for w in tot:
        tot[w] += v[w]
    if any(v[w] for w in WANT[:8]):
        print(f"{k:80s}" + "".join(f"{v[w]:9d}" for w in WANT) + f"{v['instr']:9d}")
    else:
        rest += 1
print(f"{'TOTAL (' + str(len(kern)) + ' kernels, ' + str(rest) + ' not listed)':80s}" + "".join(f"{tot[w]:9d}" for w in WANT) + f"{tot['instr']:9d}")
