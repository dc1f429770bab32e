#!/usr/bin/env python
"""Summarise an `ncu --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,
sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active` log: one line per launch + per-kernel totals."""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
d = OrderedDict()
for r in rows:
    d.setdefault(int(r[0]), {"name": r[4].split("(")[0].replace("void ", "")[:26], "grid": r[8]})[r[12]] = float(r[14].replace(",", ""))
first = min(d)
tot = {}
for i, v in d.items():
    t = v.get("gpu__time_duration.sum", 0) / 1000
    rd, wr = v.get("dram__bytes_read.sum", 0) / 1e6, v.get("dram__bytes_write.sum", 0) / 1e6
    tp = v.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0)
    print(f"{i - first:3d} {v['name']:26s} {v['grid']:16s} {t:9.1f} us  rd {rd:8.1f} MB  wr {wr:8.1f} MB  tensor {tp:5.1f}%")
    a = tot.setdefault(v["name"], [0, 0.0])
    a[0] += 1
    a[1] += t
print("# totals")
for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"# {k:26s} {n:3d} launches {t / 1000:8.3f} ms")
print(f"# all {sum(t for _, t in tot.values()) / 1000:.3f} ms")
