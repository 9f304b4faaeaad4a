#!/usr/bin/env python
"""Per-kernel sums of an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...).
    python tools/launch_summary.py gpurun_out/launches.csv ["header text"] > profiles/launches_summary.txt"""
import csv
import re
import sys
from collections import OrderedDict

rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
rd = csv.DictReader(rows)
acc = OrderedDict()
total, n = 0.0, 0
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*$", "", r["Kernel Name"]).replace("vv::", "").replace("void ", "")
    us = float(r["Metric Value"].replace(",", "")) / (1e3 if r["Metric Unit"] == "ns" else 1.0)
    a = acc.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += us
    total += us
    n += 1
if len(sys.argv) > 2:
    print(sys.argv[2])
print(f"total {total:.1f} us over {n} launches")
for name, (k, us) in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:46s} n={k:3d} {us:10.1f} us {100 * us / total:5.1f}%")
