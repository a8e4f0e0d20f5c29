#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel: launches, total, share."""
import csv
import re
import sys
from collections import defaultdict

path, last_n = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = []
with open(path, newline="") as f:
    lines = [l for l in f if not l.startswith("==")]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
    rows.append((re.sub(r"\(.*", "", r["Kernel Name"]), v * scale))
if last_n:
    rows = rows[-last_n:]
tot = sum(v for _, v in rows)
agg = defaultdict(lambda: [0, 0.0])
for k, v in rows:
    agg[k][0] += 1
    agg[k][1] += v
print("launches %d, total %.1f us" % (len(rows), tot))
print("%-28s %8s %12s %8s" % ("kernel", "launches", "total_us", "share"))
for k, (n, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-28s %8d %12.1f %7.1f%%" % (k[:28], n, v, 100 * v / tot))
