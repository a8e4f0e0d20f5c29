#!/usr/bin/env python
"""Top stall rows per kernel from `ncu -i X.ncu-rep --page source --csv` (stdin): sections split at 'Kernel Name' rows."""
import collections
import csv
import sys

top_n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
rows = list(csv.reader(sys.stdin))
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}
        sections.append(cur)
    elif cur is not None and cur["hdr"] is None and "Source" in r:
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] is not None and len(r) == len(cur["hdr"]):
        cur["data"].append(r)
for s in sections:
    hdr, data = s["hdr"], s["data"]
    isamp, iex, isrc, ith = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("Avg. Threads Executed")
    cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    agg = collections.Counter()
    for r in data:
        for i in cols:
            if r[i].isdigit():
                agg[hdr[i]] += int(r[i])
    print("===", s["name"][:40], "rows", len(data), "samples", sum(int(r[isamp]) for r in data), agg.most_common(5))
    for i in sorted(sorted(range(len(data)), key=lambda i: -int(data[i][isamp]))[:top_n]):
        r = data[i]
        print("  ", i, "samp", r[isamp], "exec", r[iex], "thr", r[ith], r[isrc][:90])
