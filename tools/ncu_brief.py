#!/usr/bin/env python
"""Brief per-kernel summary of an .ncu-rep (duration, occupancy, IPC, active lanes, DRAM bytes, top stall reasons)."""
import csv
import subprocess
import sys

for rep in sys.argv[1:]:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    want = ["Kernel Name", "gpu__time_duration.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_elapsed",
            "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread", "launch__grid_size", "dram__bytes_read.sum",
            "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed"]
    stall = [h for h in hdr if "smsp__average_warps_issue_stalled" in h and "per_issue_active" in h and "not_issued" not in h]
    for r in rows[2:]:
        print("=====", rep)
        for w in want:
            if w in hdr:
                print("  ", w, r[hdr.index(w)], units[hdr.index(w)])
        st = sorted(((float(r[hdr.index(h)].replace(",", "")), h) for h in stall), reverse=True)[:8]
        for v, h in st:
            print("      stall %.2f %s" % (v, h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
