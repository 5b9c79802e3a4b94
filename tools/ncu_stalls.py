#!/usr/bin/env python3
"""Top warp-stall reasons (warps stalled per issue slot) of every kernel in an ncu --set full report.
usage: tools/ncu_stalls.py report.ncu-rep"""
import csv
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
ki = hdr.index("Kernel Name")
cols = [i for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
seen = set()
for r in rows[2:]:
    k = r[ki].split("(")[0]
    if k in seen:
        continue
    seen.add(k)
    vals = sorted(((float(r[i]), hdr[i][len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for i in cols if r[i]), reverse=True)[:6]
    print(f"{k:18s}", "  ".join(f"{n}={v:.2f}" for v, n in vals))
