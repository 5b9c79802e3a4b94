#!/usr/bin/env python3
"""Instruction / stall-sample share of source-line ranges of one kernel (needs -lineinfo).
usage: tools/ncu_phase.py report.ncu-rep kernel_regex name:lo-hi[,lo-hi] ..."""
import csv
import subprocess
import sys


def main(rep, kern, specs):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = next(r for r in rows if r and r[0] == "Line No")
    ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    agg = {}
    for r in rows:
        if len(r) > ii and r[0].isdigit():
            try:
                a = agg.setdefault(int(r[0]), [0, 0])
                a[0] += int(r[ii]); a[1] += int(r[si])
            except ValueError:
                pass
    ti = sum(a[0] for a in agg.values()) or 1
    ts = sum(a[1] for a in agg.values()) or 1
    used = set()
    print(f"total warp instructions {ti}, samples {ts}")
    for sp in specs:
        name, rng = sp.split(":")
        inst = smp = 0
        for part in rng.split(","):
            lo, hi = map(int, part.split("-"))
            for ln, a in agg.items():
                if lo <= ln <= hi:
                    inst += a[0]; smp += a[1]; used.add(ln)
        print(f"{name:12s} {100*inst/ti:5.1f}% inst {100*smp/ts:5.1f}% samples")
    inst = sum(a[0] for ln, a in agg.items() if ln not in used)
    smp = sum(a[1] for ln, a in agg.items() if ln not in used)
    print(f"{'other':12s} {100*inst/ti:5.1f}% inst {100*smp/ts:5.1f}% samples")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3:])
