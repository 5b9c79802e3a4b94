#!/usr/bin/env python3
"""Per-source-line instruction / stall-sample shares of one kernel in an ncu report (needs -lineinfo).
usage: tools/ncu_hot_lines.py report.ncu-rep kernel_regex [top_n]"""
import csv
import subprocess
import sys


def main(rep, kern, top=40):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = next(r for r in rows if r and r[0] == "Line No")
    ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    lines = []
    for r in rows:
        if len(r) > ii and r[0].isdigit():
            try:
                lines.append((int(r[0]), r[1], int(r[ii]), int(r[si])))
            except ValueError:
                pass
    # several kernel instances repeat the listing: aggregate by line
    agg = {}
    for ln, src, inst, smp in lines:
        a = agg.setdefault(ln, [src, 0, 0])
        a[1] += inst
        a[2] += smp
    ti = sum(a[1] for a in agg.values()) or 1
    ts = sum(a[2] for a in agg.values()) or 1
    print(f"total warp instructions {ti}, stall samples {ts}")
    for ln, (src, inst, smp) in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
        print(f"{ln:5d} {100 * inst / ti:5.1f}% inst {100 * smp / ts:5.1f}% samples | {src.strip()[:110]}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40)
