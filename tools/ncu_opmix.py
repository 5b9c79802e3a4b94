#!/usr/bin/env python3
"""Executed-instruction mix of one kernel in an ncu report, per opcode and per barrier-delimited phase (SASS page).
This is the view that found k_fast's fixed per-thread costs, the 3-instruction __vsub2, k_describe's address arithmetic and
k_blur's constant-bank loads (profiles/r2_ncu_fast_v3.md, profiles/experiments/README.md).
usage: tools/ncu_opmix.py report.ncu-rep kernel_regex [launch_skip] [pixels_per_launch]"""
import collections
import csv
import subprocess
import sys

ALU = {"PRMT", "VIMNMX", "VIMNMX3", "LOP3", "SHF", "IADD3", "ISETP", "SEL", "VIADD", "LEA", "VIADDMNMX", "POPC", "FLO", "P2R",
       "R2P", "PLOP3", "VOTE", "IABS", "VABSDIFF", "VABSDIFF4"}
FMA = {"IMAD", "FFMA", "FMUL", "FADD", "HFMA2", "HADD2", "IDP"}


def main(rep, kern, skip="0", px=None):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--launch-skip", skip,
                          "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    ia, ie, iss = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
    data = []
    for r in rows[2:]:
        if r and r[0] == "Kernel Name":             # the next launch of the same kernel
            break
        try:
            data.append((r[ia].strip(), int(r[ie]), int(r[iss])))
        except (ValueError, IndexError):
            pass
    tot = sum(d[1] for d in data) or 1
    tots = sum(d[2] for d in data) or 1
    print(rows[0][1][:60], "warp instructions %.4e" % tot, "SASS", len(data))
    ops = collections.Counter()
    phases = [[0, 0, 0, 0]]
    for s, e, st in data:
        op = (s.split()[1] if s.startswith("@") else s.split()[0])
        ops[op] += e
        base = op.split(".")[0]
        p = phases[-1]
        p[0] += e; p[1] += st
        p[2] += e if base in ALU else 0
        p[3] += e if base in FMA else 0
        if "BAR.SYNC" in s:
            phases.append([0, 0, 0, 0])
    for i, p in enumerate(phases):
        per_px = "  %.1f thread instr / px" % (p[0] * 32 / float(px)) if px else ""
        print("phase %d: %5.1f %% of the instructions, %5.1f %% of the samples, ALU-pipe %2.0f %%, FMA-pipe %2.0f %%%s"
              % (i, 100 * p[0] / tot, 100 * p[1] / tots, 100 * p[2] / max(p[0], 1), 100 * p[3] / max(p[0], 1), per_px))
    for k, v in ops.most_common(24):
        print("  %-24s %5.1f %%" % (k, 100 * v / tot))


if __name__ == "__main__":
    main(*sys.argv[1:5])
