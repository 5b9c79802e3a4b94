#!/usr/bin/env python3
"""Summarise an ncu report (raw page) into a small markdown table for profiles/.
usage: tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r1_ncu_<tag>.md ["note"]"""
import csv
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "time"), ("launch__grid_size", "grid"), ("launch__registers_per_thread", "regs"),
        ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma %"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu %"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu %"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex %"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2 %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"),
        ("smsp__inst_executed.sum", "warp inst")]


def main(rep, out, note=""):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    with open(out, "w") as f:
        f.write(f"# ncu --set full summary ({rep.split('/')[-1]})\n\n{note}\n\n")
        f.write("| kernel | " + " | ".join(n for _, n in COLS) + " |\n|---|" + "---|" * len(COLS) + "\n")
        for r in rows[2:]:
            cells = []
            for key, _ in COLS:
                if key in hdr:
                    i = hdr.index(key)
                    v = r[i]
                    try:
                        v = f"{float(v):.4g}"
                    except ValueError:
                        pass
                    cells.append(f"{v} {units[i]}".strip())
                else:
                    cells.append("-")
            f.write(f"| {r[ki].split('(')[0]} | " + " | ".join(cells) + " |\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
