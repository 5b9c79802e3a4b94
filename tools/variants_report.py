"""Prints the per-variant lines of the last tools/variants_bench.sh gpurun call (from gpurun_out/variants.log)."""
import ast
import re
import sys

base = None
for line in open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/variants.log"):
    m = re.match(r"build_variants/libviorb_(\w+)\.so (\d+) (\{.*\})", line)
    if not m:
        continue
    st = ast.literal_eval(m.group(3))
    if base is None:
        base = st
    diff = {k: round(v - base[k], 3) for k, v in st.items() if abs(v - base[k]) >= 0.004}
    print(f"{m.group(1):10s} {m.group(2):>6s}  total {sum(st.values()):.3f}  diff vs first: {diff}")
