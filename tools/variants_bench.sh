#!/bin/bash
# Runs the short bench once per library in build_variants/ and prints throughput + the stage table.
for lib in build_variants/libviorb_*.so; do
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --variant-lib $lib 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$lib', round(d['value']), {k: round(v, 3) for k, v in d['stages_ms_per_step'].items()})"
done
