#!/bin/bash
# End-to-end rate against the number of H2D / kernel / D2H chunks per viorb_extract_batch call (VIORB_CHUNKS knob).
for c in 1 2 3 4 6 8; do VIORB_CHUNKS=$c python bench.py --steps 12 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('chunks $c', 'e2e', round(d['e2e']['value']), 'ms', round(d['e2e']['ms_per_step'],3), 'resident', round(d['value']))"; done
