#!/bin/bash
# End-to-end rate over pipeline workers x chunks per batch call.
for w in 2 3 4 5; do for c in 1 2; do VIORB_CHUNKS=$c python bench.py --steps 16 --warmup 3 --no-cpu-baseline --e2e-workers $w 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('workers $w chunks $c', 'e2e', round(d['e2e']['value']), 'ms', round(d['e2e']['ms_per_step'],3), 'steps', d['e2e']['steps'], 'resident', round(d['value']))"; done; done
