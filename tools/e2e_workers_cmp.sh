for w in 1 2 3; do python bench.py --steps 12 --warmup 3 --no-cpu-baseline --e2e-workers $w 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('workers', d['e2e'].get('workers'), 'e2e', round(d['e2e']['value']), 'ms', round(d['e2e']['ms_per_step'],3), 'resident', round(d['value']))"; done
