for cfg in "3 12" "3 30" "4 32" "2 30"; do set -- $cfg; python bench.py --steps $2 --warmup 3 --no-cpu-baseline --e2e-workers $1 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('workers', d['e2e'].get('workers'), 'steps', d['steps'], 'e2e', round(d['e2e']['value']), 'ms', round(d['e2e']['ms_per_step'],3), 'resident', round(d['value']))"; done
