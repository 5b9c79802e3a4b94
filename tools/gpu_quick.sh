#!/bin/bash
# Parity tests of the extraction / matching path + a short bench; prints throughput and the stage table.
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'])
print({k: round(v, 3) for k, v in d['stages_ms_per_step'].items()})"
