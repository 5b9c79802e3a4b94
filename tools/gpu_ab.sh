#!/bin/bash
# Extraction parity tests, then the default bench (resident + e2e) and, with BENCH2, a second workload.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_ref_pin.py tests/test_f2f.py -x -q -m gpu 2>&1 | tail -2
for w in kitti_stereo ${BENCH2}; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-knn-sweep ${BENCH_ARGS} 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('[$w]', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']) if d['e2e'] else None)
print({k: round(v, 3) for k, v in d['stages_ms_per_step'].items()})"
done
