#!/bin/bash
# 1-GPU box: all parity tests, then one bench line per workload (BASELINE.json configs[0..3]; configs[4] rides along as knn_sweep).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for w in ${WORKLOADS:-kitti_stereo kitti_mono_f2f zed_stereo_4000kp kitti_batch_4096}; do
  timeout 900 python bench.py --workload $w --steps ${STEPS:-10} --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w exit $?"
  tail -2 gpurun_out/bench_$w.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_$w.json").read().strip().splitlines()[-1])
    print(round(d["value"]), d["unit"], "e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 3), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 4),
          "cpu", round(d["cpu_baseline"]["value"], 1), "cv2", d["cpu_baseline"]["cv2_floor"].get("value"))
    print({k: round(v, 3) for k, v in d["stages_ms_per_step"].items()})
    print({k: v for k, v in d["config"].items() if "per" in k})
except Exception as e:
    print("no line:", e)
PY
done
