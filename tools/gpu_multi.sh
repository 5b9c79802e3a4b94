#!/bin/bash
# 8-GPU box: the bench as the driver launches it at N = 1, 2, 4, 8 (default workload: weak scaling; knn_sweep = configs[4]
# rides along), then the 4096-frame strong-scaling batch (configs[3]) at N = 1 and 8.
mkdir -p gpurun_out
run() {  # N workload tag
  local n=$1 w=$2 tag=$3
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${tag}_1gpu.json 2> gpurun_out/bench_${tag}_1gpu.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --workload $w --steps 10 --warmup 3 > gpurun_out/bench_${tag}_${n}gpu.json 2> gpurun_out/bench_${tag}_${n}gpu.err
  fi
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_${tag}_${n}gpu.json").read().strip().splitlines()[-1])
    ks = d.get("knn_sweep", {}).get("sizes", [])
    print("$tag N=$n", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms/step", round(d["ms_per_step"], 3),
          "knn", [(s["workload"][13:], round(s["ms"], 3), s["match"]) for s in ks])
except Exception as e:
    print("$tag N=$n: no line:", e)
PY
}
for n in ${NS:-1 2 4 8}; do run $n kitti_stereo stereo; done
for n in ${NS2:-1 8}; do run $n kitti_batch_4096 batch4096; done
