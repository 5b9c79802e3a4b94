#!/bin/bash
# 2-GPU box: NCCL parity test of the row-sharded matcher, then the bench as the driver launches it at N=2 (knn_sweep leg included).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm --format=csv > gpurun_out/gpu2.txt 2>&1
timeout 900 python -m pytest tests/test_sharded_nccl.py -x -q -m gpu -s > gpurun_out/pytest_nccl.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_nccl.log
tail -12 gpurun_out/pytest_nccl.log
N=${N:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
echo "bench exit $?"; tail -3 gpurun_out/bench_${N}gpu.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_${N}gpu.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], json.dumps(d.get("knn_sweep")))
PY
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_1gpu_sweep.json 2> gpurun_out/bench_1gpu_sweep.err
python - <<PY
import json
d = json.loads(open("gpurun_out/bench_1gpu_sweep.json").read().strip().splitlines()[-1])
print(d["value"], json.dumps(d.get("knn_sweep")))
PY
