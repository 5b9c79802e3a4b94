#!/bin/bash
# tcgen05 kNN kernel: correctness against the oracle, then A/B timing against the warp-MMA kernel.
mkdir -p gpurun_out
timeout 300 python tools/knn_tc5_check.py > gpurun_out/tc5_check.log 2>&1; echo "check exit $?" >> gpurun_out/tc5_check.log
tail -20 gpurun_out/tc5_check.log
if grep -q "check exit 0" gpurun_out/tc5_check.log; then
  timeout 300 python tools/knn_sweep.py 2000 8000 32000 100000 200000 > gpurun_out/tc5_sweep.jsonl 2>&1
  VIORB_KNN=mma timeout 300 python tools/knn_sweep.py 2000 32000 200000 >> gpurun_out/tc5_sweep.jsonl 2>&1
  cat gpurun_out/tc5_sweep.jsonl
  bash tools/gpu_quick.sh
fi
