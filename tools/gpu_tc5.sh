#!/bin/bash
# tcgen05 kNN kernel: correctness against the oracle (default split choice and forced splits), the N x N sweep, the short bench.
mkdir -p gpurun_out
timeout 300 python tools/knn_tc5_check.py > gpurun_out/tc5_check.log 2>&1; echo "check exit $?" >> gpurun_out/tc5_check.log
for s in 1 2 5; do VIORB_KNN_SPLIT=$s timeout 300 python tools/knn_tc5_check.py 2>&1 | tail -1 | sed "s/^/split $s: /" >> gpurun_out/tc5_check.log; done
tail -8 gpurun_out/tc5_check.log
if grep -q "check exit 0" gpurun_out/tc5_check.log; then
  timeout 300 python tools/knn_sweep.py 2000 4000 8000 16000 25000 32000 100000 200000 > gpurun_out/tc5_sweep.jsonl 2>&1
  VIORB_KNN_SPLIT=1 timeout 300 python tools/knn_sweep.py 2000 8000 32000 >> gpurun_out/tc5_sweep.jsonl 2>&1
  cut -c1-120 gpurun_out/tc5_sweep.jsonl
  bash tools/gpu_quick.sh
fi
