#!/bin/bash
# CTA-pair form of the tcgen05 kNN kernel (VIORB_KNN=tc5x2): correctness against the oracle, then timing against the default.
mkdir -p gpurun_out
timeout 200 python tools/knn_tc5_check.py 2>&1 | tail -1
VIORB_KNN=tc5x2 timeout 90 python tools/knn_tc5_check.py > gpurun_out/tc5x2_check.log 2>&1; echo "pair check exit $?"
tail -16 gpurun_out/tc5x2_check.log | cut -c1-200
if grep -q "mismatching shapes: 0" gpurun_out/tc5x2_check.log; then
  VIORB_KNN=tc5x2 timeout 200 python tools/knn_shard_shape.py | cut -c1-120
  VIORB_KNN=tc5x2 timeout 200 python tools/knn_sweep.py 32000 100000 200000 | cut -c1-120
  timeout 200 python tools/knn_sweep.py 200000 | cut -c1-120
fi
