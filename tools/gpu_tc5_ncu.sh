#!/bin/bash
# ncu --set full capture of the tcgen05 kNN kernel on a 32000 x 32000 sweep (plain run first).
mkdir -p gpurun_out
timeout 300 python tools/knn_sweep.py 32000 > gpurun_out/tc5_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_knn2_f4 -s 1 -c 1 -o gpurun_out/prof_tc5 -f python tools/knn_sweep.py 32000 > gpurun_out/ncu_tc5.log 2>&1
echo "rc $?"; tail -3 gpurun_out/ncu_tc5.log; cat gpurun_out/tc5_plain.log
