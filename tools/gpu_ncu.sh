#!/bin/bash
# ncu evidence: launch list of a short bench run, then a full capture of selected kernels.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --pairs 64 --no-cpu-baseline --no-e2e --no-knn-sweep ${BENCH_ARGS}"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launches rc $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-k_fast|k_blur}" -s ${SKIP:-6} -c ${COUNT:-2} -o gpurun_out/prof -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc $?"
tail -3 gpurun_out/ncu_full.log
