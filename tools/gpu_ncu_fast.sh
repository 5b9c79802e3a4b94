#!/bin/bash
# ncu --set full capture of k_fast only (plain run first); BENCH_ARGS / --variant-lib as in gpu_ncu.sh
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --pairs 64 --no-cpu-baseline --no-e2e --no-knn-sweep ${BENCH_ARGS}"
$CMD > gpurun_out/plain_fast.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_fast -s ${SKIP:-4} -c ${COUNT:-1} -o gpurun_out/${OUT:-prof_fast} -f $CMD > gpurun_out/ncu_fast.log 2>&1
echo "full rc $?"
tail -2 gpurun_out/ncu_fast.log
