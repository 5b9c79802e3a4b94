#!/bin/bash
# ncu --set full capture of one kernel (KREGEX) of a short bench run (plain run first): KREGEX=k_octree_dense BENCH_ARGS="--workload zed_stereo_4000kp" OUT=prof_oct
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-knn-sweep ${BENCH_ARGS}"
$CMD > gpurun_out/plain_k.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-k_fast} -s ${SKIP:-4} -c ${COUNT:-1} -o gpurun_out/${OUT:-prof_k} -f $CMD > gpurun_out/ncu_k.log 2>&1
echo "full rc $?"
tail -2 gpurun_out/ncu_k.log
