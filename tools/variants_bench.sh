#!/bin/bash
# Runs the short bench once per library in build_variants/ (base first) and logs throughput + the stage table.
mkdir -p gpurun_out; : > gpurun_out/variants.log
for lib in build_variants/libviorb_base.so $(ls build_variants/libviorb_*.so | grep -v _base.so); do
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-knn-sweep ${BENCH_ARGS} --variant-lib $lib 2>/dev/null | python tools/variants_line.py $lib | tee -a gpurun_out/variants.log
done
