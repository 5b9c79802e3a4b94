#!/bin/bash
# A/B of libviorb variants (build_variants/) on the tcgen05 kNN kernel: correctness check, then the sweep sizes in $SIZES.
mkdir -p gpurun_out; : > gpurun_out/tc5_variants.log
for lib in build_variants/libviorb_*.so; do
  echo "== $lib" | tee -a gpurun_out/tc5_variants.log
  VIORB_TOOL_LIB=$lib timeout 300 python tools/knn_tc5_check.py 2>&1 | tail -2 | tee -a gpurun_out/tc5_variants.log
  VIORB_TOOL_LIB=$lib timeout 300 python tools/knn_sweep.py ${SIZES:-2000 32000 200000} 2>&1 | cut -c1-130 | tee -a gpurun_out/tc5_variants.log
done
