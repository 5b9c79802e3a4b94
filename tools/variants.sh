#!/bin/bash
# Builds libviorb variants with extra -D flags into build_variants/ (git-ignored) for A/B runs on the GPU box:
#   tools/variants.sh name1 "-DX=1 -DY=2" name2 "-DX=3" ...
# then on the box: VIORB_LIBRARY=build_variants/libviorb_name1.so python bench.py ...
set -e
cd "$(dirname "$0")/.."
mkdir -p build_variants
while [ $# -ge 2 ]; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $2 \
      -o build_variants/libviorb_$1.so vi_slam_b200/csrc/viorb_api.cu &
  shift 2
done
wait
ls -la build_variants
