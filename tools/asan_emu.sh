#!/bin/bash
# AddressSanitizer pass over the kernels and host code on the CPU emulation build (no GPU needed; compute-sanitizer is
# not available on the GPU pool).  Exit code 0 and "ASAN RUN DONE" = clean.
set -e
cd "$(dirname "$0")/.."
mkdir -p /tmp/viorb_asan
g++ -O1 -g -fsanitize=address -fno-omit-frame-pointer -std=c++17 -fPIC -shared -DVIORB_EMU -D_GNU_SOURCE -Itests/cuda_emu \
    -x c++ vi_slam_b200/csrc/viorb_api.cu -o /tmp/viorb_asan/libviorb_emu_asan.so
ASAN_OPTIONS=detect_leaks=0:detect_stack_use_after_return=0 LD_PRELOAD=$(g++ -print-file-name=libasan.so) \
    python tools/asan_emu.py /tmp/viorb_asan/libviorb_emu_asan.so
