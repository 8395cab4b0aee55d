#!/bin/bash
# Build an experimental libccgpu variant: scripts/build_variant.sh <name> [-DMACRO=VALUE ...]
# The library lands in variants/<name>.so (travels to the GPU box; select it with CCGPU_LIB).
NAME=$1; shift
mkdir -p variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -cudart static \
  "$@" -Xptxas -v -o variants/$NAME.so consensuscruncher_b200/csrc/ccgpu.cu 2>&1 | grep -A2 "vote_small_kernelILi8ELi1ELi[0-9]ELb0" | grep -E "spill|Used"
