#!/bin/bash
# Builds A/B variants of the library under build_ab/ (git-ignored, travels with gpurun): build_ab.sh name "-Dflags" ...
# "head" = the committed sources (git stash around the build).
set -e
cd "$(dirname "$0")/../rust-ray-tracing_b200/csrc"
mkdir -p ../../build_ab
while [ $# -gt 0 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -std=c++17 -Xcompiler -fPIC $flags -shared -o ../../build_ab/librtb200_$name.so rtb200.cu
  echo "built $name ($flags)"
done
