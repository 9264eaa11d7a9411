#!/bin/bash
# Developer tool: build/exp/<name>.so = the library with hist.cu compiled with extra -D flags (kernel experiments).
# usage: tools/build_variant.sh <name> [-DSB_... ...]; run with tools/perf_probe.py --lib build/exp/<name>.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build/exp
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC,-fvisibility=hidden "$@" -c -o build/exp/hist_$name.o shifu_b200/csrc/hist.cu
objs=$(ls build/*.o | grep -v build/hist.o)
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o build/exp/$name.so $objs build/exp/hist_$name.o
echo build/exp/$name.so
