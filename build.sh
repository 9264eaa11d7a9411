#!/bin/bash
# Builds libshifu_b200.so (sm_100a only) and the CPU oracle.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
SRC="shifu_b200/csrc/api.cu shifu_b200/csrc/rowlist.cu shifu_b200/csrc/hist.cu shifu_b200/csrc/rows.cu shifu_b200/csrc/split.cu shifu_b200/csrc/noalloc.cu shifu_b200/csrc/prep.cu shifu_b200/csrc/score.cu shifu_b200/csrc/host_master.cpp"
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false \
    -Xcompiler -fPIC,-fvisibility=hidden -shared -o shifu_b200/libshifu_b200.so $SRC "$@"
make -s -C oracle
# a host program on the bare C ABI (no Python): see tools/train_gbt.cpp
g++ -O2 -std=c++17 -o tools/train_gbt tools/train_gbt.cpp -Lshifu_b200 -lshifu_b200 -Wl,-rpath,'$ORIGIN/../shifu_b200'
