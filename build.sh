#!/bin/bash
# Builds libshifu_b200.so (sm_100a only), the CPU oracle and tools/train_gbt; see Makefile.
set -e
cd "$(dirname "$0")"
make -j"$(nproc)" "$@"
