#!/bin/bash
# Builds acinoset_b200/libacino_b200.so for sm_100a (nvcc cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 ${NVCC_EXTRA} \
     -shared -Xcompiler -fPIC -o acinoset_b200/libacino_b200.so acinoset_b200/csrc/acino.cu -ldl
