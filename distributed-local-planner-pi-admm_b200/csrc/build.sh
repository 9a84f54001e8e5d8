#!/bin/bash
# Builds libpiadmm.so (sm_100a only) in-tree.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC \
  -Xptxas -v ${NVCC_EXTRA} -o ${PIADMM_OUT:-libpiadmm.so} piadmm.cu obca_consensus.cu piadmm_static.cu 2> build.log || { cat build.log; exit 1; }
grep -E "error|warning|Used|spill" build.log | grep -v "^$" | head -60
