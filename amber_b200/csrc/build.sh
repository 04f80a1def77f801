#!/bin/sh
# Build libamber_cuda.so for sm_100a (in-tree; travels to the GPU box with gpurun).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -Xcompiler -fPIC,-Wall,-Wno-unused-function,-pthread -shared -cudart static \
  ${AMBER_NVCC_FLAGS} -o ${AMBER_OUT:-libamber_cuda.so} amber_cuda.cu -lpthread
