#!/bin/bash
# builds the tcgen05.mma rate micro-benchmark (tools/mma_rate.cu) for sm_100a
set -e
cd "$(dirname "$0")"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o mma_rate mma_rate.cu
