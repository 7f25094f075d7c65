#!/bin/bash
# tcgen05.mma rate micro-benchmark
mkdir -p gpurun_out
timeout 120 tools/mma_rate > gpurun_out/mma_rate.log 2>&1; echo "exit $?" >> gpurun_out/mma_rate.log
cat gpurun_out/mma_rate.log
