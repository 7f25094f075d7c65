#!/usr/bin/env bash
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-attention"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 6 -c 2 -o gpurun_out/prof_gemm4 $CMD > gpurun_out/ncu_gemm4.log 2>&1
echo "exit=$?"
