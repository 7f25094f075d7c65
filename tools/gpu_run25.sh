#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 300 python tools/gpu_check.py attn_clock > gpurun_out/run25.log 2>&1
cat gpurun_out/run25.log
