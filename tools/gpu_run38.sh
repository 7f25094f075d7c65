#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 300 python tools/gpu_check.py attn_bwd_time 2>&1 | grep -v "^$" > gpurun_out/run38.log
cat gpurun_out/run38.log
