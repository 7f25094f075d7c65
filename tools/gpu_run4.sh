#!/usr/bin/env bash
# 1 GPU: attention v2 + tests + GEMM raster sweep
mkdir -p gpurun_out
L=gpurun_out/run4.log
: > $L
for cfg in "attn" "attn_time" "gemm_group"; do
  echo "### $cfg" >> $L
  timeout 400 python tools/gpu_check.py $cfg >> $L 2>&1
  echo "exit=$?" >> $L
done
echo "### pytest gpu" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -30 >> $L
tail -100 $L
