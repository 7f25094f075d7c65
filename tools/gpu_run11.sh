#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run11.log
: > $L
for cfg in "attn_time --g 2" "attn_time --g 3" "gemm_time"; do
  echo "### $cfg" >> $L
  timeout 400 python tools/gpu_check.py $cfg 2>&1 | grep -v "bn=128" >> $L
  echo "exit=$?" >> $L
done
echo "### pytest gpu" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 >> $L
tail -60 $L
