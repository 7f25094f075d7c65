#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run12.log
: > $L
echo "### pytest gpu gemm" >> $L
timeout 900 python -m pytest tests -x -q -m gpu -k "gemm or ffn or dense or dot" 2>&1 | tail -5 >> $L
for cfg in "gemm_time" "gemm_stages"; do
  echo "### $cfg" >> $L
  timeout 400 python tools/gpu_check.py $cfg 2>&1 | grep -v "bn=128" >> $L
  echo "exit=$?" >> $L
done
tail -60 $L
