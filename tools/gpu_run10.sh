#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run10.log
: > $L
for cfg in "attn --g 6" "attn_time --g 6"; do
  echo "### $cfg" >> $L
  timeout 400 python tools/gpu_check.py $cfg >> $L 2>&1
  echo "exit=$?" >> $L
done
echo "### pytest gpu attention" >> $L
timeout 900 python -m pytest tests -x -q -m gpu -k "attention or mha or case6" 2>&1 | tail -4 >> $L
tail -40 $L
