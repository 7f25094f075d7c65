#!/usr/bin/env bash
# attention v7 (alternate-tile softmax groups): correctness, parity tests, timing against v5
mkdir -p gpurun_out
L=gpurun_out/run23.log
: > $L
timeout 120 python tools/gpu_check.py attn --g 7 >> $L 2>&1; echo "attn g7 exit=$?" >> $L
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "attention" -x 2>&1 | tail -5 >> $L
for g in 7 5 7 5; do
  echo "== attn_time variant $g" >> $L
  timeout 200 python tools/gpu_check.py attn_time --g $g 2>&1 | grep -v "^$" >> $L
done
cat $L
