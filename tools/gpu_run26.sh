#!/usr/bin/env bash
# attention v8 (persistent v7): correctness, parity tests, timing
mkdir -p gpurun_out
L=gpurun_out/run26.log
: > $L
timeout 120 python tools/gpu_check.py attn --g 8 >> $L 2>&1; rc=$?; echo "attn g8 exit=$rc" >> $L
if [ $rc -eq 0 ]; then
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "attention" -x 2>&1 | tail -5 >> $L
for g in 8 7; do
  echo "== attn_time variant $g" >> $L
  timeout 200 python tools/gpu_check.py attn_time --g $g 2>&1 | grep -v "^$" >> $L
done
timeout 300 python tools/gpu_check.py attn_clock >> $L 2>&1
fi
cat $L
