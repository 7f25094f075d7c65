#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run28.log
: > $L
timeout 120 python tools/gpu_check.py attn --g 9 >> $L 2>&1; rc=$?; echo "attn g9 exit=$rc" >> $L
if [ $rc -eq 0 ]; then
  timeout 200 python tools/gpu_check.py attn_time --g 9 2>&1 | grep -v "^$" >> $L
fi
cat $L
