#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run10.log
: > $L
echo "### pytest gpu" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
echo "### attn_time auto" >> $L
timeout 400 python tools/gpu_check.py attn_time >> $L 2>&1
tail -20 $L
