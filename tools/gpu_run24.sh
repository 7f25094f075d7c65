#!/usr/bin/env bash
# phase profile of attention variants (library built with -DB200_ATTN_PROFILE=1); VARIANTS="9 7"
mkdir -p gpurun_out
L=gpurun_out/run24.log
: > $L
for g in ${VARIANTS:-9}; do
  timeout 200 python tools/gpu_check.py attn_phases --g $g 2>&1 | grep -v "^$" >> $L
done
cat $L
