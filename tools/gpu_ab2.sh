#!/usr/bin/env bash
# A/B of library variants on an arbitrary gpu_check mode: MODE=attn_bwd_time bash tools/gpu_ab2.sh
mkdir -p gpurun_out
L=gpurun_out/ab2.log
: > $L
pkg=learning-jax-sharding_b200
cp $pkg/libb200sl.so /tmp/libb200sl_default.so
for f in /tmp/libb200sl_default.so $pkg/variants/libb200sl_*.so; do
  cp $f $pkg/libb200sl.so
  echo "== $(basename $f)" >> $L
  timeout 200 python tools/gpu_check.py ${MODE:-attn_bwd_time} 2>&1 | grep "attn bwd\|rror" >> $L
done
cp /tmp/libb200sl_default.so $pkg/libb200sl.so
cat $L
