#!/usr/bin/env bash
# A/B run of the library variants under learning-jax-sharding_b200/variants/ (tools/build_variants.sh); G = attention variant
mkdir -p gpurun_out
L=gpurun_out/ab.log
: > $L
pkg=learning-jax-sharding_b200
cp $pkg/libb200sl.so /tmp/libb200sl_default.so
for round in 1 2; do
  for f in /tmp/libb200sl_default.so $pkg/variants/libb200sl_*.so; do
    cp $f $pkg/libb200sl.so
    echo "== $(basename $f) round $round" >> $L
    if [ $round -eq 1 ]; then timeout 120 python tools/gpu_check.py attn --g ${G:-9} 2>&1 | grep "RESULT\|FAIL\|rror" >> $L; fi
    timeout 120 python tools/gpu_check.py attn_quick --g ${G:-9} 2>&1 | grep "QUICK\|rror" >> $L
  done
done
cp /tmp/libb200sl_default.so $pkg/libb200sl.so
cat $L
