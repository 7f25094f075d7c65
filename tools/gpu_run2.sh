#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run2.log
: > $L
for cfg in "attn" "attn_time"; do
  echo "### $cfg" >> $L
  timeout 300 python tools/gpu_check.py $cfg >> $L 2>&1
  echo "exit=$?" >> $L
done
echo "### pytest gpu" >> $L
timeout 900 python -m pytest tests -q -m gpu 2>&1 | tail -60 >> $L
echo "### smoke" >> $L
timeout 300 python __graft_entry__.py smoke >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench" >> $L
timeout 600 python bench.py > gpurun_out/bench_r01.json 2>> $L
echo "exit=$?" >> $L
cat gpurun_out/bench_r01.json >> $L
echo "### bench reference" >> $L
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 >> $L 2>&1
tail -120 $L
