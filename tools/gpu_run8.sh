#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run8.log
: > $L
echo "### pytest gpu" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -30 >> $L
echo "### bench" >> $L
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_r01d.json 2>> $L
echo "exit=$?" >> $L
cut -c1-3000 gpurun_out/bench_r01d.json >> $L
echo "### bench case5" >> $L
timeout 600 python bench.py --workload case5_attention_dense --steps 10 >> $L 2>&1
echo "### bench case6" >> $L
timeout 600 python bench.py --workload case6_attention --steps 5 --warmup 3 >> $L 2>&1
tail -60 $L
