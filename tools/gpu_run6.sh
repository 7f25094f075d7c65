#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run6.log
: > $L
echo "### pytest gpu" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -30 >> $L
echo "### bench" >> $L
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_r01c.json 2>> $L
echo "exit=$?" >> $L
cut -c1-1600 gpurun_out/bench_r01c.json >> $L
echo "### gemm_time" >> $L
timeout 300 python tools/gpu_check.py gemm_time 2>&1 | grep -E "bn=256|cublas" >> $L
tail -60 $L
