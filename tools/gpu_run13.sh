#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run13.log
: > $L
echo "### pytest gpu" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 >> $L
echo "### smoke" >> $L
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2 >> $L
echo "### bench" >> $L
timeout 600 python bench.py > gpurun_out/bench_r01e.json 2>> $L
echo "exit=$?" >> $L
cut -c1-3200 gpurun_out/bench_r01e.json >> $L
echo "### ncu launches" >> $L
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-attention"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 6 -c 4 --csv --log-file gpurun_out/launches2.csv $CMD > gpurun_out/ncu_launches2.log 2>&1
echo "exit=$?" >> $L
echo "### ncu full gemm" >> $L
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 6 -c 2 -o gpurun_out/prof_gemm2 $CMD > gpurun_out/ncu_gemm2.log 2>&1
echo "exit=$?" >> $L
tail -30 $L
