#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run21.log
: > $L
echo "### pytest gpu" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
echo "### smoke" >> $L
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1 >> $L
echo "### bench reference" >> $L
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 2>&1 | grep "^{" | cut -c1-400 >> $L
echo "### bench" >> $L
timeout 600 python bench.py > gpurun_out/bench_r01_final.json 2>> $L
echo "exit=$?" >> $L
CMD2="python bench.py --workload case5_attention_dense --steps 2 --warmup 3"
timeout 300 $CMD2 > gpurun_out/plain4.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemm_bf16|attention" -s 9 -c 6 --csv --log-file gpurun_out/launches_attn.csv $CMD2 > gpurun_out/ncu_launches_attn.log 2>&1
echo "ncu launches exit=$?" >> $L
tail -12 $L | cut -c1-600
