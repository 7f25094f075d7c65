#!/usr/bin/env bash
# ncu: full capture of the attention v9 kernel + launch list of the case5 attention workload (each after a plain run)
mkdir -p gpurun_out
L=gpurun_out/run30.log
: > $L
CMD="python tools/gpu_check.py attn_prof"
timeout 300 $CMD > gpurun_out/plain_attn9.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:attention -s 1 -c 1 -o gpurun_out/prof_attn9 $CMD > gpurun_out/ncu_attn9.log 2>&1
echo "ncu exit=$?" >> $L
CMD2="python bench.py --workload case5_attention_dense --steps 2 --warmup 3"
timeout 300 $CMD2 > gpurun_out/plain_case5.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 9 -c 6 --csv --log-file gpurun_out/launches_attn9.csv $CMD2 > gpurun_out/ncu_launches_attn9.log 2>&1
echo "ncu launches exit=$?" >> $L
grep "^{" gpurun_out/plain_case5.log | cut -c1-900 >> $L
tail -5 $L
