#!/usr/bin/env bash
# serpentine K order in the GEMM: correctness + timing, then DRAM traffic under ncu; plus the attention-workload launch list
mkdir -p gpurun_out
L=gpurun_out/run32.log
: > $L
timeout 300 python tools/gpu_check.py gemm_serp >> $L 2>&1; echo "gemm_serp exit=$?" >> $L
timeout 300 python tools/gpu_check.py gemm_traffic > gpurun_out/plain_traffic.log 2>&1 && \
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:gemm_bf16 --csv --log-file gpurun_out/gemm_traffic_serp.csv python tools/gpu_check.py gemm_traffic > gpurun_out/ncu_traffic_serp.log 2>&1
echo "ncu traffic exit=$?" >> $L
bash tools/gpu_run31.sh >> $L 2>&1
cat $L
