#!/usr/bin/env bash
# ncu launch list of the case5 attention workload (own kernels only), after a plain run
mkdir -p gpurun_out
CMD2="python bench.py --workload case5_attention_dense --steps 2 --warmup 3"
timeout 300 $CMD2 > gpurun_out/plain_case5.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'gemm_bf16|attention_fwd' -s 9 -c 6 --csv --log-file gpurun_out/launches_attn9.csv $CMD2 > gpurun_out/ncu_launches_attn9.log 2>&1
echo "ncu launches exit=$?"
tail -3 gpurun_out/launches_attn9.csv | cut -c1-300
