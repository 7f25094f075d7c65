#!/usr/bin/env bash
# streaming e2e + new attention defaults: parity tests, default bench, attention workload bench
mkdir -p gpurun_out
L=gpurun_out/run29.log
: > $L
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 >> $L
timeout 600 python bench.py > gpurun_out/bench_r01h.json 2>gpurun_out/bench_r01h.err; echo "bench exit=$?" >> $L
cut -c1-1800 gpurun_out/bench_r01h.json >> $L
timeout 600 python bench.py --workload case5_attention_dense --steps 20 --warmup 3 > gpurun_out/bench_attn_r01h.json 2>gpurun_out/bench_attn_r01h.err; echo "attn bench exit=$?" >> $L
cut -c1-1500 gpurun_out/bench_attn_r01h.json >> $L
cat $L
