#!/usr/bin/env bash
# final 1-GPU validation of the round: full GPU test suite, smoke, reference arm, default bench; then ncu of the backward kernel
mkdir -p gpurun_out
L=gpurun_out/run39.log
: > $L
timeout 1200 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 >> $L
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 >> $L
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | grep '^{' | cut -c1-400 >> $L
timeout 600 python bench.py > gpurun_out/bench_r01_final2.json 2>gpurun_out/bench_r01_final2.err; echo "bench exit=$?" >> $L
cut -c1-2600 gpurun_out/bench_r01_final2.json >> $L
CMD="python tools/gpu_check.py attn_bwd_time"
timeout 300 $CMD > gpurun_out/plain_bwd.log 2>&1; echo "bwd_time exit=$?" >> $L; grep "attn bwd" gpurun_out/plain_bwd.log >> $L
cat $L
