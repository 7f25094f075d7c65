#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run15.log
: > $L
echo "### pytest gpu gemm" >> $L
timeout 900 python -m pytest tests -x -q -m gpu -k "gemm or ffn or dense or dot or zz" 2>&1 | tail -3 >> $L
echo "### gemm_stages" >> $L
timeout 400 python tools/gpu_check.py gemm_stages 2>&1 | grep -E "stages=4|cublas" >> $L
echo "### bench" >> $L
timeout 600 python bench.py --no-cpu-baseline --no-attention > gpurun_out/bench_r01f.json 2>> $L
python - <<'PY' >> $L
import json
d=json.load(open('gpurun_out/bench_r01f.json'))
print('value',round(d['value'],1),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value'],1),'frac',round(d['roofline']['frac'],3),'clocks',d['clocks'])
PY
echo "### ncu full gemm" >> $L
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-attention"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 6 -c 2 -o gpurun_out/prof_gemm3 $CMD > gpurun_out/ncu_gemm3.log 2>&1
echo "exit=$?" >> $L
tail -30 $L
