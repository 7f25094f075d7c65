#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run3.log
: > $L
echo "### zz alone" >> $L
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -15 >> $L
echo "### pytest gpu all (-x)" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 >> $L
echo "### bench" >> $L
timeout 600 python bench.py > gpurun_out/bench_r01b.json 2>> $L
echo "exit=$?" >> $L
cat gpurun_out/bench_r01b.json >> $L
echo "### ncu launches" >> $L
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "exit=$?" >> $L
echo "### ncu full gemm" >> $L
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 6 -c 2 -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "exit=$?" >> $L
echo "### ncu full attention" >> $L
CMD2="python tools/gpu_check.py attn_prof"
timeout 300 $CMD2 > gpurun_out/plain3.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:attention -s 1 -c 1 -o gpurun_out/prof_attn $CMD2 > gpurun_out/ncu_attn.log 2>&1
echo "exit=$?" >> $L
tail -5 gpurun_out/ncu_gemm.log gpurun_out/ncu_attn.log >> $L
tail -60 $L
