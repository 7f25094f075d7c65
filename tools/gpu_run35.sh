#!/usr/bin/env bash
# Dense train_step slice (8f row 1): parity tests, kernel timings
mkdir -p gpurun_out
L=gpurun_out/run35.log
: > $L
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "tn or adam or train" 2>&1 | tail -15 >> $L
timeout 300 python tools/gpu_check.py train_time 2>&1 | grep -v "^$" >> $L
cat $L
