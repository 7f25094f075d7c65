#!/usr/bin/env bash
mkdir -p gpurun_out
L=gpurun_out/run37.log
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "attention_backward or train_step" 2>&1 | tail -40 > $L
cat $L
