#!/usr/bin/env bash
N=${1:-4}
mkdir -p gpurun_out
L=gpurun_out/run36_n$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566"
timeout 600 $TR tools/mgpu_check.py > $L 2>&1; echo "exit=$?" >> $L
grep "OK\|FAIL\|RESULT\|rror\|exit=" $L | grep -v "rank [1-9]" | cut -c1-300
