#!/usr/bin/env bash
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/run9_n$N.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
echo "### mgpu_check" >> $L
timeout 600 $TR tools/mgpu_check.py >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench tp n=$N" >> $L
timeout 600 $TR bench.py --gpus $N --steps 50 --warmup 5 >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench fsdp peer n=$N" >> $L
timeout 600 $TR bench.py --gpus $N --steps 50 --warmup 5 --workload case3_fully_sharded >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench case6 n=$N" >> $L
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --workload case6_attention >> $L 2>&1
echo "exit=$?" >> $L
grep -v "Warning\|warn\|^W1\|OMP_NUM\|^\*\*\*" $L | grep -v "^$" | cut -c1-1500 | tail -60
