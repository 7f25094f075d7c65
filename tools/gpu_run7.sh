#!/usr/bin/env bash
N=${1:-8}
mkdir -p gpurun_out
L=gpurun_out/run7_n$N.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
echo "### mgpu_check" >> $L
timeout 600 $TR tools/mgpu_check.py >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench tp fused n=$N" >> $L
timeout 600 $TR bench.py --gpus $N --steps 50 --warmup 5 >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench tp nccl n=$N" >> $L
B200_TP_COLLECTIVE=nccl timeout 600 $TR bench.py --gpus $N --steps 50 --warmup 5 >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench fsdp n=$N" >> $L
timeout 600 $TR bench.py --gpus $N --steps 50 --warmup 5 --workload case3_fully_sharded >> $L 2>&1
echo "exit=$?" >> $L
grep -v "Warning\|warn\|^W1\|OMP_NUM\|^\*\*\*" $L | cut -c1-600 | tail -60
