#!/usr/bin/env bash
# N GPUs: multi-GPU bring-up + bench under torchrun
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/run5_n$N.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
echo "### mgpu_check" >> $L
timeout 600 $TR tools/mgpu_check.py >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench tp n=$N" >> $L
timeout 600 $TR bench.py --gpus $N --steps 30 --warmup 5 >> $L 2>&1
echo "exit=$?" >> $L
echo "### bench fsdp n=$N" >> $L
timeout 600 $TR bench.py --gpus $N --steps 30 --warmup 5 --workload case3_fully_sharded >> $L 2>&1
echo "exit=$?" >> $L
grep -v "^W1\|^\[W\|warn" $L | tail -60
