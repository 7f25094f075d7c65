#!/usr/bin/env bash
N=${1:-8}
mkdir -p gpurun_out
L=gpurun_out/run19_n$N.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
for MODE in peer nccl; do
  echo "### bench fsdp $MODE n=$N" >> $L
  B200_FSDP_COLLECTIVE=$MODE timeout 600 $TR bench.py --gpus $N --steps 50 --warmup 5 --workload case3_fully_sharded 2>&1 | grep "^{" >> $L
done
echo "### bench tp n=$N" >> $L
timeout 600 $TR bench.py --gpus $N --steps 50 --warmup 5 2>&1 | grep "^{" >> $L
