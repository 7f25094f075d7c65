#!/usr/bin/env bash
N=${1:-4}
mkdir -p gpurun_out
L=gpurun_out/run17_n$N.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
echo "### pytest gpu (rank-less, 1 GPU of the box)" >> $L
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
echo "### mgpu_check" >> $L
timeout 600 $TR tools/mgpu_check.py 2>&1 | grep -E "FAIL|RESULT|Error|error" >> $L
echo "### bench reference n=$N" >> $L
timeout 600 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 2>&1 | grep "^{" | cut -c1-300 >> $L
for W in case4_gspmd_ff case3_fully_sharded case6_attention; do
  echo "### bench $W n=$N" >> $L
  ST=50; [ $W = case6_attention ] && ST=5
  timeout 600 $TR bench.py --gpus $N --steps $ST --warmup 5 --workload $W 2>&1 | grep "^{" >> $L
done
cut -c1-400 $L | tail -30
