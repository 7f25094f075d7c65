#!/usr/bin/env bash
# lean N-GPU scaling point: default (TP) bench, FSDP default, case6 attention
N=${1:-4}
mkdir -p gpurun_out
L=gpurun_out/run40_n$N.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577"
summ='import sys,json
d=json.loads(sys.stdin.read())
e=d.get("e2e") or {}
print(round(d["value"],1), "TFLOP/s", round(d["ms_per_step"],3), "ms | e2e", round(e.get("value",0),1), "|", d["config"]["parallelism"])'
for wl in case4_gspmd_ff case3_fully_sharded; do
  echo "### $wl" >> $L
  timeout 300 $TR bench.py --gpus $N --steps 50 --warmup 5 --workload $wl 2>&1 | grep '^{' | python -c "$summ" >> $L 2>&1
done
echo "### case6_attention" >> $L
timeout 300 $TR bench.py --gpus $N --steps 5 --warmup 3 --workload case6_attention 2>&1 | grep '^{' | python -c "$summ" >> $L 2>&1
cat $L
