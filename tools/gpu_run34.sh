#!/usr/bin/env bash
# N-GPU validation of the round's late changes: mgpu_check, default (TP) bench, FSDP variants, case6 attention
N=${1:-8}
mkdir -p gpurun_out
L=gpurun_out/run34_n$N.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555"
echo "### mgpu_check" >> $L
timeout 600 $TR tools/mgpu_check.py 2>&1 | grep "OK\|FAIL\|RESULT\|rror" | grep -v "rank [1-9]" >> $L
summ='import sys,json
d=json.loads(sys.stdin.read())
e=d.get("e2e") or {}
print(round(d["value"],1), "TFLOP/s", round(d["ms_per_step"],3), "ms | e2e", round(e.get("value",0),1), "|", d["config"]["parallelism"])'
run() {
  wl=$1; shift
  echo "### $wl $*" >> $L
  env "$@" timeout 300 $TR bench.py --gpus $N --steps 50 --warmup 5 --workload $wl 2>&1 | grep '^{' | tee -a gpurun_out/run34_n${N}_lines.jsonl | python -c "$summ" >> $L 2>&1
}
: > gpurun_out/run34_n${N}_lines.jsonl
run case4_gspmd_ff X=1
run case3_fully_sharded B200_FSDP_COLLECTIVE=nccl B200_FSDP_PREFETCH=0
run case3_fully_sharded B200_FSDP_COLLECTIVE=nccl
run case3_fully_sharded B200_FSDP_COLLECTIVE=nccl B200_FSDP_GEMM_CTAS=132 NCCL_MAX_NCHANNELS=16
run case3_fully_sharded B200_FSDP_COLLECTIVE=nccl B200_FSDP_CHUNKS=2
run case3_fully_sharded B200_FSDP_COLLECTIVE=peer
echo "### case6" >> $L
timeout 300 $TR bench.py --gpus $N --steps 5 --warmup 3 --workload case6_attention 2>&1 | grep '^{' | tee -a gpurun_out/run34_n${N}_lines.jsonl | python -c "$summ" >> $L 2>&1
cat $L
