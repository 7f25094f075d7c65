#!/usr/bin/env bash
# FSDP (case3): forward prefetch on/off, chunk count, GEMM grid bound, for NCCL all-gather and copy-engine pulls
N=${1:-2}
mkdir -p gpurun_out
L=gpurun_out/run33_n$N.log
: > $L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
echo "### mgpu_check" >> $L
timeout 600 $TR tools/mgpu_check.py 2>&1 | grep "FSDP\|RESULT\|rror" >> $L
run() {
  echo "### $*" >> $L
  env "$@" timeout 300 $TR bench.py --gpus $N --steps 50 --warmup 5 --workload case3_fully_sharded 2>&1 | grep '^{' | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'TFLOP/s', round(d['ms_per_step'],3), 'ms | e2e', round(d['e2e']['value'],1), '|', d['config']['parallelism'])" >> $L 2>&1
}
run B200_FSDP_COLLECTIVE=nccl B200_FSDP_PREFETCH=0
run B200_FSDP_COLLECTIVE=nccl
run B200_FSDP_COLLECTIVE=nccl B200_FSDP_CHUNKS=2
run B200_FSDP_COLLECTIVE=nccl B200_FSDP_GEMM_CTAS=132 NCCL_MAX_NCHANNELS=16
run B200_FSDP_COLLECTIVE=peer B200_FSDP_PREFETCH=0
run B200_FSDP_COLLECTIVE=peer
run B200_FSDP_COLLECTIVE=peer B200_FSDP_CHUNKS=2
cat $L
