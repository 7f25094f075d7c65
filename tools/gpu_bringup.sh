#!/usr/bin/env bash
# First-contact bring-up on the GPU box: every configuration in its own process under a timeout.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/bringup.log 2>&1
for cfg in "gemm --g 1 --bn 256 --layout nk" "gemm --g 1 --bn 256 --layout kn" "gemm --g 1 --bn 128 --layout kn" \
           "gemm --g 1 --bn 64 --layout kn" "gemm --g 2 --bn 256 --layout nk" "gemm --g 2 --bn 256 --layout kn" \
           "gemm --g 2 --bn 128 --layout kn" "f32" "gemm_time"; do
  echo "### $cfg" >> gpurun_out/bringup.log
  timeout 240 python tools/gpu_check.py $cfg >> gpurun_out/bringup.log 2>&1
  echo "exit=$?" >> gpurun_out/bringup.log
done
tail -150 gpurun_out/bringup.log
