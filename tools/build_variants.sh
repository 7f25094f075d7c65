#!/usr/bin/env bash
# Builds A/B variants of the library for one gpurun call: tools/build_variants.sh name1="-DFLAG=1" name2="-DX=2 -DY=3" ...
# Each lands in learning-jax-sharding_b200/variants/libb200sl_<name>.so; the default build is restored at the end.
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
pkg="$here/learning-jax-sharding_b200"
mkdir -p "$pkg/variants"
files="${VARIANT_FILES:-attention9.cu}"
for spec in "$@"; do
  name="${spec%%=*}"; flags="${spec#*=}"
  for f in $files; do touch "$pkg/csrc/$f"; done
  B200_EXTRA_NVCC_FLAGS="$flags" bash "$pkg/build.sh" > /dev/null
  cp "$pkg/libb200sl.so" "$pkg/variants/libb200sl_$name.so"
  echo "built variant $name ($flags)"
done
for f in $files; do touch "$pkg/csrc/$f"; done
bash "$pkg/build.sh" > /dev/null
echo "default build restored"
