#!/usr/bin/env bash
# Compiles csrc/xla_ffi.cu against tools/ffi_stub (a stand-in for jaxlib's xla/ffi/api/ffi.h, see its header) so that the
# XLA typed-FFI handlers go through a compiler at least once; the object file is discarded.  Not a jaxlib verification.
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
out="$(mktemp /tmp/b200_ffi_check.XXXXXX.o)"
"$NVCC" -gencode arch=compute_100a,code=sm_100a -std=c++17 -I"$here/tools/ffi_stub" -c \
  "$here/learning-jax-sharding_b200/csrc/xla_ffi.cu" -o "$out"
n=$(nm "$out" | grep -c " T B200" || true)
rm -f "$out"
echo "xla_ffi.cu compiles against the FFI stand-in: $n handler symbols"
[[ "$n" -ge 9 ]]
