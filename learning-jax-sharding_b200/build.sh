#!/usr/bin/env bash
# Builds libb200sl.so (the C-ABI library of include/b200_sharded.h) in-tree for sm_100a.
# nvcc cross-compiles without a GPU; the .so is git-ignored but travels with gpurun snapshots.
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
src="$here/csrc"
out="$here/libb200sl.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC
       --expt-relaxed-constexpr -Xptxas -v ${B200_EXTRA_NVCC_FLAGS:-})
# --clean: forget every object file first (what __graft_entry__.build() does, so that "does it build" really compiles)
if [[ "${1:-}" == "--clean" ]]; then rm -rf "$here/build"; fi
mkdir -p "$here/build"
objs=()
pids=()
for f in "$src"/*.cu; do
  o="$here/build/$(basename "${f%.cu}").o"
  objs+=("$o")
  if [[ ! -f "$o" || "$f" -nt "$o" || "$src/sm100_ptx.cuh" -nt "$o" || "$src/host_common.h" -nt "$o" \
        || "$here/../include/b200_sharded.h" -nt "$o" ]]; then
    "$NVCC" "${FLAGS[@]}" -c "$f" -o "$o" > "$o.log" 2>&1 &
    pids+=("$!:$o")
  fi
done
fail=0
for pe in "${pids[@]:-}"; do
  [[ -z "$pe" ]] && continue
  pid="${pe%%:*}"; o="${pe#*:}"
  if ! wait "$pid"; then echo "nvcc failed for $o:"; cat "$o.log"; fail=1; rm -f "$o"; fi
done
[[ $fail -eq 0 ]] || exit 1
"$NVCC" -shared -o "$out" "${objs[@]}" -gencode arch=compute_100a,code=sm_100a
echo "built $out"
