#!/bin/bash
# Builds libghr.so (all CUDA kernels + the C ABI) in-tree for sm_100a.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
OUT="${GHR_OUT:-$HERE/libghr.so}"   # GHR_OUT + GHR_NVCC_EXTRA=-D...: tuning variants (tools/), never shipped
OBJ="$HERE/_obj${GHR_OUT:+_$(basename "$GHR_OUT" .so)}"
SRCS=("$HERE"/ghr_*.cu)
newest=$(ls -t "$HERE"/ghr_*.cu "$HERE"/*.cuh "$HERE"/*.h "$HERE/../../include/ghr.h" "$0" | head -1)
if [ -f "$OUT" ] && [ "$OUT" -nt "$newest" ] && [ -z "${GHR_FORCE_BUILD:-}" ]; then exit 0; fi
mkdir -p "$OBJ"
pids=()
for src in "${SRCS[@]}"; do
  obj="$OBJ/$(basename "${src%.cu}").o"
  "$NVCC" -c -O3 -std=c++17 -lineinfo -ccbin /usr/bin/g++ -Xcompiler -fPIC \
    -gencode arch=compute_100a,code=sm_100a ${GHR_NVCC_EXTRA:-} "$src" -o "$obj" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
"$NVCC" -shared -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a "$OBJ"/ghr_*.o -o "$OUT"
echo "built $OUT"
