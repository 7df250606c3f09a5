#!/bin/bash
# Compiles the UNMODIFIED reference CUDA sources, from where they lie under
# /root/reference, plus our C-ABI wrapper oracle/ref_shim.cu, into
# oracle/_ref/libref_gauhuman.so.  Nothing is copied out of /root/reference.
# -include cstdint / cfloat: gcc-13 header hygiene (the reference uses
# std::uintptr_t and FLT_MAX without including them); no source edits.
# Flags mirror the reference setup.py (no -use_fast_math, default -fmad) with
# TORCH_CUDA_ARCH_LIST=10.0.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${GAUHUMAN_REFERENCE:-/root/reference}"
DGR="$REF/submodules/diff-gaussian-rasterization"
KNN="$REF/submodules/simple-knn"
if [ ! -d "$DGR/cuda_rasterizer" ]; then
  echo "build_ref: $DGR not present; keeping any prebuilt oracle/_ref" >&2
  exit 0
fi
mkdir -p "$HERE/_ref"
OUT="$HERE/_ref/libref_gauhuman.so"
SRCS=("$DGR/cuda_rasterizer/rasterizer_impl.cu" "$DGR/cuda_rasterizer/forward.cu" "$DGR/cuda_rasterizer/backward.cu" "$KNN/simple_knn.cu" "$HERE/ref_shim.cu")
newest=$(ls -t "${SRCS[@]}" "$0" | head -1)
if [ -f "$OUT" ] && [ "$OUT" -nt "$newest" ]; then exit 0; fi
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -shared -Xcompiler -fPIC -std=c++17 -O3 -w \
  -ccbin /usr/bin/g++ \
  -gencode arch=compute_100,code=sm_100 \
  -include cstdint -include cfloat \
  -I"$DGR/third_party/glm" -I"$DGR/cuda_rasterizer" -I"$KNN" \
  "${SRCS[@]}" -o "$OUT"
echo "built $OUT"
