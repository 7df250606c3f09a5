#!/bin/bash
# tools/build_variant.sh TAG "-DGHR_FWD_STAGES=8 ..."  ->  variants/libghr_TAG.so  (tuning only; use with GHR_LIB_PATH)
set -euo pipefail
cd "$(dirname "$0")/.."
mkdir -p variants
GHR_FORCE_BUILD=1 GHR_OUT="$PWD/variants/libghr_$1.so" GHR_NVCC_EXTRA="$2" bash mygauhuman_b200/csrc/build.sh
