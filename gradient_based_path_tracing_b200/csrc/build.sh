#!/usr/bin/env bash
# Builds libgdpt.so (the C-ABI library) for sm_100a, in tree.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../libgdpt.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr --extended-lambda -Xcompiler -fPIC -Xcompiler -O3 ${GDPT_NVCC_EXTRA:-}"
mkdir -p "$HERE/../_build"
pids=()
for f in gdpt_api.cu gdpt_api_validate.cu; do
  $NVCC $FLAGS -c "$HERE/$f" -o "$HERE/../_build/${f%.cu}.o" &
  pids+=($!)
done
g++ -O3 -std=c++17 -fPIC -c "$HERE/gdpt_bvh.cpp" -o "$HERE/../_build/gdpt_bvh.o" &
pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT" "$HERE/../_build/gdpt_api.o" "$HERE/../_build/gdpt_api_validate.o" "$HERE/../_build/gdpt_bvh.o" -lcudart
echo "built $OUT"
