#!/usr/bin/env bash
# Builds libgdpt.so (the C-ABI library) for sm_100a, in tree.  The two arithmetic modes and
# the validation launchers are separate translation units so they compile in parallel.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../libgdpt.so"
B="$HERE/../_build"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr --extended-lambda -Xcompiler -fPIC -Xcompiler -O3 ${GDPT_NVCC_EXTRA:-}"
mkdir -p "$B"
SRCS="gdpt_api.cu gdpt_api_validate.cu gdpt_wave_f64.cu gdpt_wave_f32.cu gdpt_path_f64.cu gdpt_path_f32.cu gdpt_validate_f64.cu gdpt_validate_f32.cu gdpt_poisson.cu gdpt_poisson_l1.cu gdpt_poisson_noalloc.cu gdpt_multi.cu gdpt_bvh_gpu.cu gdpt_textures.cu gdpt_shift_f64.cu gdpt_shift_f32.cu gdpt_smallgdpt_f64.cu"
pids=(); OBJS=()
newest_hdr=$(ls -t "$HERE"/*.cuh "$HERE"/*.h "$HERE/../../include/gdpt.h" | head -1)
for f in $SRCS; do
  o="$B/${f%.cu}.o"; OBJS+=("$o")
  if [ ! -f "$o" ] || [ "$HERE/$f" -nt "$o" ] || [ "$newest_hdr" -nt "$o" ] || [ -n "${GDPT_FORCE:-}" ]; then
    # exact mode mirrors the reference's double arithmetic (gcc, x86-64, no FMA contraction):
    # keep multiply and add separate there so ties and branch decisions fall the same way
    EXTRA=""; case "$f" in *_f64.cu) EXTRA="-fmad=false";; esac
    $NVCC $FLAGS $EXTRA -c "$HERE/$f" -o "$o" &
    pids+=($!)
  fi
done
o="$B/gdpt_bvh.o"; OBJS+=("$o")
if [ ! -f "$o" ] || [ "$HERE/gdpt_bvh.cpp" -nt "$o" ] || [ "$HERE/gdpt_bvh.h" -nt "$o" ]; then
  g++ -O3 -std=c++17 -fPIC -c "$HERE/gdpt_bvh.cpp" -o "$o" &
  pids+=($!)
fi
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT" "${OBJS[@]}" -lcudart
echo "built $OUT"
# the C++ host over the C ABI: lajolla_b200 (container reader, flattening, PFM writers)
CLI="$HERE/../lajolla_b200"
if [ ! -f "$CLI" ] || [ "$OUT" -nt "$CLI" ] || [ -n "$(find "$HERE/host" -newer "$CLI" -type f | head -1)" ]; then
  g++ -O2 -std=c++17 -I "$HERE/../../include" "$HERE/host/lajolla_b200.cpp" "$HERE/host/gdpt_container.cpp" -o "$CLI" \
      -L "$HERE/.." -lgdpt -lz -Wl,-rpath,'$ORIGIN'
  echo "built $CLI"
fi
# the reference's second program as a host over the C ABI (same command line and output files as small_gdpt.cpp)
SG="$HERE/../smallgdpt_b200"
if [ ! -f "$SG" ] || [ "$OUT" -nt "$SG" ] || [ "$HERE/host/smallgdpt_b200.cpp" -nt "$SG" ]; then
  g++ -O2 -std=c++17 -I "$HERE/../../include" "$HERE/host/smallgdpt_b200.cpp" -o "$SG" -L "$HERE/.." -lgdpt -Wl,-rpath,'$ORIGIN'
  echo "built $SG"
fi
