#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the reference's small_gdpt.cpp (unmodified, from where it lies) into
# oracle/_ref/smallgdpt_ref.  The program's own build line (small_gdpt.cpp:5) is `g++ -fopenmp -O3 ... -lfftw3`;
# FFTW is the stand-in already used for the lajolla reference.  -Wno-narrowing: the row seed `rng[3]={0,0,y*y*y}`
# (small_gdpt.cpp:462) narrows int to unsigned short in a braced initialiser.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${GDPT_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/../_ref"
if [ ! -f "$REF/small_gdpt.cpp" ]; then echo "build_smallgdpt: $REF/small_gdpt.cpp not found; skipping" >&2; exit 0; fi
mkdir -p "$OUT/obj_smallgdpt"
gcc -O3 -w -I"$REF/src/fftw-3.3.10/api" -c "$HERE/../ref_shims/fftw_standin.c" -o "$OUT/obj_smallgdpt/sg_fftw_standin.o"
gcc -O3 -w -c "$HERE/../dct1.c" -o "$OUT/obj_smallgdpt/sg_dct1.o"
g++ -O3 -fopenmp -w -Wno-narrowing -I"$REF" -I"$REF/src/fftw-3.3.10/api" "$HERE/smallgdpt_driver.cpp" \
    "$OUT/obj_smallgdpt/sg_fftw_standin.o" "$OUT/obj_smallgdpt/sg_dct1.o" -static-libstdc++ -static-libgcc -o "$OUT/smallgdpt_ref" -lm
echo "build_smallgdpt: ok -> $OUT/smallgdpt_ref"
