#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the REAL reference (lajolla + the GDPT add-on) from the
# sources where they lie under /root/reference into oracle/_ref/ (git-ignored):
#   oracle/_ref/liblajolla_ref.so   reference + stand-ins + C hooks (ctypes)
#   oracle/_ref/lajolla             the reference CLI (src/main.cpp), for "./lajolla scene.xml" timing
# Nothing from /root/reference is copied into the repository.  Two stand-ins replace the
# absent third-party binaries (Embree 4.3.0, FFTW 3.3.10): ref_shims/rtc_standin.cpp and
# ref_shims/fftw_standin.c (+ dct1.c).  One file is patched on the fly (sed, output only in
# oracle/_ref/gen/): src/render.cpp —
#   (1) `int spp = 1000` / `int spp = 256`  ->  scene.options.samples_per_pixel
#       (the hard-coded counts at render.cpp:104,293; the XML sampleCount is otherwise ignored)
#   (2) a capture call after fourierSolve (render.cpp:353) so c, cx, cy, out can be read back.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${GDPT_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then echo "build_ref: $REF/src not found; skipping" >&2; exit 0; fi
mkdir -p "$OUT/obj" "$OUT/gen"
CXX="${CXX:-g++}"; CC="${CC:-gcc}"
CXXFLAGS="-std=c++17 -O3 -g0 -fPIC -w -pthread -I$REF/src -I$REF/old_embree/embree/include -I$REF/src/fftw-3.3.10/api"
CFLAGS="-O3 -fPIC -w -I$REF/src/fftw-3.3.10/api"

sed -e 's|int spp = 1000 *;//scene.options.samples_per_pixel;|int spp = scene.options.samples_per_pixel;|' \
    -e 's|int spp = 256;//scene.options.samples_per_pixel;|int spp = scene.options.samples_per_pixel;|' \
    -e 's|^\(\s*\)fourierSolve(w, h, (double\*)c, (double\*)cx, (double\*)cy, 0.04, (double\*)out);|&\n\1gdpt_ref_dump_hook(w, h, (double*)c, (double*)cx, (double*)cy, (double*)out);|' \
    "$REF/src/render.cpp" > "$OUT/gen/render_patched.cpp"
grep -q "gdpt_ref_dump_hook" "$OUT/gen/render_patched.cpp" || { echo "build_ref: dump-hook splice failed" >&2; exit 1; }
if grep -q "int spp = 1000" "$OUT/gen/render_patched.cpp"; then echo "build_ref: spp patch failed" >&2; exit 1; fi
echo 'extern "C" void gdpt_ref_dump_hook(int, int, const double*, const double*, const double*, const double*);' > "$OUT/gen/hook_decl.h"

pids=()
compile() { # src obj extra-flags
  if [ ! -f "$2" ] || [ "$1" -nt "$2" ]; then
    case "$1" in
      *.c) $CC $CFLAGS $3 -c "$1" -o "$2" & ;;
      *) $CXX $CXXFLAGS $3 -c "$1" -o "$2" & ;;
    esac
    pids+=($!)
    if [ ${#pids[@]} -ge 8 ]; then wait "${pids[0]}"; pids=("${pids[@]:1}"); fi
  fi
}
OBJS=()
for f in "$REF"/src/*.cpp "$REF"/src/parsers/*.cpp "$REF"/src/3rdparty/pugixml.cpp "$REF"/src/3rdparty/miniz.c; do
  b="$(basename "$f")"
  [ "$b" = "render.cpp" ] && continue
  [ "$b" = "main.cpp" ] && continue
  o="$OUT/obj/${b%.*}.o"; OBJS+=("$o"); compile "$f" "$o" ""
done
o="$OUT/obj/render_patched.o"; OBJS+=("$o")
compile "$OUT/gen/render_patched.cpp" "$o" "-include $OUT/gen/hook_decl.h -I$REF/src"
for f in "$HERE/ref_shims/rtc_standin.cpp" "$HERE/ref_shims/ref_hooks.cpp" "$HERE/ref_shims/fftw_standin.c" "$HERE/dct1.c" "$HERE/pcg_oracle.c"; do
  b="$(basename "$f")"; o="$OUT/obj/shim_${b%.*}.o"; OBJS+=("$o"); compile "$f" "$o" ""
done
compile "$REF/src/main.cpp" "$OUT/obj/main.o" ""
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done

$CXX -shared -pthread -static-libstdc++ -static-libgcc -Wl,-Bsymbolic -Wl,--exclude-libs,ALL -o "$OUT/liblajolla_ref.so" "${OBJS[@]}"
$CXX -pthread -o "$OUT/lajolla" "$OUT/obj/main.o" "${OBJS[@]}"
echo "build_ref: ok -> $OUT/liblajolla_ref.so, $OUT/lajolla"
