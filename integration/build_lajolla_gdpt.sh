#!/usr/bin/env bash
# Builds the reference's OWN command-line program with the GPU binding in its `gradpath` arm:
#   oracle/_ref/lajolla_gdpt = /root/reference/src/main.cpp + the reference's scene parser, loaders and every other
#   translation unit (the objects oracle/build_ref.sh already compiled from the sources where they lie)
#   + integration/render_b200.cpp (the file a maintainer adds) + render.cpp with the ONE call-site edit of
#   INTEGRATION.md §1 applied on the fly (sed, output only under oracle/_ref/gen/) + libgdpt.so.
# `oracle/_ref/lajolla_gdpt [-t N] [-o out] scene.xml` then behaves like `./lajolla scene.xml` (src/main.cpp:11-51) with
# <integrator type="gradpath"> rendered on the GPU.  The other integrators keep the reference's CPU code (on the Embree /
# FFTW stand-ins of oracle/ref_shims, since those libraries are absent here).
# Test / integration infrastructure: nothing from /root/reference is copied into the repository; outputs are git-ignored.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$HERE/.."
REF="${GDPT_REFERENCE_DIR:-/root/reference}"
OUT="$ROOT/oracle/_ref"
if [ ! -d "$REF/src" ]; then echo "build_lajolla_gdpt: $REF/src not found; skipping" >&2; exit 0; fi
[ -f "$OUT/obj/main.o" ] || bash "$ROOT/oracle/build_ref.sh" >/dev/null
CXX="${CXX:-g++}"
CXXFLAGS="-std=c++17 -O3 -g0 -fPIC -w -pthread -I$REF/src -I$REF/old_embree/embree/include -I$REF/src/fftw-3.3.10/api -I$ROOT/include"
mkdir -p "$OUT/gen" "$OUT/obj"
# the GradPath arm of render() calls the binding; the spp edits are the ones build_ref.sh makes for the
# CPU integrators that stay (path_render reads the XML sample count instead of its hard-coded 256)
sed -e 's|int spp = 1000 *;//scene.options.samples_per_pixel;|int spp = scene.options.samples_per_pixel;|' \
    -e 's|int spp = 256;//scene.options.samples_per_pixel;|int spp = scene.options.samples_per_pixel;|' \
    -e 's|return gradient_path_render(scene);|return gradient_path_render_b200(scene);|' \
    "$REF/src/render.cpp" > "$OUT/gen/render_b200_patched.cpp"
grep -q "return gradient_path_render_b200(scene);" "$OUT/gen/render_b200_patched.cpp" || { echo "call-site edit failed" >&2; exit 1; }
echo '#include "image.h"
struct Scene; Image3 gradient_path_render_b200(const Scene &scene);' > "$OUT/gen/b200_decl.h"
$CXX $CXXFLAGS -include "$OUT/gen/b200_decl.h" -c "$OUT/gen/render_b200_patched.cpp" -o "$OUT/obj/render_b200_patched.o" &
$CXX $CXXFLAGS -c "$HERE/render_b200.cpp" -o "$OUT/obj/render_b200.o" &
wait
OBJS=()
for o in "$OUT"/obj/*.o; do
  case "$(basename "$o")" in
    render_patched.o|render_b200_patched.o|render_b200.o|main.o|shim_ref_hooks.o) ;;   # ref_hooks is the ctypes oracle, not part of the CLI
    *) OBJS+=("$o") ;;
  esac
done
$CXX -pthread -o "$OUT/lajolla_gdpt" "$OUT/obj/main.o" "$OUT/obj/render_b200_patched.o" "$OUT/obj/render_b200.o" "${OBJS[@]}" \
    -L"$ROOT/gradient_based_path_tracing_b200" -lgdpt -Wl,-rpath,'$ORIGIN/../../gradient_based_path_tracing_b200'
echo "build_lajolla_gdpt: ok -> $OUT/lajolla_gdpt"
