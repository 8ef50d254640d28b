#!/usr/bin/env bash
# Test infrastructure (NOT product code).
# Builds a TRIMMED CPU build of the reference renderer itself -- libpbrt_ref.so + the `pbrt` command -- from the
# sources where they lie under /root/reference, into oracle/_ref/ (git-ignored).  Nothing is copied out of
# /root/reference; only generated files (soac output, rgb2spec tables), objects, .so files and binaries land here.
# The reference's own cmake build is not runnable on this box (GLFW/X11, OIDN/ISPC, Imath fetched from the network,
# two source blobs missing: SURVEY.md 8c), so this script lists the translation units itself:
#   kept     src/pbrt/*.cpp, src/pbrt/cpu/*.cpp, src/pbrt/util/*.cpp (no *_test.cpp), cmd/pbrt.cpp, and the vendored
#            double-conversion, utf8proc, libdeflate, ptex, lodepng, rply, stb, qoi, filesystem, nanovdb headers
#   dropped  wavefront/, gpu/ (the CUDA/OptiX path), util/gui.cpp (GLFW/OpenGL: no graphics stack here)
#   stubbed  <pbrt/util/gui.h>        -> oracle/stubs/pbrt_shadow (GUI::Initialize / GetResolution only)
#            OpenImageDenoise/oidn.hpp -> oracle/stubs (identity filter; never executed on the load-from-file path)
#            OpenEXR (Imf*)           -> oracle/stubs/exr (own minimal scanline reader/writer; Imath is not vendored)
#            util/bluenoise.cpp, util/pmj02tables.cpp (missing blobs) -> zero tables (PNG dithering / pmj02bn sampler only)
#            RenderWavefront          -> aborts
# Build options mirror the reference's CMakeLists.txt defaults: PBRT_WITH_PATH_GUIDING, PBRT_WITH_OIDN,
# PBRT_RGB_RENDERING, PTEX_STATIC; floating point as for the OpenPGL build (build_ref.sh): -msse4.2 and
# -ffp-contract=off, i.e. PBRT_BUILD_NATIVE_EXECUTABLE=OFF (no FMA contraction, so results do not depend on the
# build machine).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${VSPG_REFERENCE_ROOT:-/root/reference}"
if [ ! -d "$REF/src/pbrt" ]; then
  echo "build_ref_pbrt: $REF not present; keeping prebuilt oracle/_ref (if any)"; exit 0
fi
S="$REF/src"; EXT="$S/ext"; PGL="$EXT/openpgl"
OUT="$HERE/_ref"; OBJ="$OUT/obj"; GEN="$OUT/gen"
mkdir -p "$OBJ/pbrt" "$OBJ/ext" "$GEN"
JOBS="${JOBS:-$(nproc)}"
if [ -x /usr/bin/g++ ]; then CXX="${VSPG_REF_CXX:-/usr/bin/g++}"; CC="${VSPG_REF_CC:-/usr/bin/gcc}"; else CXX="${VSPG_REF_CXX:-g++}"; CC="${VSPG_REF_CC:-gcc}"; fi
[ -f "$OUT/libopenpgl_ref.so" ] || bash "$HERE/build_ref.sh"

DEFS="-DPBRT_IS_LINUX -DPBRT_RGB_RENDERING -DPBRT_WITH_PATH_GUIDING -DPBRT_WITH_OIDN -DPTEX_STATIC -DNDEBUG -DPBRT_HAVE_MMAP -DPBRT_HAVE_POSIX_MEMALIGN -DNANOVDB_USE_ZIP -DPBRT_NOINLINE=__attribute__((noinline)) -DPBRT_RESTRICT=__restrict__"
INC="-I$HERE/stubs/pbrt_shadow -I$HERE/stubs -I$HERE/stubs/exr -I$GEN -I$S -I$EXT -I$EXT/stb -I$EXT/qoi -I$EXT/libdeflate \
-I$EXT/filesystem -I$EXT/ptex/src/ptex -I$EXT/double-conversion -I$EXT/openvdb/nanovdb -I$EXT/utf8proc \
-I$HERE/stubs/ptex -I$HERE/stubs/openpgl -I$PGL/openpgl/include -I$PGL/openpgl/include/openpgl -I$EXT/tbb/include"
FLAGS="-std=c++17 -O2 -fPIC -msse4.2 -ffp-contract=off -w"

# ---- generated sources --------------------------------------------------------
if [ ! -f "$GEN/pbrt_soa.h" ]; then
  $CXX -std=c++17 -O1 -w -I$S -I$EXT $DEFS "$S/pbrt/cmd/soac.cpp" -o "$GEN/soac"
  "$GEN/soac" "$S/pbrt/pbrt.soa" > "$GEN/pbrt_soa.h"
fi
if [ ! -f "$GEN/rgbspectrum_aces.cpp" ]; then
  $CXX -std=c++17 -O2 -w -I$S -I$EXT $DEFS "$S/pbrt/cmd/rgb2spec_opt.cpp" -o "$GEN/rgb2spec_opt" -lpthread
  "$GEN/rgb2spec_opt" 64 "$GEN/rgbspectrum_srgb.cpp" sRGB
  "$GEN/rgb2spec_opt" 64 "$GEN/rgbspectrum_dci_p3.cpp" DCI_P3
  "$GEN/rgb2spec_opt" 64 "$GEN/rgbspectrum_rec2020.cpp" REC2020
  "$GEN/rgb2spec_opt" 64 "$GEN/rgbspectrum_aces.cpp" ACES2065_1
fi

# ---- translation units ----------------------------------------------------------
LIST="$OBJ/pbrt_units.txt"; : > "$LIST"
for f in "$S"/pbrt/*.cpp "$S"/pbrt/cpu/*.cpp "$S"/pbrt/util/*.cpp; do
  case "$f" in *_test.cpp|*/util/gui.cpp) continue;; esac
  echo "$f" >> "$LIST"
done
for f in "$GEN"/rgbspectrum_*.cpp "$HERE/ref_harness/pbrt_missing_tables.cpp" "$EXT/lodepng/lodepng.cpp" "$EXT/rply/rply.cpp" "$EXT/gtest/gtest-all.cc" \
         "$EXT"/double-conversion/double-conversion/*.cc "$EXT"/ptex/src/ptex/*.cpp; do
  echo "$f" >> "$LIST"
done
compile_one() {
  src="$1"; o="$OBJ/pbrt/$(echo "${src#$REF/}" | tr '/.' '__').o"
  if [ ! -f "$o" ] || [ "$src" -nt "$o" ]; then
    $CXX $FLAGS $DEFS $INC -c "$src" -o "$o" 2> "$o.log" || { echo "FAILED $src"; head -30 "$o.log"; rm -f "$o"; return 1; }
  fi
}
export -f compile_one; export CXX FLAGS DEFS INC OBJ REF
xargs -a "$LIST" -P "$JOBS" -I{} bash -c 'compile_one "$@"' _ {}
# C sources: utf8proc, libdeflate
for f in "$EXT/utf8proc/utf8proc.c" "$EXT"/libdeflate/lib/*.c "$EXT"/libdeflate/lib/x86/*.c; do
  o="$OBJ/ext/$(basename "$f" .c).o"
  [ -f "$o" ] || $CC -O2 -fPIC -w -I"$EXT/libdeflate" -I"$EXT/utf8proc" -DUTF8PROC_STATIC -c "$f" -o "$o"
done

# ---- link -------------------------------------------------------------------------
$CXX -shared -o "$OUT/libpbrt_ref.so" "$OBJ"/pbrt/*.o "$OBJ"/ext/*.o -L"$OUT" -lopenpgl_ref -ltbb_ref -lz -lpthread -ldl -Wl,-rpath,'$ORIGIN' -Wl,--no-undefined
$CXX $FLAGS $DEFS $INC -o "$OUT/pbrt" "$S/pbrt/cmd/pbrt.cpp" -L"$OUT" -lpbrt_ref -lopenpgl_ref -ltbb_ref -lpthread -ldl -Wl,-rpath,'$ORIGIN'
if [ -f "$HERE/ref_harness/pbrt_harness.cpp" ]; then
  $CXX $FLAGS $DEFS $INC -shared -o "$OUT/libpbrtharness.so" "$HERE/ref_harness/pbrt_harness.cpp" -L"$OUT" -lpbrt_ref -lopenpgl_ref -ltbb_ref -lpthread -ldl -Wl,-rpath,'$ORIGIN'
fi
# .nvdb files for the NanoVDB medium tests come from the reference's vendored NanoVDB itself (header-only)
if [ -f "$HERE/ref_harness/nvdb_writer.cpp" ]; then
  $CXX -std=c++17 -O2 -w -DNANOVDB_USE_ZIP -I"$EXT/openvdb/nanovdb" -o "$OUT/nvdb_writer" "$HERE/ref_harness/nvdb_writer.cpp" -lz -lpthread
fi
echo "build_ref_pbrt: ok -> $OUT"
