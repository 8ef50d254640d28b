#!/usr/bin/env bash
# Test infrastructure (NOT product code).
# Builds the reference's OWN OpenPGL 0.7.0 (VSP branch) query library, from the
# sources where they lie under /root/reference, into oracle/_ref/ (git-ignored).
# Nothing is copied out of /root/reference; only object files / .so land here.
#   oracle/_ref/libtbb_ref.so      vendored oneTBB  (src/ext/tbb/src/tbb/*.cpp)
#   oracle/_ref/libopenpgl_ref.so  OpenPGL api + CPU_4 device (src/ext/openpgl/openpgl/CMakeLists.txt:16-31)
#   oracle/_ref/libpglharness.so   our C-callable probe harness (oracle/ref_harness/pgl_harness.cpp)
# Build options mirror src/ext/CMakeLists.txt:262-273 (VSP guiding ON, ISGB ON,
# radiance caches OFF, compression OFF) and the integrator's PGL_DEVICE_TYPE_CPU_4
# (cpu/guidedvolpathvspgintegrator.cpp:111).  OIDN is replaced by the identity stub
# in oracle/stubs (needs ISPC, absent); the cmake-generated defines.h/version.h are
# hand-written in oracle/stubs/openpgl.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${VSPG_REFERENCE_ROOT:-/root/reference}"
if [ ! -d "$REF/src/ext/openpgl" ]; then
  echo "build_ref: $REF not present; keeping prebuilt oracle/_ref (if any)"; exit 0
fi
EXT="$REF/src/ext"; PGL="$EXT/openpgl"; TBB="$EXT/tbb"
OUT="$HERE/_ref"; OBJ="$OUT/obj"; mkdir -p "$OBJ/tbb" "$OBJ/pgl"
JOBS="${JOBS:-$(nproc)}"
# /usr/bin/g++ links libstdc++ dynamically; the image's /opt/gcc wrapper links it statically, and one private copy of
# libstdc++ per .so breaks iostream locale facets across the three libraries (seen as a crash in std::cout << bool).
if [ -x /usr/bin/g++ ]; then CXX="${VSPG_REF_CXX:-/usr/bin/g++}"; else CXX="${VSPG_REF_CXX:-g++}"; fi
# ---- oneTBB ------------------------------------------------------------------
if [ ! -f "$OUT/libtbb_ref.so" ]; then
  ls "$TBB"/src/tbb/*.cpp | xargs -P "$JOBS" -I{} sh -c \
    "$CXX -std=c++17 -O2 -fPIC -mwaitpkg -mrtm -w -D__TBB_BUILD -D__TBB_USE_ITT_NOTIFY=0 -I$TBB/include -I$TBB/src -c {} -o $OBJ/tbb/\$(basename {} .cpp).o"
  $CXX -shared -o "$OUT/libtbb_ref.so" "$OBJ"/tbb/*.o -lpthread -ldl
fi
# ---- OpenPGL -----------------------------------------------------------------
DEF="-include $HERE/stubs/openpgl_build_defs.h"
INC="-I$HERE/stubs -I$HERE/stubs/openpgl -I$PGL -I$PGL/openpgl -I$PGL/openpgl/include -I$PGL/openpgl/include/openpgl -I$PGL/third-party -I$PGL/third-party/embreeSrc -I$TBB/include"
# -ffp-contract=off: the reference builds for plain SSE4.2 (no FMA); keep gcc from fusing.
FLAGS="-std=c++17 -O2 -fPIC -msse4.2 -ffp-contract=off -w"
SRCS="openpgl/api/api.cpp openpgl/api/deviceCPU4.cpp openpgl/directional/ISurfaceSamplingDistribution.cpp openpgl/directional/IVolumeSamplingDistribution.cpp openpgl/directional/vmm/VMMPhaseFunctions.cpp third-party/embreeSrc/common/simd/sse.cpp third-party/embreeSrc/common/sys/alloc.cpp third-party/embreeSrc/common/sys/barrier.cpp third-party/embreeSrc/common/sys/condition.cpp third-party/embreeSrc/common/sys/mutex.cpp third-party/embreeSrc/common/sys/estring.cpp third-party/embreeSrc/common/sys/sysinfo.cpp third-party/embreeSrc/common/sys/thread.cpp"
if [ ! -f "$OUT/libopenpgl_ref.so" ]; then
  for s in $SRCS; do echo "$s"; done | xargs -P "$JOBS" -I{} sh -c \
    "$CXX $FLAGS $DEF $INC -c $PGL/{} -o $OBJ/pgl/\$(echo {} | tr '/' '_').o"
  $CXX -shared -o "$OUT/libopenpgl_ref.so" "$OBJ"/pgl/*.o -L"$OUT" -ltbb_ref -lpthread -ldl -Wl,-rpath,'$ORIGIN'
fi
# ---- harness -------------------------------------------------------------------
if [ -f "$HERE/ref_harness/pgl_harness.cpp" ]; then
  $CXX $FLAGS -DOPENPGL_VSP_GUIDING -DOPENPGL_IMAGE_SPACE_GUIDING_BUFFER $INC -shared -o "$OUT/libpglharness.so" "$HERE/ref_harness/pgl_harness.cpp" "$HERE/ref_harness/pgl_hg_table.cpp" \
     -L"$OUT" -lopenpgl_ref -ltbb_ref -lpthread -ldl -Wl,-rpath,'$ORIGIN'
fi
echo "build_ref: ok -> $OUT"
