#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the *unmodified* reference (uberpixel/Rayne) hot-path
# sources where they lie under /root/reference into oracle/_ref/ (git-ignored).
# Nothing is copied into the repo: the config header is generated from the reference's own
# Source/RayneConfig.h.in with the values CMake/CompilerCheck.cmake:28-101 would choose,
# except RN_PLATFORM_LINUX 0 / RN_PLATFORM_POSIX 1 (avoids <xcb/xcb.h>, Source/Base/RNBase.h:53-55).
# The reference's own cmake build is NOT run (it needs XCB/X11 and Vendor/OIS, absent here).
#
#   oracle/build_ref.sh [strict|release|shim|all]
#     strict : -O2 -ffp-contract=off                      -> parity goldens (librayne_strict.a, ref_harness_strict)
#     release: the reference's flags (CMakeLists.txt:70)  -> timed CPU baseline (ref_harness_release)
set -euo pipefail
REF=${RN_REFERENCE_ROOT:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
MODE=${1:-all}
[ -d "$REF/Source" ] || { echo "reference not present at $REF - keeping prebuilt oracle/_ref"; exit 0; }
mkdir -p "$OUT/cfg"

# 1. config header generated from the reference template
sed -e 's/\${VERSION_MAJOR}/2/; s/\${VERSION_MINOR}/0/; s/\${VERSION_PATCH}/0/; s/\${VERSION_ABI}/10/' \
    -e 's/\${HAVE_STDINT_H}/1/; s/\${HAVE_STDDEF_H}/1/' \
    -e 's/\${RAYNE_HAS_VTUNE}/0/; s/\${RAYNE_ENABLE_VTUNE}/0/' \
    -e 's/\${RAYNE_FUNCTION_SIGNATURE}/__PRETTY_FUNCTION__/' \
    -e 's/\${RAYNE_EXPECT_TRUE}/__builtin_expect(!!(x), 1)/; s/\${RAYNE_EXPECT_FALSE}/__builtin_expect(!!(x), 0)/' \
    -e 's/\${RAYNE_ALIGNAS}/alignas(x)/; s/\${RAYNE_NOEXCEPT}/noexcept/g; s/\${RAYNE_CONSTEXPR}/constexpr/' \
    -e 's/\${RAYNE_NORETURN}/[[noreturn]]/; s/\${RAYNE_INLINE}/inline __attribute__((__always_inline__))/' \
    -e 's/\${RAYNE_NOINLINE}/__attribute__((noinline))/; s/\${RAYNE_UNUSED}/[[maybe_unused]]/' \
    -e 's/\${RAYNE_SUPPORTS_TRIVIALLY_COPYABLE}/1/; s/\${RAYNE_RNAPI_EXPORT}//; s/\${RAYNE_RNAPI_IMPORT}//' \
    -e 's/\${RAYNE_PLATFORM_OSX}/0/; s/\${RAYNE_PLATFORM_IOS}/0/; s/\${RAYNE_PLATFORM_VISIONOS}/0/' \
    -e 's/\${RAYNE_PLATFORM_WINDOWS}/0/; s/\${RAYNE_PLATFORM_LINUX}/0/; s/\${RAYNE_PLATFORM_POSIX}/1/; s/\${RAYNE_PLATFORM_ANDROID}/0/' \
    -e 's/\${RAYNE_INT8}/int8_t/; s/\${RAYNE_UINT8}/uint8_t/; s/\${RAYNE_INT16}/int16_t/; s/\${RAYNE_UINT16}/uint16_t/' \
    -e 's/\${RAYNE_INT32}/int32_t/; s/\${RAYNE_UINT32}/uint32_t/; s/\${RAYNE_INT64}/int64_t/; s/\${RAYNE_UINT64}/uint64_t/' \
    "$REF/Source/RayneConfig.h.in" > "$OUT/cfg/RayneConfig.h"
if grep -q '\${' "$OUT/cfg/RayneConfig.h"; then echo "unsubstituted config variable"; grep -n '\${' "$OUT/cfg/RayneConfig.h"; exit 1; fi

# 2. TU list = Source/CMakeLists.txt:19-121 minus the 7 that need generated vendor headers
#    (pnglibconf.h / jansson_config.h / zipconf.h) and minus Base/RNBase.cpp, which the harness
#    #includes so that the file-static RN::__functionPool can be created (RNBase.cpp:164,174).
SKIP='Assets/RNPNGAssetWriter.cpp|Assets/RNPNGAssetLoader.cpp|Assets/RNASTCAssetLoader.cpp|Assets/RNDDSAssetLoader.cpp|Rendering/RNTexture.cpp|Objects/RNJSONSerialization.cpp|System/RNFileManager.cpp|Base/RNBase.cpp'
TUS=$(sed -n '19,121p' "$REF/Source/CMakeLists.txt" | grep -oE '[A-Za-z/]+/RN[A-Za-z0-9]+\.cpp' | sort -u | grep -vE "$SKIP")

INC="-I$OUT/cfg -I$REF/Source -I$REF/Vendor -I$REF/Vendor/concurrentqueue -I$REF/Vendor/utf8_v2_3_4/source"
COMMON="-std=c++17 -DRN_BUILD_LIBRARY=1 -DNDEBUG -include cstring -w -fPIC -pthread $INC"
STRICT_FLAGS="-O2 -ffp-contract=off"
# reference release flags: CMakeLists.txt:70, Source/CMakeLists.txt:334-335
RELEASE_FLAGS="-O3 -fno-math-errno -funsafe-math-optimizations -march=core2 -mfpmath=sse"

build_one() {
    local name=$1 flags=$2
    local obj=$OUT/obj_$name
    mkdir -p "$obj"
    echo "$TUS" | xargs -P "$(nproc)" -I{} bash -c '
        src="$1"; o="$2/$(echo "$1" | tr / _ | sed s/.cpp$/.o/)"
        if [ ! -f "$o" ] || [ "$3/Source/$src" -nt "$o" ]; then
            g++ '"$COMMON $flags"' -c "$3/Source/$src" -o "$o" || { echo "FAILED $src"; exit 255; }
        fi' _ {} "$obj" "$REF"
    rm -f "$OUT/librayne_$name.a"
    ar rcs "$OUT/librayne_$name.a" "$obj"/*.o
    g++ $COMMON $flags -g -fno-access-control -DRN_REF_ROOT="\"$REF\"" \
        "$HERE/ref_harness.cpp" "$OUT/librayne_$name.a" \
        -Wl,--unresolved-symbols=ignore-all -ldl -o "$OUT/ref_harness_$name"
    echo "built $OUT/ref_harness_$name"
}
# RN::SceneCUDA (integration/RNSceneCUDA.cpp, the Rayne-side binding of librn_cuda) + its twin-scene test, linked against
# the strict build of the reference and against rayne_b200/librn_cuda.so (found at run time through $ORIGIN)
build_shim() {
    local root; root=$(cd "$HERE/.." && pwd)
    [ -f "$OUT/librayne_strict.a" ] || build_one strict "$STRICT_FLAGS"
    [ -f "$root/rayne_b200/librn_cuda.so" ] || { echo "build librn_cuda.so first (python -m rayne_b200.build)"; exit 1; }
    g++ $COMMON $STRICT_FLAGS -g -fno-access-control -I"$root/include" -I"$root/integration" \
        "$root/integration/shim_test.cpp" "$root/integration/RNSceneCUDA.cpp" "$OUT/librayne_strict.a" \
        -L"$root/rayne_b200" -lrn_cuda -Wl,-rpath,'$ORIGIN/../../rayne_b200' \
        -Wl,--unresolved-symbols=ignore-all -ldl -o "$OUT/shim_test"
    echo "built $OUT/shim_test"
}
case "$MODE" in
    shim)    build_shim ;;
    strict)  build_one strict "$STRICT_FLAGS" ;;
    release) build_one release "$RELEASE_FLAGS" ;;
    all)     build_one strict "$STRICT_FLAGS"; build_one release "$RELEASE_FLAGS"; build_shim ;;
esac
