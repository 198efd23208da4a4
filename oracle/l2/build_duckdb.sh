#!/usr/bin/env bash
# oracle/l2/build_duckdb.sh -- TEST INFRASTRUCTURE ONLY (SURVEY 8(c) "L2": the whole reference library).
#
# Builds the UNMODIFIED reference (DuckDB fork + AdaCom) out of tree from the sources where they lie under
# $REFERENCE (read-only) into oracle/_ref/l2build/ (git-ignored; the intermediate objects are gpurun-ignored, the
# drivers linked from them under oracle/_ref/l2/ travel to the GPU box).  Recipe = SURVEY.md 8(c):
#   * SDSL: the reference ships no prebuilt third_party/sdsl/lib/*.a, so libsdsl.a is made from the objects
#     oracle/Makefile compiles out of third_party/sdsl-lite/lib and handed to the link lines;
#   * targets duckdb_static + tpch_extension + jemalloc_extension (BUILD_UNITTESTS=0: the unit-test and
#     benchmark_runner targets need third_party/sqlite/sqlite3.c, a stripped blob in this tree).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REFERENCE="${REFERENCE:-/root/reference}"
OUT="$HERE/../_ref"
BUILD="$OUT/l2build"
if [ ! -f "$REFERENCE/CMakeLists.txt" ]; then
  echo "oracle/l2: $REFERENCE not present -- keeping prebuilt oracle/_ref/l2 (if any)"; exit 0
fi
make -C "$HERE/.." ref >/dev/null
mkdir -p "$BUILD"
ar rcs "$OUT/libsdsl.a" "$OUT"/sdsl_*.o
if [ ! -f "$BUILD/build.ninja" ]; then
  cmake -S "$REFERENCE" -B "$BUILD" -GNinja -DCMAKE_BUILD_TYPE=Release \
    -DBUILD_TPCH_EXTENSION=1 -DBUILD_UNITTESTS=0 -DBUILD_SHELL=0 -DCMAKE_POLICY_VERSION_MINIMUM=3.5 \
    -DCMAKE_EXPORT_COMPILE_COMMANDS=1 \
    -DCMAKE_EXE_LINKER_FLAGS="$OUT/libsdsl.a" -DCMAKE_SHARED_LINKER_FLAGS="$OUT/libsdsl.a" > "$BUILD/configure.log" 2>&1
fi
ninja -C "$BUILD" -j "${L2_JOBS:-8}" duckdb_static tpch_extension jemalloc_extension
echo "oracle/l2: libduckdb_static.a ready in $BUILD/src"
