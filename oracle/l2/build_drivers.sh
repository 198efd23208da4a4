#!/usr/bin/env bash
# oracle/l2/build_drivers.sh -- TEST INFRASTRUCTURE ONLY.  Links oracle/l2/l2_driver.cpp twice:
#   oracle/_ref/l2/l2_driver_ref : against the UNMODIFIED reference library (oracle/l2/build_duckdb.sh)
#   oracle/_ref/l2/l2_driver_gpu : against the reference with integration/duckdb/reference.patch applied to COPIES of
#       src/storage/table/column_segment.cpp and src/function/compression_config.cpp (in oracle/_ref/l2build_gpu/),
#       + integration/duckdb/succinct_gpu.cpp + libadacom_gpu.so -- the drop-in run of SURVEY 8(f) row 1.
# Only the two unity objects that contain the patched files are recompiled (with the commands cmake generated for
# them); they replace their members in a copy of libduckdb_static.a.  Nothing under /root/reference is written.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REPO="$(cd "$HERE/../.." && pwd)"
REFERENCE="${REFERENCE:-/root/reference}"
OUT="$REPO/oracle/_ref"
BUILD="$OUT/l2build"
GPU="$OUT/l2build_gpu"
BIN="$OUT/l2"
if [ ! -f "$REFERENCE/CMakeLists.txt" ]; then
  echo "oracle/l2: $REFERENCE not present -- keeping prebuilt oracle/_ref/l2 (if any)"; exit 0
fi
"$HERE/build_duckdb.sh" > /dev/null
mkdir -p "$GPU/src/storage/table" "$GPU/src/function" "$BIN"
CXX="${CXX:-g++}"
INC="-I$REFERENCE/src/include -I$REFERENCE/extension/tpch/include -I$REFERENCE/third_party/sdsl/include -I$REFERENCE/benchmark/micro/succinct"
LIBS_COMMON="$BUILD/extension/tpch/libtpch_extension.a $BUILD/extension/jemalloc/libjemalloc_extension.a \
  $BUILD/third_party/fsst/libduckdb_fsst.a $BUILD/third_party/fmt/libduckdb_fmt.a $BUILD/third_party/libpg_query/libduckdb_pg_query.a \
  $BUILD/third_party/re2/libduckdb_re2.a $BUILD/third_party/miniz/libduckdb_miniz.a $BUILD/third_party/utf8proc/libduckdb_utf8proc.a \
  $BUILD/third_party/hyperloglog/libduckdb_hyperloglog.a $BUILD/third_party/fastpforlib/libduckdb_fastpforlib.a \
  $BUILD/third_party/mbedtls/libduckdb_mbedtls.a $OUT/libsdsl.a"

# ---- the driver object (same for both arms) ----
$CXX -std=c++11 -O2 -DNDEBUG -w $INC -c "$HERE/l2_driver.cpp" -o "$GPU/l2_driver.o"

# ---- stock arm ----
$CXX -o "$BIN/l2_driver_ref" "$GPU/l2_driver.o" -Wl,--start-group "$BUILD/src/libduckdb_static.a" $LIBS_COMMON -Wl,--end-group \
  -lpthread -ldl -static-libstdc++ -static-libgcc

# ---- GPU arm: patched copies, recompiled unity objects, swapped archive members ----
cp "$REFERENCE/src/storage/table/column_segment.cpp" "$GPU/src/storage/table/column_segment.cpp"
cp "$REFERENCE/src/function/compression_config.cpp" "$GPU/src/function/compression_config.cpp"
chmod u+w "$GPU/src/storage/table/column_segment.cpp" "$GPU/src/function/compression_config.cpp"
patch -s -p1 -d "$GPU" < "$REPO/integration/duckdb/reference.patch"
python3 - "$BUILD" "$GPU" "$REFERENCE" "$REPO" <<'PY'
import json, os, shlex, subprocess, sys
build, gpu, ref, repo = sys.argv[1:5]
cc = json.load(open(os.path.join(build, "compile_commands.json")))
jobs = {"ub_duckdb_storage_table.cpp": ("src/storage/table/column_segment.cpp", "ub_duckdb_storage_table.cpp.o"),
        "ub_duckdb_function.cpp": ("src/function/compression_config.cpp", "ub_duckdb_function.cpp.o")}
for entry in cc:
    name = os.path.basename(entry["file"])
    if name not in jobs or "/src/function/ub_" not in entry["file"] and name == "ub_duckdb_function.cpp":
        continue
    rel, obj = jobs[name]
    unity = open(entry["file"]).read()
    want = f'#include <{ref}/{rel}>'
    if want not in unity:
        want = f'#include "{ref}/{rel}"'
    assert want in unity, (entry["file"], want)
    patched_unity = os.path.join(gpu, name)
    open(patched_unity, "w").write(unity.replace(want, f'#include "{gpu}/{rel}"'))
    args = shlex.split(entry["command"])
    o = args.index("-o")
    args[o + 1] = os.path.join(gpu, obj)
    args[args.index("-c") + 1] = patched_unity
    # the patched files sit outside their source directory: keep their relative includes working
    args[1:1] = [f"-I{repo}/integration/duckdb", f"-I{repo}/include", f"-I{ref}/{os.path.dirname(rel)}"]
    subprocess.check_call(args, cwd=entry["directory"])
    print("oracle/l2: recompiled", obj)
PY
FLAGS="$(python3 - "$BUILD" <<'PY'
import json, os, shlex, sys
cc = json.load(open(os.path.join(sys.argv[1], "compile_commands.json")))
e = next(c for c in cc if c["file"].endswith("ub_duckdb_storage_table.cpp"))
a = shlex.split(e["command"])
print(" ".join(x for x in a[1:] if x.startswith(("-D", "-I", "-O", "-std", "-f"))))
PY
)"
$CXX $FLAGS -I"$REPO/integration/duckdb" -I"$REPO/include" -c "$REPO/integration/duckdb/succinct_gpu.cpp" -o "$GPU/succinct_gpu.o"
cp "$BUILD/src/libduckdb_static.a" "$GPU/libduckdb_gpu.a"
ar r "$GPU/libduckdb_gpu.a" "$GPU/ub_duckdb_storage_table.cpp.o" "$GPU/ub_duckdb_function.cpp.o"
CSRC="$REPO/duckdb_adaptive_compression_b200/csrc"
$CXX -o "$BIN/l2_driver_gpu" "$GPU/l2_driver.o" "$GPU/succinct_gpu.o" -Wl,--start-group "$GPU/libduckdb_gpu.a" $LIBS_COMMON -Wl,--end-group \
  -L"$CSRC" -ladacom_gpu -Wl,-rpath,'$ORIGIN/../../../duckdb_adaptive_compression_b200/csrc' -lpthread -ldl -static-libstdc++ -static-libgcc
# ---- oracle for the neighbour codecs: the reference's bitpacking.cpp / rle.cpp + fastpforlib ----
$CXX $FLAGS -I"$REFERENCE" -shared -o "$OUT/libadacom_codecs_ref.so" "$HERE/codecs_ref.cpp" \
  -Wl,--allow-multiple-definition -Wl,--start-group "$BUILD/src/libduckdb_static.a" $LIBS_COMMON -Wl,--end-group \
  -lpthread -ldl -static-libstdc++ -static-libgcc
ls -la "$BIN" "$OUT/libadacom_codecs_ref.so"
