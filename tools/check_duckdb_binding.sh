#!/bin/bash
# Syntax-checks integration/duckdb/succinct_gpu.cpp against the reference's own headers (DuckDB fork + vendored SDSL).
# Only meaningful where the reference tree is mounted (this container); prints SKIP elsewhere.
REF=${ADACOM_REFERENCE:-/root/reference}
HERE=$(cd "$(dirname "$0")/.." && pwd)
if [ ! -d "$REF/src/include/duckdb" ]; then
	echo "SKIP: no reference tree at $REF"
	exit 0
fi
g++ -std=c++11 -fsyntax-only -Wall -Wno-unused-parameter -isystem "$REF/src/include" -isystem "$REF/third_party/sdsl/include" \
	-I"$HERE/include" -I"$HERE/integration/duckdb" "$HERE/integration/duckdb/succinct_gpu.cpp" && echo "OK: binding compiles against the reference headers"
