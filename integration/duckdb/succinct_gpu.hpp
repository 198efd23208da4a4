//===----------------------------------------------------------------------===//
// succinct_gpu.hpp -- the reference-side binding of the B200 succinct codec.
//
// This translation unit is compiled AGAINST THE REFERENCE'S OWN HEADERS (DuckDB as forked by
// leonwind/duckdb-adaptive-compression: src/include + third_party/sdsl/include) and links libadacom_gpu.so.
// It provides, with the reference's exact slot signatures (src/include/duckdb/function/compression_function.hpp:
// 65-177), a CompressionFunction for CompressionType::COMPRESSION_SUCCINCT whose segments live on the GPU, plus the
// three bodies the policy side calls (ColumnSegment::BitCompressFromSuccinct / BitCompressFromUncompressed /
// UncompressSuccinct, src/storage/table/column_segment.cpp:348-506).
//
// Registration (one row, src/function/compression_config.cpp:20):
//     {CompressionType::COMPRESSION_SUCCINCT, SuccinctGpuFun::GetFunction, SuccinctGpuFun::TypeIsSupported},
//
// No reference header has to change: the device handle of a segment is kept in ColumnSegment::segment_state, the
// CompressedSegmentState that the ColumnSegment constructor obtains from the `init_segment` slot
// (column_segment.cpp:107-108) -- a slot SuccinctFun leaves empty.
//
// tools/check_duckdb_binding.sh syntax-checks this file against /root/reference where that tree exists.
//===----------------------------------------------------------------------===//
#pragma once

#include "duckdb/function/compression_function.hpp"

namespace duckdb {

struct SuccinctGpuFun {
	//! succinct.cpp:345-366 (same eight physical types, InternalException otherwise)
	static CompressionFunction GetFunction(PhysicalType type);
	//! succinct.cpp:368-382
	static bool TypeIsSupported(PhysicalType type);

	//! Bodies for the policy-side members of ColumnSegment (they are private members of the reference class, so the
	//! maintainer forwards to these from column_segment.cpp:348, :385 and :458).
	//! Returns the new ColumnSegment::GetDataSize() = 9 + 8 * ceil(count * width / 64).
	static idx_t Compact(ColumnSegment &segment, bool zero_extended_minmax);
	static idx_t Uncompact(ColumnSegment &segment);
	//! ColumnSegment::SuccinctSize() (column_segment.cpp:204-222) for a GPU-resident segment
	static idx_t DataSize(ColumnSegment &segment);
};

} // namespace duckdb
