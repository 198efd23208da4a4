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
// (column_segment.cpp:107-108) -- a slot SuccinctFun leaves empty.  The edits to the two reference SOURCE files
// (registry row + the forwards inside ColumnSegment) are integration/duckdb/reference.patch; oracle/l2/build_drivers.sh
// applies it to a copy, links this file + libadacom_gpu.so into the reference and runs the reference's own benchmark
// bodies through it (oracle/l2/l2_driver.cpp).
//===----------------------------------------------------------------------===//
#pragma once

#include "duckdb/function/compression_function.hpp"

namespace duckdb {

struct SuccinctGpuFun {
	//! succinct.cpp:345-366 (same eight physical types, InternalException otherwise)
	static CompressionFunction GetFunction(PhysicalType type);
	//! succinct.cpp:368-382
	static bool TypeIsSupported(PhysicalType type);

	//! Bodies for the policy-side members of ColumnSegment (private members of the reference class: reference.patch
	//! forwards to these from column_segment.cpp:348, :385 and :458).
	//! BitCompressFromSuccinct / BitCompressFromUncompressed: analyze + encode on the GPU; returns the new
	//! ColumnSegment::GetDataSize() = 9 + 8 * ceil(count * width / 64).
	static idx_t Compact(ColumnSegment &segment, bool zero_extended_minmax);
	//! adaptive path (column_segment.cpp:385-456): the segment was UNCOMPRESSED in a buffer-managed block; its values
	//! move to the device and the returned state becomes ColumnSegment::segment_state
	static unique_ptr<CompressedSegmentState> InitSegmentFromBlock(ColumnSegment &segment, data_ptr_t block_ptr);
	//! UncompressSuccinct (column_segment.cpp:458-506): decode into the freshly allocated flat block
	static void UncompactToBlock(ColumnSegment &segment, data_ptr_t block_ptr);
	//! ColumnSegment::GetDataSize() / SuccinctSize() (column_segment.cpp:204-222) for a GPU-resident segment
	static idx_t DataSize(ColumnSegment &segment);
	//! sdsl::size_in_bytes of the slot-width vector of a fresh segment (column_segment.cpp:101-105)
	static idx_t SlotWidthSize(idx_t segment_size, idx_t type_size);
};

} // namespace duckdb
