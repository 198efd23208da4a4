// adacom_kernels.cuh -- host-callable launchers of the codec kernels.  Each launcher enqueues exactly
// ONE kernel on `stream` and returns the cudaError_t of the launch.
#pragma once

#include "adacom_device.cuh"

namespace adacom {

enum DecodeMode : int { kModeFlat = 0, kModeSum = 1, kModeFilter = 2, kModePoint = 3 };

struct TypeInfo {
	int size;        // 1, 2, 4, 8
	bool is_signed;
};

struct DecodeArgs {
	const SegDesc *desc = nullptr;
	const Job *jobs = nullptr;       // (segment, k) work units
	uint32_t n_jobs = 0;
	// device-built chunk table: when chunk_first != nullptr the kernel takes the jobs
	// [chunk_first[seg_b], chunk_first[seg_e]) of `jobs` (n_jobs is ignored)
	const uint32_t *chunk_first = nullptr;
	uint32_t seg_b = 0, seg_e = 0;
	uint32_t tile_granularity = 0;   // kVectorSize -> 2048-value tiles; 0 -> scan chunks (scan_chunk_values(w))
	const uint8_t *raw = nullptr;    // raw arena (source for segments that are not repacked)
	int literal = 0;                 // ADACOM_LITERAL: add min even when the segment was not repacked (Q1)
	int uncompact = 0;               // UncompressSuccinct mode: skip segments that are not compacted
	// kModeFlat
	void *out = nullptr;
	uint64_t out_row0 = 0;           // out[(row_start + i) - out_row0]
	// kModeSum / kModeFilter
	Agg *agg = nullptr;
	uint32_t op_mask = 0;
	uint64_t constant = 0;
	uint32_t *bitmap32 = nullptr;    // optional, bit (row - out_row0)
	// kModePoint
	const uint32_t *job_query = nullptr;
	const int64_t *q_const = nullptr;
	unsigned long long *q_count = nullptr;
	unsigned long long *q_first = nullptr;
};

// TMA-staged unpack of packed tiles (+ fused SUM / filter).  grid = persistent CTAs.
cudaError_t launch_decode(TypeInfo t, DecodeMode mode, const DecodeArgs &args, int sm_count, cudaStream_t stream);

// width-specialised fused scan+SUM over re-packed segments whose decoded interval is sign/truncation linear
struct ScanSumArgs {
	const SegDesc *desc = nullptr;
	const Job *jobs = nullptr;            // the column's device-built chunk table (Job.k bit 31 = fast path)
	const uint32_t *chunk_first = nullptr;  // first chunk of every segment (+ total at [n_segs])
	uint32_t seg_b = 0, seg_e = 0;        // segment range to scan
	const uint8_t *raw = nullptr;         // raw arena (source of segments that are not re-packed)
	int literal = 0;
	int type_size = 8;
	int is_signed = 1;
	Agg *agg = nullptr;
};

constexpr uint32_t kJobFastBit = 0x80000000u;

// chunk table of the whole column, built on the device from the current descriptors (no host round trip)
cudaError_t launch_build_chunks(const SegDesc *desc, uint32_t n_segs, int type_size, int is_signed, Job *jobs_out,
                                uint32_t *chunk_first, cudaStream_t stream);
cudaError_t launch_scan_sum_fields(const ScanSumArgs &args, int sm_count, cudaStream_t stream);

// scan_vector / scan_partial of one segment, arbitrary start (generic unaligned path, no staging)
cudaError_t launch_scan_range(TypeInfo t, const SegDesc *desc, uint32_t seg, const uint8_t *raw, uint64_t start,
                              uint64_t scan_count, int literal, void *result, uint64_t result_offset,
                              cudaStream_t stream);

// batched fetch_row gather
cudaError_t launch_fetch_rows(TypeInfo t, const SegDesc *desc, const uint8_t *raw, uint64_t n, const uint32_t *seg_idx,
                              const uint32_t *row_in_seg, int literal, void *out, cudaStream_t stream);

// analyze: per-segment unsigned min/max (+ optional validity / NullValue fill); one CTA per segment
cudaError_t launch_analyze(TypeInfo t, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint8_t *raw,
                           const uint64_t *validity, int zero_ext, cudaStream_t stream);

// layout: widths, repack decision, 128-byte aligned word offsets (exclusive scan), totals[0] = words
cudaError_t launch_layout(TypeInfo t, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int flags, uint32_t extent,
                          unsigned long long *totals, cudaStream_t stream);

// bind packed_ptr = base + word_offset for the repacked segments of the range
cudaError_t launch_bind_extent(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint64_t *base, uint32_t extent,
                               cudaStream_t stream);

// after an uncompact decode: segments of the range that were compacted go back to slot width / UNCOMPRESSED
cudaError_t launch_reset_uncompacted(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int type_size,
                                     cudaStream_t stream);

// encode: TMA-staged raw tiles -> FOR subtract + bit-pack into the packed extent
cudaError_t launch_encode(TypeInfo t, const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw,
                          uint32_t extent, int sm_count, cudaStream_t stream);

// synthetic generator: raw[i] = T(splitmix64(seed + first_index + i) & mask)
cudaError_t launch_fill_splitmix(TypeInfo t, void *raw, uint64_t n, uint64_t seed, uint64_t first_index, uint64_t mask,
                                 cudaStream_t stream);

}  // namespace adacom
