// adacom_kernels.cuh -- host-callable launchers of the codec kernels.  Each launcher enqueues exactly
// ONE kernel on `stream` and returns the cudaError_t of the launch.
#pragma once

#include "adacom_device.cuh"
#include "filter_range.hpp"

#include <string>

namespace adacom {

enum DecodeMode : int { kModeFlat = 0, kModeSum = 1, kModeFilter = 2, kModePoint = 3 };

struct TypeInfo {
	int size;        // 1, 2, 4, 8
	bool is_signed;
};

struct DecodeArgs {
	const SegDesc *desc = nullptr;
	// kModeFlat: (segment, tile) pairs of 2048 values from the column's static tile table
	const Job *jobs = nullptr;
	uint32_t n_jobs = 0;
	// fused scans: n_units segments, seg_b + unit (unit_seg == nullptr) or unit_seg[unit]; the kernel enumerates
	// the <= kScanParts chunks of <= 16 KiB of every segment from its descriptor (their number depends on the
	// current width, so there is no table)
	uint32_t n_units = 0;
	uint32_t seg_b = 0;
	const uint32_t *unit_seg = nullptr;
	const uint32_t *parts = nullptr;      // device word: chunks per segment the column currently needs (<= kScanParts)
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
	// kModePoint: unit == query
	const int64_t *q_const = nullptr;
	unsigned long long *q_count = nullptr;
	unsigned long long *q_first = nullptr;
};

// TMA-staged unpack of packed tiles (+ fused SUM / filter).  grid = persistent CTAs.
cudaError_t launch_decode(TypeInfo t, DecodeMode mode, const DecodeArgs &args, int sm_count, cudaStream_t stream);

// width-specialised fused scan+SUM over re-packed segments whose decoded interval is sign/truncation linear
struct ScanSumArgs {
	const SegDesc *desc = nullptr;
	uint32_t seg_b = 0, seg_e = 0;        // segment range to scan (chunks are derived from the descriptors)
	const uint32_t *parts = nullptr;      // device word: chunks per segment the column currently needs (<= kScanParts)
	const uint8_t *raw = nullptr;         // raw arena (source of segments that are not re-packed)
	int literal = 0;
	int type_size = 8;
	int is_signed = 1;
	Agg *agg = nullptr;                   // zeroed by the caller, one atomic triple per CTA
};

cudaError_t launch_scan_sum_fields(const ScanSumArgs &args, int sm_count, cudaStream_t stream);

// ---- fused multi-column scan (conjunctive range filters + SUM of affine products), kernels_query.cu ----
constexpr int kQueryMaxCols = 8;
constexpr int kQueryMaxPieces = 3;   // segments of one column a row block may touch
constexpr int kQueryMaxAggs = 8;
constexpr int kQueryMaxFactors = 3;
constexpr int kQueryMaxGroups = 8;
constexpr int kQueryConsumerWarps = 4;   // consumer warps per CTA of scan_columns_kernel (+ one producer warp)
constexpr uint32_t kQueryMaxBlockRows = 4096;  // a consumer lane owns one 32-row group of the block

struct QueryCol {
	const SegDesc *desc = nullptr;
	const uint32_t *vec_seg = nullptr;  // segment holding row 2048*v of the column, v = 0 .. ceil(rows / 2048)
	const uint8_t *raw = nullptr;
	uint32_t n_segs = 0;
	uint32_t type_size = 8;
	uint32_t is_signed = 1;
	uint32_t has_pred = 0;
	int64_t lo = 0, hi = 0;             // inclusive range on the decoded T value (compared as T)
};

struct QueryAgg {  // sum over the selected rows of prod_k (add[k] + mul[k] * value(col[k])); 0 factors = COUNT(*)
	uint32_t n_factors = 0;
	uint32_t col[kQueryMaxFactors] = {0, 0, 0};
	int64_t add[kQueryMaxFactors] = {0, 0, 0};
	int64_t mul[kQueryMaxFactors] = {0, 0, 0};
};

// The sum expressions of a query flattened into one factor program (scan_columns_dense_kernel): per step
// x = add + mul * value(col); term = fresh ? x : term * x; if (emit) accumulator[emit - 1] += term.  The planner orders
// the aggregates so that one whose factor list extends another's continues that term (TPC-H Q1: price, price*(100-d),
// price*(100-d)*(100+t) are three steps, not six), and the kernel walks the program in a real loop: the unrolled
// kQueryMaxAggs x kQueryMaxFactors form issued every predicated-off slot for every row.
struct QueryStep {
	int64_t add = 0, mul = 0;
	uint16_t col = 0;
	uint8_t fresh = 1;
	uint8_t emit = 0;   // 0: keep going; else index of the aggregate + 1
	uint32_t reserved = 0;
};
constexpr int kQueryMaxSteps = kQueryMaxAggs * kQueryMaxFactors;

// a run of table rows inside which every referenced column is ONE segment (scan_columns_dense_kernel)
struct DenseRange {
	uint64_t row_begin;
	uint32_t rows;
	uint32_t first_block;            // index of the range's first 1024-row block
	uint32_t seg[kQueryMaxCols];     // segment of every column
};
constexpr uint32_t kDenseBlockRows = 1024;
constexpr uint32_t kDenseThreads = 128;

struct QueryArgs {
	QueryCol cols[kQueryMaxCols];
	QueryAgg aggs[kQueryMaxAggs];
	QueryStep steps[kQueryMaxSteps];
	uint32_t n_steps = 0;
	uint32_t n_cols = 0, n_aggs = 0;
	uint32_t factor_mask = 0;             // columns some sum expression reads
	uint64_t row_begin = 0, row_end = 0;  // table rows; row_begin is a multiple of 2048
	uint32_t block_rows = 2048;           // 2048 | 4096 | 8192
	uint32_t stages = 2;
	uint32_t stage_bytes = 0;             // multiple of 128: worst-case staged bytes of one row block
	int literal = 0;
	uint32_t narrow_values = 0;           // every value and every factor (add + mul * value) fits int32 (dense kernel)
	uint32_t narrow_products = 0;         // every prod_k (add + mul * value) provably fits int64: 64-bit products
	uint32_t group_col = 0;               // GROUP BY: column whose decoded value (0 .. n_groups-1) is the group key
	uint32_t n_groups = 0;                // 0: no grouping; rows whose key is outside the range are dropped
	uint32_t dense_groups = 0;            // grouped launches: per-thread int64 accumulators in shared memory are safe
	                                      // (every partial sum provably fits) -> dense selections skip the lists
	const DenseRange *dense_ranges = nullptr;  // grouped launches on field-comparable columns: the uniform ranges (device)
	uint32_t dense_n_ranges = 0, dense_blocks = 0;
	uint32_t dense_window = 0;            // bytes of shared memory a warp needs for all the columns of 1024 rows
	uint32_t sparse_window = 0;           // ungrouped launches: bytes of shared memory a warp needs for the predicate
	                                      // columns of 1024 rows (0: use the pipelined kernel)
	const uint32_t *in_bitmap = nullptr;  // optional selection (bit r - row_begin), AND-ed
	uint32_t *out_bitmap = nullptr;       // optional: the resulting selection
	Agg *out = nullptr;                   // n_aggs (x n_groups, group-major) aggregates, zeroed by the caller;
	                                      // count = selected rows (of the group)
};

static_assert(sizeof(QueryArgs) <= 4000, "QueryArgs travels as a kernel parameter");
cudaError_t launch_scan_columns(const QueryArgs &args, int sm_count, cudaStream_t stream);
// extra dynamic shared memory of a grouped launch (bytes); dense: + the per-thread accumulators
uint32_t scan_columns_group_smem(uint32_t n_groups, uint32_t n_aggs, uint32_t dense);

// scan_vector / scan_partial of one segment, arbitrary start (generic unaligned path, no staging)
cudaError_t launch_scan_range(TypeInfo t, const SegDesc *desc, uint32_t seg, const uint8_t *raw, uint64_t start,
                              uint64_t scan_count, int literal, void *result, uint64_t result_offset,
                              cudaStream_t stream);

// batched fetch_row gather
cudaError_t launch_fetch_one(TypeInfo t, const SegDesc *desc, const uint8_t *raw, uint32_t seg, uint32_t row, int literal,
                             void *out, cudaStream_t stream);
cudaError_t launch_fetch_rows(TypeInfo t, const SegDesc *desc, const uint8_t *raw, uint64_t n, const uint32_t *seg_idx,
                              const uint32_t *row_in_seg, int literal, void *out, cudaStream_t stream, uint32_t n_segs = 0,
                              unsigned long long *bad = nullptr);  // bad: lowest out-of-range query index (atomicMin), or NULL

// batched equality look-ups whose queries share the chunks of their segment (byte-aligned widths 8 / 16 / 32 of
// re-packed segments only; the host falls back to launch_decode(kModePoint) otherwise).  Queries are sorted by segment;
// q_const / q_count / q_first are in that order.
constexpr uint32_t kPointSlice = 64;  // queries per job
struct PointJob {
	uint32_t seg;
	uint32_t chunk;       // 16 KiB chunk of the segment's payload
	uint32_t q_lo, q_hi;  // slice of the sorted queries
};
cudaError_t launch_point_shared(const SegDesc *desc, const PointJob *jobs, uint32_t n_jobs, const int64_t *q_const,
                                unsigned long long *q_count, unsigned long long *q_first, int type_size, int sm_count,
                                cudaStream_t stream);

// analyze: per-segment unsigned min/max (+ optional validity / NullValue fill); one CTA per segment
cudaError_t launch_analyze(TypeInfo t, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint8_t *raw,
                           const uint64_t *validity, int zero_ext, cudaStream_t stream,
                           const uint32_t *seg_list = nullptr /* device: the segments, instead of the range */);

// layout: widths, repack decision, 128-byte aligned word offsets (exclusive scan); totals (host-mapped) =
// {words, ticket}
cudaError_t launch_layout(TypeInfo t, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int flags, uint32_t extent,
                          unsigned long long *totals, unsigned long long ticket, uint32_t *parts, int whole_column,
                          unsigned long long *scratch /* 128 words of device memory */, cudaStream_t stream,
                          const uint32_t *seg_list = nullptr);

// bind packed_ptr = base + word_offset for the repacked segments of the range
cudaError_t launch_bind_extent(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint64_t *base, uint32_t extent,
                               cudaStream_t stream, const uint32_t *seg_list = nullptr);

// rollback of launch_layout when the encode behind it cannot complete: unbound descriptors of `extent` get their old width / state
cudaError_t launch_revert_layout(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint32_t extent, uint32_t *parts,
                                 uint32_t raw_parts, cudaStream_t stream, const uint32_t *seg_list = nullptr);
// after an uncompact decode: segments of the range that were compacted go back to slot width / UNCOMPRESSED
cudaError_t launch_reset_uncompacted(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int type_size,
                                     uint32_t *parts, uint32_t raw_parts, cudaStream_t stream,
                                     const uint32_t *seg_list = nullptr);

// encode: FOR subtract + width-specialised bit-pack of the tiles whose segment the preceding layout assigned to
// `extent` (payload at base + word_offset; launch_bind_extent publishes the pointers afterwards)
cudaError_t launch_encode(TypeInfo t, const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw,
                          uint32_t extent, uint64_t *base, int sm_count, cudaStream_t stream);

// synthetic generator: raw[i] = T(splitmix64(seed + first_index + i) & mask)
cudaError_t launch_fill_splitmix(TypeInfo t, void *raw, uint64_t n, uint64_t seed, uint64_t first_index, uint64_t mask,
                                 cudaStream_t stream);

// fused decode + compare + SUM/COUNT (+ bitmap) in the field domain, for ranges whose re-packed payloads are all
// <= 32 bits wide (kernels_filter.cu); seg_capacity: upper bound of the values per segment
cudaError_t launch_filter_fields(TypeInfo t, const SegDesc *desc, uint32_t seg_b, uint32_t n_segs, uint32_t seg_capacity,
                                 const uint8_t *raw, int literal, uint32_t op_mask, uint64_t constant, uint32_t *bitmap32,
                                 uint64_t out_row0, Agg *agg, int sm_count, cudaStream_t stream);

// ---- general single-column table filter (kernels_select.cu): conjunction of comparisons / IS [NOT] NULL, validity
// mask, selection bitmap in / out; per-value decode, any width
// (FilterRange, SelectPred, kSelectMaxPreds: filter_range.hpp)
struct SelectArgs {
	const SegDesc *desc = nullptr;
	const Job *jobs = nullptr;             // 2048-value tiles of the segment range (static tile table)
	uint32_t n_jobs = 0;
	const uint8_t *raw = nullptr;
	int literal = 0;
	uint32_t n_preds = 0;
	uint32_t conj_or = 0;
	SelectPred preds[kSelectMaxPreds];
	const uint64_t *validity = nullptr;    // column-wide validity words (bit = column row), device; nullptr = all valid
	const uint64_t *in_bitmap = nullptr;   // incoming selection, bit (row - out_row0)
	uint32_t *out_bitmap32 = nullptr;      // resulting selection (zeroed by the caller)
	uint64_t out_row0 = 0;
	Agg *agg = nullptr;                    // sum over selected valid rows, count = selected rows, reserved = selected valid rows
};
cudaError_t launch_select(TypeInfo t, const SelectArgs &args, int sm_count, cudaStream_t stream);

// ---- neighbour codecs: DuckDB bitpacking / RLE segment images (kernels_codecs.cu) ----
struct CodecSeg {
	uint64_t image_offset;  // of the segment's block image inside the device arena (multiple of 16)
	uint64_t row_start;     // first row of the segment inside the column
	uint32_t bytes;         // image size
	uint32_t count;         // tuples
	uint32_t first_run;     // RLE: index of the segment's first run in the run_start array
	uint32_t reserved;
};
// One record per (segment, 2048-value group), resolved ONCE when the column is created (like the tile table of the
// succinct codec): the scan kernels start from it instead of walking job -> segment -> metadata entry -> group header,
// four dependent global loads per 2048 values.
struct CodecGroup {
	uint64_t payload;       // bitpacking: byte offset of the group's packed payload in the image arena; RLE: of the run values
	uint64_t out_row;       // column row of the group's first value
	uint64_t frame, constant, delta_offset;  // bitpacking group header (bitpacking.cpp:576-632)
	uint32_t mode, width;   // bitpacking
	uint32_t count;         // values in the group
	uint32_t run_lo, run_hi;  // RLE: run_start indices of the run holding the group's first row / the next group's first row
	uint32_t first;         // RLE: row of the group's first value inside its segment
};
static_assert(sizeof(CodecGroup) == 64, "CodecGroup is one 64-byte record");
struct CodecArgs {
	const uint8_t *images = nullptr;
	const CodecSeg *segs = nullptr;
	const Job *jobs = nullptr;      // (segment, 2048-value group) -- or (segment, tile of 2048 runs) for the RLE SUM
	const CodecGroup *groups = nullptr;  // records of the same groups (same slice as jobs; not for the RLE SUM tiles)
	uint32_t n_jobs = 0;
	void *out = nullptr;            // decode: flat T values, out[row - out_row0]; fetch: one value per query
	uint64_t out_row0 = 0;
	Agg *agg = nullptr;
	uint32_t *run_start = nullptr;  // RLE: exclusive prefix sums of the run lengths
};
// codec 0 = bitpacking, 1 = RLE; mode 0 = decode to flat values, 1 = fused SUM/COUNT
// text behind adacom_last_error(NULL) (adacom_capi.cu), for entry points that live outside adacom_capi.cu
void set_process_error(const std::string &msg);
cudaError_t launch_codec_scan(TypeInfo t, int codec, int mode, const CodecArgs &args, int sm_count, cudaStream_t stream);
cudaError_t launch_codec_fetch(TypeInfo t, int codec, const CodecArgs &args, const uint32_t *seg_idx,
                               const uint32_t *row_in_seg, uint64_t n, cudaStream_t stream);
cudaError_t launch_rle_run_starts(const CodecArgs &args, uint32_t seg_b, uint32_t n_segs, int sm_count, cudaStream_t stream);
// fills groups_out[j] for the jobs of args (after launch_rle_run_starts for RLE)
cudaError_t launch_codec_groups(TypeInfo t, int codec, const CodecArgs &args, CodecGroup *groups_out, cudaStream_t stream);

// the same operator, warp per 2048-value tile (static tile table), lane per 32-row group (pred32); handles every kind of
// segment (field-comparable ones in the field domain, the others per value); wmax: widest live payload of the column
// valid32 (bit = column row) / in32 (bit = row - out_row0): optional masks AND-ed into the selection -- NULL rows and
// rows outside the incoming selection neither pass nor enter the SUM
cudaError_t launch_filter_words(TypeInfo t, const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw,
                                int literal, const FilterRange &range, const uint32_t *valid32, const uint32_t *in32,
                                uint32_t *bitmap32, uint64_t out_row0, Agg *agg, uint32_t wmax, int sm_count,
                                cudaStream_t stream);

// ---- checkpoint image (kernels_checkpoint.cu) ----
constexpr uint32_t kBlockAllocSize = 262144;  // Storage::BLOCK_ALLOC_SIZE: 8-byte checksum + Storage::BLOCK_SIZE payload

// One payload inside one block.  checkpoint: addr = source (device address of the packed words / the slot-width
// values, only sizeof(T)-aligned for the latter); restore: addr = destination.
struct CkPiece {
	uint64_t addr;
	uint32_t block_off;  // byte offset inside the block payload (multiple of 8)
	uint32_t bytes;      // payload bytes (8 * words, or count * sizeof(T))
	uint32_t pad_words;  // restore into the packed extent: words up to the next 128-byte boundary (zero filled)
	uint32_t elem_size;  // sizeof(T) for slot-width payloads, 8 for packed words
};
static_assert(sizeof(CkPiece) == 24, "CkPiece layout");

// payloads -> 256 KiB block images with DuckDB block checksums; pieces of block b = [piece_begin[b], piece_begin[b+1])
cudaError_t launch_checkpoint_blocks(const CkPiece *pieces, const uint32_t *piece_begin, uint8_t *blocks,
                                     uint32_t n_blocks, int sm_count, cudaStream_t stream);
// block images -> payloads, every block checksum verified in the same pass (*bad_block = lowest failing block id,
// initialised to ~0 by the caller)
cudaError_t launch_restore_blocks(const CkPiece *pieces, const uint32_t *piece_begin, const uint8_t *blocks,
                                  uint32_t n_blocks, unsigned long long *bad_block, int sm_count, cudaStream_t stream);

}  // namespace adacom
