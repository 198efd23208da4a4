/*
 * adacom_gpu.h -- C ABI of the B200-native succinct column-segment codec (libadacom_gpu.so).
 *
 * This is the drop-in boundary for AdaCom's succinct hot path.  Plain pointers and sizes only; every
 * entry point returns an adacom_status and never throws.  The C++ shim in
 * duckdb_adaptive_compression_b200/host/ wraps these calls behind the reference's own
 * CompressionFunction / ColumnSegment / ColumnSegmentCatalog surface; INTEGRATION.md shows the
 * binding a maintainer of the reference would add.
 *
 * Reference interfaces replaced (paths relative to the reference repository root):
 *   plugin vtable ............ src/include/duckdb/function/compression_function.hpp:65-177
 *   SuccinctFun::GetFunction . src/storage/compression/succinct.cpp:335-366
 *   append + min/max ......... src/storage/compression/succinct.cpp:271-322          -> adacom_analyze
 *   BitCompressFromSuccinct .. src/storage/table/column_segment.cpp:348-383          -> adacom_encode
 *   BitCompressFromUncompr. .. src/storage/table/column_segment.cpp:385-456          -> adacom_encode(ADACOM_ZERO_EXT_MINMAX)
 *   SuccinctScan/ScanPartial . src/storage/compression/succinct.cpp:123-144,232-240  -> adacom_scan_partial, adacom_decode
 *   filter after scan ........ src/storage/table/column_segment.cpp:575-658          -> adacom_scan_filter (fused)
 *   UncompressSuccinct ....... src/storage/table/column_segment.cpp:458-506          -> adacom_uncompact
 *   SuccinctFetchRow ......... src/storage/compression/succinct.cpp:244-260          -> adacom_fetch_rows
 *   GetDataSize .............. src/storage/table/column_segment.cpp:204-214          -> adacom_column_data_size
 *
 * Object model.  An adacom_ctx owns one GPU (device ordinal, streams, scratch).  An adacom_column is a
 * batch of column segments of one physical type resident in that GPU's HBM: a raw arena (the slot-width
 * int_vector payloads == flat T values, segments back to back), a packed arena (FOR + fixed-width
 * SDSL-layout words, every segment starting on a 128-byte boundary) and a descriptor table.  All batched
 * calls take a [seg_begin, seg_end) range and run as ONE kernel launch over that range.
 *
 * There is no CPU fallback: every compute entry point fails with ADACOM_ERR_CUDA / ADACOM_ERR_NO_DEVICE
 * when no sm_100 device is usable.
 */
#ifndef ADACOM_GPU_H
#define ADACOM_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADACOM_ABI_VERSION 1

typedef enum {
	ADACOM_OK = 0,
	ADACOM_ERR_INVALID = 1,     /* bad argument (the shim raises duckdb::InternalException) */
	ADACOM_ERR_CUDA = 2,        /* a CUDA call failed; adacom_last_error() has the text */
	ADACOM_ERR_NOMEM = 3,       /* device or pinned-host allocation failed (reference: std::bad_alloc) */
	ADACOM_ERR_UNSUPPORTED = 4, /* physical type not supported by SuccinctFun::TypeIsSupported */
	ADACOM_ERR_STATE = 5,       /* call not valid in the column's current state */
	ADACOM_ERR_NO_DEVICE = 6,   /* no CUDA device / not an sm_100 part */
	ADACOM_ERR_CORRUPT = 7      /* checkpoint image fails its checksums (reference: IOException "Corrupt database
	                               file", src/common/file_buffer.cpp:95-106) or is malformed */
} adacom_status;

/* duckdb::PhysicalType values of the supported types (src/include/duckdb/common/types.hpp) */
typedef enum {
	ADACOM_UINT8 = 2,
	ADACOM_INT8 = 3,
	ADACOM_UINT16 = 4,
	ADACOM_INT16 = 5,
	ADACOM_UINT32 = 6,
	ADACOM_INT32 = 7,
	ADACOM_UINT64 = 8,
	ADACOM_INT64 = 9
} adacom_ptype;

/* comparison operators of TemplatedFilterSelection (column_segment.cpp:593-655) */
typedef enum {
	ADACOM_CMP_EQ = 0,
	ADACOM_CMP_NE = 1,
	ADACOM_CMP_LT = 2,
	ADACOM_CMP_GT = 3,
	ADACOM_CMP_LE = 4,
	ADACOM_CMP_GE = 5
} adacom_cmp;

/* flags (mirror DBConfig's succinct switches, src/include/duckdb/main/config.hpp:189-197) */
enum {
	ADACOM_PADDED = 1,          /* succinct_padded_to_next_byte_enabled: w = (w + 7) & ~7 */
	ADACOM_LITERAL = 2,         /* reproduce reference quirks Q1/Q2 literally */
	ADACOM_ZERO_EXT_MINMAX = 4  /* adaptive encoder: min/max over zero-extended raw bytes (Q2b) */
};

/* per-segment state bits */
enum {
	ADACOM_SEG_COMPACTED = 1, /* ColumnSegment::compacted */
	ADACOM_SEG_REPACKED = 2,  /* the width really shrank: payload lives in the packed arena */
	ADACOM_SEG_ANALYZED = 4   /* min_factor / max_factor are valid */
};

/* Host view of one segment == the codec-relevant ColumnSegment members
 * (src/include/duckdb/storage/table/column_segment.hpp:60,190-214). 64 bytes. */
typedef struct {
	uint64_t row_start;     /* index of the segment's first value inside the column (SegmentBase::start) */
	uint64_t count;         /* SegmentBase::count */
	uint64_t word_offset;   /* first 64-bit word of the payload in the packed arena (multiple of 16) */
	uint64_t words;         /* payload words = ceil(count * width / 64) when repacked, else 0 */
	uint64_t min_factor;    /* UINT64_MAX = unset */
	uint64_t max_factor;
	uint64_t size_in_bytes; /* sdsl::size_in_bytes(succinct_vec) == ColumnSegment::GetDataSize() */
	uint8_t width;          /* succinct_vec.width() */
	uint8_t state;          /* ADACOM_SEG_* */
	uint8_t type_size;      /* 1, 2, 4, 8 */
	uint8_t is_signed;
	uint32_t capacity;      /* slots of the segment: segment_size / type_size */
} adacom_segment_meta;

/* 128-bit two's-complement aggregate (DuckDB SUM over integers -> HUGEINT) */
typedef struct {
	uint64_t sum_lo;
	uint64_t sum_hi;
	uint64_t count;
	uint64_t reserved;     /* adacom_scan_select / adacom_scan_sum_valid: selected rows that are not NULL; else 0 */
} adacom_agg;

typedef struct adacom_ctx adacom_ctx;
typedef struct adacom_column adacom_column;

/* ---- context -------------------------------------------------------------------------------------- */
int adacom_abi_version(void);
adacom_status adacom_ctx_create(int device, adacom_ctx **out);
void adacom_ctx_destroy(adacom_ctx *ctx);
/* text of the last error raised on this context (or on the process when ctx == NULL) */
const char *adacom_last_error(const adacom_ctx *ctx);
adacom_status adacom_ctx_synchronize(adacom_ctx *ctx);
/* the compute stream (cudaStream_t) every kernel of this context is launched on */
void *adacom_ctx_stream(adacom_ctx *ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
uint64_t adacom_ctx_launch_count(const adacom_ctx *ctx);
/* CUDA events on the compute stream, for timing the kernels where they are launched */
adacom_status adacom_event_create(adacom_ctx *ctx, void **event_out);
adacom_status adacom_event_record(adacom_ctx *ctx, void *event);
adacom_status adacom_event_elapsed_ms(adacom_ctx *ctx, void *start, void *stop, float *ms_out); /* syncs on stop */
void adacom_event_destroy(adacom_ctx *ctx, void *event);
/* pinned host memory for staging (cudaHostAlloc) */
adacom_status adacom_host_alloc(adacom_ctx *ctx, size_t bytes, void **out);
void adacom_host_free(adacom_ctx *ctx, void *ptr);
/* device memory on the context's GPU for callers that do not link the CUDA runtime themselves: the selection bitmaps
 * adacom_scan_filter / adacom_scan_select exchange (1 bit per row, 8-byte words), device-side fetch batches */
adacom_status adacom_device_alloc(adacom_ctx *ctx, size_t bytes, void **out);
void adacom_device_free(adacom_ctx *ctx, void *ptr);
/* write `bytes` of zeros to a scratch buffer > L2 (benchmark hygiene: flush the 126 MB L2) */
adacom_status adacom_flush_l2(adacom_ctx *ctx);

/* ---- column life cycle ---------------------------------------------------------------------------- */
/* A column of n_segments segments; segment g holds seg_counts[g] values and has seg_capacity slots
 * (Storage::BLOCK_SIZE / sizeof(T) = 262136/S by default when seg_capacity == 0).  Allocates the raw
 * arena (sum(seg_counts) * S bytes) and the descriptor table.  Fresh segments are SUCCINCT at slot width
 * 8*S, not compacted (column_segment.cpp:101-105). */
adacom_status adacom_column_create(adacom_ctx *ctx, adacom_ptype type, uint64_t n_segments,
                                   const uint64_t *seg_counts, uint32_t seg_capacity, adacom_column **out);
void adacom_column_destroy(adacom_column *col);
uint64_t adacom_column_num_values(const adacom_column *col);
uint64_t adacom_column_num_segments(const adacom_column *col);
/* device pointer of the raw arena (T values, segments back to back; NULL once released) */
void *adacom_column_raw_device_ptr(adacom_column *col);
/* device pointer of the packed arena (uint64 words; NULL before the first encode) */
void *adacom_column_packed_device_ptr(adacom_column *col);
/* copy host T values [first_value, first_value + n) into the raw arena (async on the compute stream) */
adacom_status adacom_column_upload_raw(adacom_column *col, const void *host_values, uint64_t first_value, uint64_t n);
/* free the raw arena once every segment is repacked (ERR_STATE otherwise) */
adacom_status adacom_column_release_raw(adacom_column *col);
/* Build a column directly from packed images (the H2D staging format): metas[] as produced by
 * adacom_column_get_meta, host_words = the packed arena (metas[g].word_offset indexes into it). */
adacom_status adacom_column_from_packed(adacom_ctx *ctx, adacom_ptype type, uint64_t n_segments,
                                        const adacom_segment_meta *metas, const uint64_t *host_words,
                                        uint64_t total_words, adacom_column **out);
/* synthetic-data generator for benchmarks: raw[i] = T(splitmix64(seed + first_index + i) & mask) */
adacom_status adacom_column_fill_splitmix(adacom_column *col, uint64_t seed, uint64_t first_index, uint64_t mask);

/* ---- analyze (SuccinctAppendLoop's min/max; optional validity mask + NullValue<T> fill) ------------- */
/* validity: DuckDB validity words for the WHOLE column (bit i = row i valid), device or host pointer per
 * validity_on_device; NULL = all valid.  NULL rows are excluded from min/max and their slot is
 * overwritten with NullValue<T>() exactly like succinct.cpp:288-292. */
adacom_status adacom_analyze(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                             const uint64_t *validity, int validity_on_device);

/* ---- encode (BitCompressFromSuccinct / BitCompressFromUncompressed) -------------------------------- */
/* width = hi(max - min) + 1 (padded to bytes with ADACOM_PADDED); segments whose width shrinks are packed
 * into the packed arena, all segments of the range become `compacted`.  Re-encoding a range replaces the
 * packed arena (epoch swap: scans never observe a half-written payload). */
adacom_status adacom_encode(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags);

/* ---- segment-list forms (the access-frequency policy) -------------------------------------------------- */
/* ColumnSegmentCatalog::CompactSegments (src/catalog/catalog_entry/column_segment_catalog.cpp:64-116) sorts the segments
 * by read count, Compact()s the coldest 90 % and Uncompact()s the rest: arbitrary subsets of a column, not ranges.
 * These calls take such a subset (host array of segment ids, strictly ascending) and run ONE launch set for it,
 * with the semantics of the range forms.  device_validity: as adacom_analyze, device pointer or NULL. */
adacom_status adacom_analyze_list(adacom_column *col, uint64_t n, const uint32_t *segs, int flags,
                                  const uint64_t *device_validity);
adacom_status adacom_encode_list(adacom_column *col, uint64_t n, const uint32_t *segs, int flags);
adacom_status adacom_uncompact_list(adacom_column *col, uint64_t n, const uint32_t *segs, int flags);

/* ---- metadata / accounting --------------------------------------------------------------------------- */
adacom_status adacom_column_get_meta(adacom_column *col, uint64_t seg_begin, uint64_t seg_end,
                                     adacom_segment_meta *out);
/* sum of ColumnSegment::GetDataSize() over the range (9 + 8*ceil(bits/64) per succinct segment) */
adacom_status adacom_column_data_size(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, uint64_t *out);
/* total payload words of the packed arena (including the 128-byte alignment gaps) */
uint64_t adacom_column_packed_words(const adacom_column *col);
/* copy the packed arena / one segment's payload words to the host */
adacom_status adacom_column_download_packed(adacom_column *col, uint64_t first_word, uint64_t n_words,
                                            uint64_t *host_out);
/* sdsl::int_vector<0>::serialize image of one segment: 8-byte bit count, 1-byte width, data words */
adacom_status adacom_segment_serialize(adacom_column *col, uint64_t seg, uint8_t *host_out, uint64_t capacity,
                                       uint64_t *bytes_out);

/* ---- decode / scan ------------------------------------------------------------------------------------ */
/* Batched decode of [seg_begin, seg_end) into a flat T array: out[row - row_start(seg_begin)].
 * out is a device pointer when out_on_device, else a host pointer (pinned or pageable). */
adacom_status adacom_decode(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags, void *out,
                            int out_on_device);
/* scan_vector / scan_partial of ONE segment: result[result_offset + i] = value(start + i), i < scan_count
 * (SuccinctScanPartial, succinct.cpp:123-144).  scan_count may exceed 2048. */
adacom_status adacom_scan_partial(adacom_column *col, uint64_t seg, uint64_t start, uint64_t scan_count,
                                  int flags, void *result, uint64_t result_offset, int result_on_device);
/* ---- pull-model scan (the vtable path) ------------------------------------------------------------------
 * ColumnData::ScanVector (src/storage/table/column_data.cpp:92-139) pulls <= 2048 values per call and walks the
 * segments of a column in order; SuccinctScanPartial (succinct.cpp:123-144) re-copies and decodes per call.  A scan
 * stream serves that loop from the device: segments [seg_begin, seg_end) are decoded `window_segments` at a time
 * (one launch per window) on a stream of the scan's own and copied into a ring of `depth` pinned host windows that
 * run ahead of the consumer, so scan_vector / scan_partial become pointer arithmetic into pinned memory (zero-copy
 * FlatVector::SetData, like FixedSizeScan, fixed_size_uncompressed.cpp).  No call of the scan holds the context
 * mutex while it waits; any number of streams (one per DuckDB scan thread) run concurrently.
 * window_segments = 0 picks a read-ahead ramp (windows of 1, 2, 4, ... 32 segments: the first vector of a short scan
 * waits for one segment, a long scan settles at 8 MiB per launch); depth = 0 picks 3 windows (depth >= 2). */
typedef struct adacom_scan_stream adacom_scan_stream;
adacom_status adacom_scan_stream_open(adacom_column *col, uint64_t seg_begin, uint64_t seg_end,
                                      uint32_t window_segments, uint32_t depth, int flags, adacom_scan_stream **out);
/* Decoded values of segment `seg` (flat T array, `*count` values) in pinned host memory.  Blocks until the window
 * holding it has arrived; touching a window issues the ones after it.  The pointer stays valid until the stream has
 * moved depth-1 windows past the segment's window or is closed.  Segments may be revisited inside the current and
 * the previous window for free; anything else restarts the pipeline at that window. */
adacom_status adacom_scan_stream_segment(adacom_scan_stream *s, uint64_t seg, const void **values, uint64_t *count);
void adacom_scan_stream_close(adacom_scan_stream *s);

/* UncompressSuccinct over a range: decode back into the raw arena, segments become not compacted */
adacom_status adacom_uncompact(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags);

/* fused decode + SUM (+COUNT) over a range; result on the host after an internal sync */
adacom_status adacom_scan_sum(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags, adacom_agg *out);
/* same, asynchronous: the aggregate is left in device memory (32 bytes, adacom_agg layout) */
adacom_status adacom_scan_sum_async(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                    adacom_agg *device_out);
/* Allocation-free fused decode + SUM/COUNT of ONE packed payload.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The call takes no context and allocates nothing, on the device or the host side:
 * no cudaMalloc / cudaFree / stream-ordered allocation, no allocating library routine.  Its only working memory is a
 * file-scope `static __device__` array of the library (kernels_noalloc.cu: 524,544 bytes, resident on every GPU that
 * loads the library until the process exits): one 32-byte partial per 4096 values plus a ticket word, which is why
 * count is limited to ADACOM_NOALLOC_MAX_VALUES -- a larger count is rejected with ADACOM_ERR_INVALID (split the call).
 *   device_words  the payload on the device: count fields of `width` bits, LSB first in 64-bit words (the packed arena
 *                 layout: adacom_column_packed_device_ptr() + 8 * meta.word_offset), ceil(count * width / 64) words
 *   min_factor    the frame of reference added to every field (adacom_segment_meta.min_factor; 0 for none)
 *   stream        a cudaStream_t of the CURRENT device (NULL = the default stream); the caller keeps that device current
 *   device_out    one adacom_agg in device-accessible memory (device or pinned host), overwritten on the stream:
 *                 sum over i of T(min_factor + field_i) sign-/zero-extended to 128 bits, count, reserved = 0
 * Same answer as adacom_scan_sum over the segment the payload belongs to.  Every call on a device shares the array
 * and starts by clearing its ticket on `stream`; calls that overlap on the GPU (different streams or threads) race on
 * it -- keep them on one stream or otherwise ordered.  The first call on a device looks the array's device address
 * up (cudaGetSymbolAddress); make that call outside a graph capture.  Errors: text in adacom_last_error(NULL). */
#define ADACOM_NOALLOC_MAX_VALUES (1ull << 26)
adacom_status adacom_packed_sum_noalloc(const uint64_t *device_words, uint64_t count, uint32_t width, uint64_t min_factor,
                                        adacom_ptype type, void *stream, adacom_agg *device_out);
/* fused decode + (value <op> constant) + COUNT/SUM of the matching rows; optional selection bitmap
 * (1 bit per row of the range, bit r - row_start(seg_begin); device pointer, zeroed by the call). */
adacom_status adacom_scan_filter(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags, adacom_cmp op,
                                 int64_t constant, adacom_agg *out, uint64_t *device_bitmap);
/* NULL semantics of the two fused single-column calls above: they take no validity mask and treat every slot as a
 * value -- a NULL row (whose slot holds NullValue<T>() = numeric_limits<T>::min(), succinct.cpp:288-292) would enter
 * SUM / COUNT.  For nullable columns, IS [NOT] NULL, OR-conjunctions and selection refinement use adacom_scan_select /
 * adacom_scan_sum_valid below. */

/* ---- table filter on one column: ColumnSegment::FilterSelection (column_segment.cpp:575-683, 685-844) -------------- */
typedef enum {
	ADACOM_PRED_COMPARE = 0,     /* TableFilterType::CONSTANT_COMPARISON: value <op> constant, NULL rows never pass */
	ADACOM_PRED_IS_NULL = 1,     /* TableFilterType::IS_NULL (TemplatedNullSelection<true>) */
	ADACOM_PRED_IS_NOT_NULL = 2  /* TableFilterType::IS_NOT_NULL */
} adacom_pred_kind;
typedef enum {
	ADACOM_CONJ_AND = 0,         /* TableFilterType::CONJUNCTION_AND (column_segment.cpp:717-723) */
	ADACOM_CONJ_OR = 1           /* TableFilterType::CONJUNCTION_OR  (column_segment.cpp:688-716) */
} adacom_conj;
typedef struct {
	int32_t kind;                /* adacom_pred_kind */
	int32_t op;                  /* adacom_cmp (ADACOM_PRED_COMPARE only) */
	int64_t constant;            /* compared as the column's type T */
} adacom_pred;
#define ADACOM_MAX_PREDS 8
/* Fused decode + FilterSelection + aggregate over [seg_begin, seg_end):
 *   selected(row) = conj over preds[0 .. n_preds) (n_preds == 0: every row) AND bit(row) of device_bitmap_in (the
 *                   incoming selection vector `sel`; NULL = all rows)
 *   validity: DuckDB validity words for the WHOLE column (bit i = column row i valid; same convention as
 *             adacom_analyze), device or host pointer per validity_on_device; NULL = all valid (mask.AllValid()).
 *   out->count    = selected rows (approved_tuple_count), out->reserved = selected rows that are valid (COUNT(col)),
 *   out->sum      = SUM over the selected valid rows (DuckDB's aggregates skip NULLs),
 *   device_bitmap_out (optional, zeroed by the call): the new selection, bit (row - row_start(seg_begin)).
 * Nested filters (AND of ORs) chain calls through the bitmaps, like the recursion of FilterSelection.
 * A conjunction that amounts to ONE value range over the valid rows (or the complement of one: any single comparison,
 * BETWEEN, an OR of two half-lines, with or without IS NOT NULL) runs on the field-domain filter kernel with the
 * validity / selection words AND-ed in; IS NULL children and conjunctions that leave two separate ranges decode per
 * value.  Same results either way. */
adacom_status adacom_scan_select(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                 const adacom_pred *preds, uint32_t n_preds, adacom_conj conj, const uint64_t *validity,
                                 int validity_on_device, const uint64_t *device_bitmap_in, uint64_t *device_bitmap_out,
                                 adacom_agg *out);
/* NULL-aware fused decode + SUM: sum / reserved over the valid rows, count = all rows of the range */
adacom_status adacom_scan_sum_valid(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                    const uint64_t *validity, int validity_on_device, adacom_agg *out);

/* batch of point filters (the Zipf look-up benchmarks): query q scans segment seg_idx[q] for
 * value == constants[q]; out_count[q] = matches, out_first_row[q] = first matching row in the segment
 * (UINT64_MAX when none).  All four arrays are host pointers. */
adacom_status adacom_point_filter_batch(adacom_column *col, uint64_t n_queries, const uint32_t *seg_idx,
                                        const int64_t *constants, int flags, uint64_t *out_count,
                                        uint64_t *out_first_row);

/* ---- fused multi-column scan --------------------------------------------------------------------------- */
/* What DuckDB runs above the codec for a filter + aggregate query (RowGroup::TemplatedScan,
 * src/storage/table/row_group.cpp:391-447: scan every referenced column vector by vector, refine the selection
 * filter by filter -- ColumnSegment::FilterSelection, column_segment.cpp:575-658, AND conjunctions :685-844 --
 * then aggregate the survivors), as ONE pass over the packed columns.  All columns hold the same table rows
 * (their segment boundaries differ: capacities are 262136/S values).
 *   selection = AND over the terms with has_predicate of (lo <= value <= hi), compared as the column's type T
 *               (>=, >, <=, <, = on integers are all inclusive ranges), AND the optional input bitmap;
 *   out[g]    = {SUM over the selected rows of prod_k (add[k] + mul[k] * value(term[k])) mod 2^128, COUNT}.
 * TPC-H Q6 = terms {shipdate in [a, b-1], discount in [5, 7], quantity <= 23, extendedprice} and one sum
 * {extendedprice * discount}.  Bitmaps: 1 bit per row, bit (row - row_begin), device pointers, row_begin a
 * multiple of 2048.  ADACOM_ERR_UNSUPPORTED when a block of rows would span more than 3 segments of a column. */
typedef struct {
	adacom_column *col;
	int32_t has_predicate; /* 0: the column is only read by the sums */
	int32_t reserved;
	int64_t lo, hi;        /* inclusive; for unsigned T the bit patterns are compared unsigned */
} adacom_scan_term;
typedef struct {
	uint32_t n_factors;    /* 0..3; 0 = COUNT(*) (sum of 1) */
	uint32_t term[3];      /* index into terms[] */
	int64_t add[3];
	int64_t mul[3];
} adacom_sum_expr;
adacom_status adacom_scan_columns(adacom_ctx *ctx, const adacom_scan_term *terms, uint32_t n_terms,
                                  const adacom_sum_expr *sums, uint32_t n_sums, uint64_t row_begin, uint64_t row_end,
                                  int flags, const uint64_t *device_bitmap_in, uint64_t *device_bitmap_out,
                                  adacom_agg *out);
/* same, asynchronous: n_sums aggregates are left in device memory */
adacom_status adacom_scan_columns_async(adacom_ctx *ctx, const adacom_scan_term *terms, uint32_t n_terms,
                                        const adacom_sum_expr *sums, uint32_t n_sums, uint64_t row_begin,
                                        uint64_t row_end, int flags, const uint64_t *device_bitmap_in,
                                        uint64_t *device_bitmap_out, adacom_agg *device_out);
/* GROUP BY a small-domain column in the same pass (TPC-H Q1: extension/tpch/dbgen/queries/q01.sql): the decoded
 * value of terms[group_term] (0 .. n_groups-1, n_groups <= 8; rows with any other value are dropped) selects the
 * group; out[g * n_sums + s] = aggregate s of group g.  out: host pointer, or device pointer when out_on_device
 * (then the call is asynchronous). */
adacom_status adacom_scan_columns_grouped(adacom_ctx *ctx, const adacom_scan_term *terms, uint32_t n_terms,
                                          uint32_t group_term, uint32_t n_groups, const adacom_sum_expr *sums,
                                          uint32_t n_sums, uint64_t row_begin, uint64_t row_end, int flags,
                                          const uint64_t *device_bitmap_in, uint64_t *device_bitmap_out,
                                          adacom_agg *out, int out_on_device);
/* ---- fetch_row ------------------------------------------------------------------------------------------ */
/* batched random access: out[q] = T(vec(seg_idx[q])[row_in_seg[q]] + min).  Host pointers unless
 * on_device (then all three are device pointers). */
adacom_status adacom_fetch_rows(adacom_column *col, uint64_t n, const uint32_t *seg_idx, const uint32_t *row_in_seg,
                                int flags, void *out, int on_device);

/* ---- checkpoint image (persisted form of a column) ------------------------------------ */
/* The reference cannot checkpoint succinct segments (column_segment.cpp:529-533; init_compression / compress /
 * compress_finalize are stubs, succinct.cpp:86-119).  This gives those slots a real layout, the one DuckDB uses for
 * every other codec: ColumnCheckpointState::FlushSegment (src/storage/table/column_checkpoint_state.cpp:91-163) puts
 * each segment payload into a 256 KiB block, small payloads share blocks under the PartialBlockManager policy
 * (src/storage/partial_block_manager.cpp:14-77: payloads <= 80 % of a block go to the partially filled block with the
 * least sufficient space left, offsets 8-byte aligned), every block carries Checksum(payload)
 * (src/common/checksum.cpp:6-25) in its first 8 bytes (FileBuffer::ChecksumAndWrite, file_buffer.cpp:113-119), and a
 * DataPointer per segment (src/include/duckdb/storage/data_pointer.hpp:20-27) records {row_start, tuple_count,
 * block_id, offset, compression_type} -- extended here by the ColumnSegment members the succinct scan needs
 * (width, min_factor, max_factor, compacted/repacked).  The payload of a segment is its int_vector data words (the
 * 9-byte serialize header is recomputable from tuple_count and width) or, for a segment that was not re-packed, its
 * tuple_count slot-width values.
 *
 * image = adacom_image_header | n_segments x adacom_data_pointer | zero padding to blocks_offset |
 *         n_blocks x { uint64 checksum, 262136 payload bytes }          (all little endian)
 * The blocks are assembled and checksummed on the GPU in one pass (checkpoint_blocks_kernel); restore verifies every
 * checksum while it scatters the payloads back into HBM (restore_blocks_kernel). */
#define ADACOM_IMAGE_MAGIC "ADACOMCK"
#define ADACOM_IMAGE_VERSION 1u
#define ADACOM_BLOCK_ALLOC_SIZE 262144u /* Storage::BLOCK_ALLOC_SIZE (constants.hpp:101-106) */
#define ADACOM_BLOCK_SIZE 262136u       /* Storage::BLOCK_SIZE = BLOCK_ALLOC_SIZE - BLOCK_HEADER_SIZE */
#define ADACOM_INVALID_BLOCK (-1)       /* INVALID_BLOCK: empty segment, nothing written */

typedef struct {
	char magic[8];            /* ADACOM_IMAGE_MAGIC */
	uint32_t version;
	uint32_t ptype;           /* adacom_ptype */
	uint64_t n_segments;
	uint64_t n_blocks;
	uint64_t n_values;
	uint64_t blocks_offset;   /* byte offset of block 0 (multiple of 4096) */
	uint64_t total_bytes;     /* blocks_offset + n_blocks * ADACOM_BLOCK_ALLOC_SIZE */
	uint64_t meta_checksum;   /* Checksum() over the header with this field zero, then the pointer table */
} adacom_image_header;        /* 64 bytes */

typedef struct {
	uint64_t row_start;       /* DataPointer::row_start */
	uint64_t tuple_count;     /* DataPointer::tuple_count */
	int64_t block_id;         /* BlockPointer::block_id, ADACOM_INVALID_BLOCK when tuple_count == 0 */
	uint32_t offset;          /* BlockPointer::offset inside the block payload (multiple of 8) */
	uint32_t payload_bytes;   /* 8 * ceil(count * width / 64) when re-packed, else count * type_size */
	uint64_t min_factor;
	uint64_t max_factor;
	uint32_t capacity;
	uint8_t compression_type; /* CompressionType: 10 = COMPRESSION_SUCCINCT, 1 = COMPRESSION_UNCOMPRESSED */
	uint8_t width;
	uint8_t state;            /* ADACOM_SEG_* */
	uint8_t type_size;
	uint64_t reserved;
} adacom_data_pointer;        /* 64 bytes */

/* size of the image the column would be checkpointed to right now */
adacom_status adacom_column_checkpoint_size(adacom_column *col, uint64_t *bytes_out);
/* write the image to `out` (host pointer, or device pointer when out_on_device; capacity in bytes);
 * ADACOM_ERR_INVALID when capacity is too small.  The column is unchanged. */
adacom_status adacom_column_checkpoint(adacom_column *col, void *out, uint64_t capacity, int out_on_device,
                                       uint64_t *bytes_out);
/* rebuild a column from an image (host or device pointer): every block checksum and the metadata checksum are
 * verified, ADACOM_ERR_CORRUPT otherwise.  Re-packed payloads land in one packed extent, slot-width payloads in the raw
 * arena; the column answers scans exactly like the one that was checkpointed. */
adacom_status adacom_column_restore(adacom_ctx *ctx, const void *image, uint64_t bytes, int image_on_device,
                                    adacom_column **out);

/* ---- neighbour codecs on the same GPU ------------------------------------------------------- */
/* DuckDB's BitpackingFun and RLEFun segments, decoded from the block images DuckDB itself writes at checkpoint time
 * (the `compress` slots of src/storage/compression/bitpacking.cpp:350-528 and rle.cpp:117-245):
 *   bitpacking: [u64 end of metadata | group payloads | u32 metadata entries (mode << 24 | offset)], 2048-value groups in
 *               CONSTANT / CONSTANT_DELTA / DELTA_FOR / FOR mode, payload packed by third_party/fastpforlib;
 *   RLE:        [u64 offset of the run lengths | T values[runs] | u16 lengths[runs]].
 * Entry points replace BitpackingScan / BitpackingScanPartial (bitpacking.cpp:729-816), BitpackingFetchRow (:823-866),
 * RLEScan / RLEScanPartial (rle.cpp:298-322) and RLEFetchRow (:327-341); the fused SUM replaces scan + aggregate.
 * Validity lives in a separate column in DuckDB, so every slot decodes to whatever the compressor stored. */
typedef enum { ADACOM_CODEC_BITPACKING = 0, ADACOM_CODEC_RLE = 1 } adacom_codec;
typedef struct adacom_codec_column adacom_codec_column;
/* n_segments block images (host pointers, segment_bytes[g] bytes each, as handed to ColumnCheckpointState::FlushSegment)
 * holding tuple_counts[g] values each; the images are copied into one device arena. */
adacom_status adacom_codec_column_create(adacom_ctx *ctx, adacom_codec codec, adacom_ptype type, uint64_t n_segments,
                                         const uint8_t *const *segment_images, const uint64_t *segment_bytes,
                                         const uint64_t *tuple_counts, adacom_codec_column **out);
void adacom_codec_column_destroy(adacom_codec_column *col);
uint64_t adacom_codec_column_num_values(const adacom_codec_column *col);
uint64_t adacom_codec_column_image_bytes(const adacom_codec_column *col);
/* batched scan of [seg_begin, seg_end) into a flat T array (device pointer when out_on_device, then asynchronous) */
adacom_status adacom_codec_decode(adacom_codec_column *col, uint64_t seg_begin, uint64_t seg_end, void *out,
                                  int out_on_device);
/* fused decode + SUM/COUNT (RLE: sum of value * run length, the rows are never expanded) */
adacom_status adacom_codec_scan_sum(adacom_codec_column *col, uint64_t seg_begin, uint64_t seg_end, adacom_agg *out);
/* batched fetch_row: out[q] = value(seg_idx[q], row_in_seg[q]); host pointers unless on_device */
adacom_status adacom_codec_fetch_rows(adacom_codec_column *col, uint64_t n, const uint32_t *seg_idx,
                                      const uint32_t *row_in_seg, void *out, int on_device);

/* ---- host-buffer convenience paths (what bench.py's e2e leg times) ------------------------------------ */
/* host T values -> analyze + encode on the GPU -> metas (+ packed words when host_words_out != NULL,
 * capacity in words; *total_words_out receives the arena size). The column stays resident in *col_out. */
adacom_status adacom_compress_host(adacom_ctx *ctx, adacom_ptype type, const void *host_values, uint64_t n_segments,
                                   const uint64_t *seg_counts, uint32_t seg_capacity, int flags,
                                   adacom_segment_meta *metas_out, uint64_t *host_words_out,
                                   uint64_t host_words_capacity, uint64_t *total_words_out,
                                   adacom_column **col_out);

#ifdef __cplusplus
}
#endif
#endif /* ADACOM_GPU_H */
