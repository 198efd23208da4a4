/*
 * succinct_oracle.h -- CPU restatement (L1) of AdaCom's succinct column-segment codec.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and only
 * as the checker or the timed CPU baseline.  The product (duckdb_adaptive_compression_b200/csrc)
 * never links or calls this code.
 *
 * Parity status: PINNED.  This restatement is cross-checked (tests/test_oracle_vs_ref.py and
 * tests/golden/make_golden.py) against oracle/_ref/libadacom_ref.so, which is compiled from the
 * reference's own vendored SDSL sources under /root/reference/third_party/{sdsl,sdsl-lite} and
 * calls the same SDSL functions in the same order as the reference's codec loops.  Golden digests
 * minted by that library are committed in tests/golden/.
 *
 * Every function cites the reference file:line it follows (paths relative to /root/reference).
 */
#ifndef SUCCINCT_ORACLE_H
#define SUCCINCT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- SDSL bit primitives: third_party/sdsl/include/sdsl/bits.hpp ---- */
uint32_t orc_bits_hi(uint64_t x);                                                   /* bits.hpp:392-414 */
uint64_t orc_lo_set(uint32_t len);                                                  /* bits.hpp:84-87   */
void orc_write_int(uint64_t *word, uint64_t x, uint8_t offset, uint8_t len);        /* bits.hpp:456-476 */
void orc_write_int_and_move(uint64_t **word, uint64_t x, uint8_t *offset, uint8_t len); /* :478-499 */
uint64_t orc_read_int(const uint64_t *word, uint8_t offset, uint8_t len);           /* bits.hpp:501-511 */
uint64_t orc_read_int_and_move(const uint64_t **word, uint8_t *offset, uint8_t len); /* :513-529 */

/* ---- sdsl::int_vector<0>: int_vector.hpp:325-327 (fields), memory_management.hpp:344-375 ---- */
typedef struct {
	uint64_t m_size;  /* number of BITS */
	uint64_t *m_data; /* 64-bit words, little endian bit order */
	uint8_t m_width;  /* bits per element */
} orc_int_vector;

void orc_iv_init(orc_int_vector *v);                      /* empty vector, width 64 (int_vector.hpp:1240ff) */
void orc_iv_free(orc_int_vector *v);
void orc_iv_set_width(orc_int_vector *v, uint8_t w);      /* int_vector.hpp:147-155, 469-472 */
void orc_iv_bit_resize(orc_int_vector *v, uint64_t bits); /* memory_management.hpp:344-375 */
void orc_iv_resize(orc_int_vector *v, uint64_t n);        /* int_vector.hpp:377-380 */
uint64_t orc_iv_size(const orc_int_vector *v);            /* int_vector.hpp:390-393 */
uint64_t orc_iv_capacity(const orc_int_vector *v);        /* int_vector.hpp:416-419 */
uint64_t orc_iv_get(const orc_int_vector *v, uint64_t i); /* int_vector.hpp:1359-1363 + bits::read_int */
void orc_iv_set(orc_int_vector *v, uint64_t i, uint64_t x);
uint64_t orc_iv_size_in_bytes(const orc_int_vector *v);   /* io.hpp:636-640 + int_vector.hpp:1546-1578 */
/* serialise exactly like int_vector<0>::serialize: 8-byte bit count, 1-byte width, capacity/64 words */
uint64_t orc_iv_serialize(const orc_int_vector *v, uint8_t *out);

/* ---- ColumnSegment stand-in: src/include/duckdb/storage/table/column_segment.hpp:60,139-214 ---- */
enum { ORC_FUNC_UNCOMPRESSED = 1, ORC_FUNC_SUCCINCT = 10 }; /* compression_type.hpp:16-29 */

typedef struct {
	orc_int_vector succinct_vec;
	uint8_t *block;        /* flat block when function == UNCOMPRESSED (segment_size bytes) */
	uint64_t segment_size; /* bytes; Storage::BLOCK_SIZE = 262136 unless first-of-collection */
	uint64_t count;        /* tuples appended (SegmentBase::count) */
	uint64_t num_elements; /* column_segment.cpp:263 */
	uint64_t min_factor;   /* starts UINT64_MAX (column_segment.cpp:96) */
	uint64_t max_factor;   /* starts 0 */
	uint8_t type_size;     /* 1,2,4,8 */
	uint8_t is_signed;
	uint8_t compacted;
	uint8_t repacked;      /* oracle bookkeeping: the width actually shrank (for the intended-mode decode) */
	int function;          /* ORC_FUNC_* */
} orc_segment;

/* flags */
enum {
	ORC_PADDED = 1,        /* DBConfig::succinct_padded_to_next_byte_enabled (config.hpp:189-197) */
	ORC_LITERAL = 2,       /* reproduce reference quirks Q1/Q2 literally instead of intended semantics */
};

/* ColumnSegment ctor for a SUCCINCT transient segment: column_segment.cpp:88-110 */
void orc_segment_init_succinct(orc_segment *s, uint8_t type_size, int is_signed, uint64_t segment_size);
/* ColumnSegment ctor for an UNCOMPRESSED transient segment (adaptive mode): column_segment.cpp:60-75 */
void orc_segment_init_uncompressed(orc_segment *s, uint8_t type_size, int is_signed, uint64_t segment_size);
void orc_segment_free(orc_segment *s);

/* SuccinctAppend<T> + SuccinctAppendLoop<T>: succinct.cpp:271-322.  values: T[]; sel: optional u32
 * selection (NULL = identity); validity: optional DuckDB validity words (NULL = all valid). */
uint64_t orc_segment_append(orc_segment *s, const void *values, const uint32_t *sel, const uint64_t *validity,
                            uint64_t offset, uint64_t count);
/* plain memcpy append into the flat block (FixedSizeAppend; used before the adaptive encoder) */
uint64_t orc_segment_append_uncompressed(orc_segment *s, const void *values, uint64_t count);

/* ColumnSegment::BitCompressFromSuccinct: column_segment.cpp:348-383 */
void orc_bit_compress_from_succinct(orc_segment *s, int flags);
/* ColumnSegment::BitCompressFromUncompressed (+ the resize in Compact): column_segment.cpp:304-316, 385-456 */
void orc_bit_compress_from_uncompressed(orc_segment *s, int flags);
/* SuccinctScanPartial<T>: succinct.cpp:123-144 */
void orc_scan_partial(const orc_segment *s, uint64_t start, uint64_t scan_count, void *result, uint64_t result_offset,
                      int flags);
/* ColumnSegment::UncompressSuccinct: column_segment.cpp:458-506 */
void orc_uncompress_succinct(orc_segment *s, int flags);
/* intended SuccinctFetchRow (== FixedSizeFetchRow semantics, fixed_size_uncompressed.cpp:177-187) */
void orc_fetch_row(const orc_segment *s, uint64_t row_id, void *result, uint64_t result_idx, int flags);
/* ColumnSegment::GetDataSize: column_segment.cpp:204-214 */
uint64_t orc_segment_data_size(const orc_segment *s);

/* ---- whole-column helpers (oracle-side conveniences used by tests and the CPU baseline) ---- */
typedef struct {
	uint64_t lo, hi; /* 128-bit two's complement sum of the decoded T values (DuckDB SUM -> HUGEINT) */
	uint64_t count;
} orc_sum128;

/* compare ops of TemplatedFilterSelection: column_segment.cpp:593-655 */
enum { ORC_CMP_EQ = 0, ORC_CMP_NE = 1, ORC_CMP_LT = 2, ORC_CMP_GT = 3, ORC_CMP_LE = 4, ORC_CMP_GE = 5 };

/* decode the whole segment (vector by vector, like ColumnData::ScanVector drives it) and sum */
void orc_segment_sum(const orc_segment *s, int flags, orc_sum128 *out);
/* decode + compare against a constant of type T (passed as the 64-bit sign/zero-extended value);
 * writes match count, the 128-bit sum of the matching values and (optional) a 1-bit-per-row bitmap */
void orc_segment_filter(const orc_segment *s, int flags, int op, uint64_t constant, orc_sum128 *out,
                        uint64_t *bitmap_words);

/* general table filter on one column (ColumnSegment::FilterSelection, column_segment.cpp:685-844) + NULL-skipping
 * aggregate: conjunction (AND / OR) of comparisons, IS NULL, IS NOT NULL over an incoming selection.
 * validity / bitmaps: bit (row0 + segment row); NULL pointers = all valid / all selected / no output.
 * out->count = selected rows, *valid_count = selected valid rows, out->lo/hi = SUM over the selected valid rows */
enum { ORC_PRED_COMPARE = 0, ORC_PRED_IS_NULL = 1, ORC_PRED_IS_NOT_NULL = 2 };
typedef struct {
	int32_t kind, op;
	int64_t constant;
} orc_pred;
void orc_segment_select(const orc_segment *s, int flags, const orc_pred *preds, uint32_t n_preds, int conj_or,
                        const uint64_t *validity, uint64_t validity_row0, const uint64_t *bitmap_in,
                        uint64_t *bitmap_out, uint64_t bitmap_row0, orc_sum128 *out, uint64_t *valid_count);

/* ---- timed CPU baseline: one task per segment, T pthreads, segments dealt round-robin ---- */
typedef struct {
	double analyze_s, encode_s, scan_s, sum_s, uncompact_s; /* wall seconds, max over threads */
	uint64_t values;
	orc_sum128 sum;
	uint64_t packed_bytes;
} orc_bench_result;

/* values: n T-values; segments of seg_capacity values each (last one short).  Runs append(analyze),
 * BitCompressFromSuccinct, a full SuccinctScanPartial scan in 2048-value vectors, scan+sum and
 * UncompressSuccinct.
 * copy_per_call != 0 reproduces quirk Q4 (deep copy of the packed segment per scan call). */
int orc_bench_column(const void *values, uint64_t n, uint8_t type_size, int is_signed, uint64_t seg_capacity,
                     int threads, int copy_per_call, orc_bench_result *out);

/* ---- opaque-handle mirror (same shape as oracle/_ref's ref_* entry points) ---- */
void *orc_segment_new_succinct(uint8_t type_size, int is_signed, uint64_t segment_size);
void *orc_segment_new_uncompressed(uint8_t type_size, int is_signed, uint64_t segment_size);
void orc_segment_delete(void *h);
/* field: 0 count, 1 width, 2 bit_size, 3 min_factor, 4 max_factor, 5 compacted, 6 repacked,
 *        7 function, 8 payload words, 9 num_elements */
uint64_t orc_segment_get(void *h, int field);
const uint64_t *orc_segment_words(void *h);
const uint8_t *orc_segment_block(void *h);
uint64_t orc_segment_serialize(void *h, uint8_t *out, uint64_t out_capacity);

#ifdef __cplusplus
}
#endif
#endif
