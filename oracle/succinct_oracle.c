/*
 * succinct_oracle.c -- CPU restatement (L1) of AdaCom's succinct column-segment codec.
 * TEST INFRASTRUCTURE ONLY -- see succinct_oracle.h for the rules and the parity status (PINNED
 * against oracle/_ref, the reference's own SDSL code compiled from /root/reference).
 *
 * File:line citations are relative to /root/reference.
 */
#include "succinct_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------------------------------- */
/* SDSL bit primitives                                                                          */
/* ------------------------------------------------------------------------------------------- */

/* bits::lo_set[len]: the `len` low bits set, lo_set[64] = all ones
 * (third_party/sdsl/include/sdsl/bits.hpp:84-87, table in third_party/sdsl-lite/lib/bits.cpp). */
uint64_t orc_lo_set(uint32_t len) {
	return len >= 64 ? ~0ULL : ((1ULL << len) - 1ULL);
}

/* bits::hi: index of the most significant set bit, hi(0) == 0 (bits.hpp:392-414). */
uint32_t orc_bits_hi(uint64_t x) {
	return x == 0 ? 0u : (uint32_t)(63 - __builtin_clzll(x));
}

/* bits::write_int (bits.hpp:456-476): store the low `len` bits of x at bit `offset` of *word,
 * spilling into word[1] when the field straddles; bits outside the field are preserved. */
void orc_write_int(uint64_t *word, uint64_t x, uint8_t offset, uint8_t len) {
	x &= orc_lo_set(len);
	unsigned end = (unsigned)offset + (unsigned)len;
	if (end < 64) {
		uint64_t keep = (~0ULL << end) | orc_lo_set(offset);
		word[0] = (word[0] & keep) | (x << offset);
	} else {
		word[0] = (word[0] & orc_lo_set(offset)) | (x << offset);
		unsigned spill = end & 0x3F;
		if (spill) {
			word[1] = (word[1] & ~orc_lo_set(spill)) | (x >> (len - spill));
		}
	}
}

/* bits::write_int_and_move (bits.hpp:478-499): as write_int, then advance the cursor by len bits. */
void orc_write_int_and_move(uint64_t **word, uint64_t x, uint8_t *offset, uint8_t len) {
	x &= orc_lo_set(len);
	unsigned off = *offset;
	unsigned end = off + (unsigned)len;
	uint64_t *w = *word;
	if (end < 64) {
		uint64_t keep = (~0ULL << end) | orc_lo_set(off);
		w[0] = (w[0] & keep) | (x << off);
		*offset = (uint8_t)end;
	} else {
		w[0] = (w[0] & orc_lo_set(off)) | (x << off);
		if (end > 64) {
			unsigned spill = end & 0x3F;
			++w;
			w[0] = (w[0] & ~orc_lo_set(spill)) | (x >> (len - spill));
			*offset = (uint8_t)spill;
		} else {
			++w;
			*offset = 0;
		}
		*word = w;
	}
}

/* bits::read_int (bits.hpp:501-511). */
uint64_t orc_read_int(const uint64_t *word, uint8_t offset, uint8_t len) {
	uint64_t lo = word[0] >> offset;
	unsigned end = (unsigned)offset + (unsigned)len;
	if (end > 64) {
		return lo | ((word[1] & orc_lo_set(end & 0x3F)) << (64 - offset));
	}
	return lo & orc_lo_set(len);
}

/* bits::read_int_and_move (bits.hpp:513-529). */
uint64_t orc_read_int_and_move(const uint64_t **word, uint8_t *offset, uint8_t len) {
	const uint64_t *w = *word;
	unsigned off = *offset;
	uint64_t lo = w[0] >> off;
	unsigned end = off + (unsigned)len;
	if (end >= 64) {
		*word = w + 1;
		*offset = (uint8_t)(end & 0x3F);
		if (end == 64) {
			return lo;
		}
		unsigned spill = end & 0x3F;
		return lo | ((w[1] & orc_lo_set(spill)) << (len - spill));
	}
	*offset = (uint8_t)end;
	return lo & orc_lo_set(len);
}

/* ------------------------------------------------------------------------------------------- */
/* sdsl::int_vector<0>                                                                          */
/* ------------------------------------------------------------------------------------------- */

void orc_iv_init(orc_int_vector *v) {
	/* int_vector(size=0, default=0, intWidth=0): width(0) clamps to 64 (int_vector.hpp:147-155,
	 * 1250-1256); resize(0) still allocates the 8-byte pad word (memory_management.hpp:351-356). */
	v->m_size = 0;
	v->m_data = NULL;
	v->m_width = 64;
	orc_iv_bit_resize(v, 0);
}

void orc_iv_free(orc_int_vector *v) {
	free(v->m_data);
	v->m_data = NULL;
	v->m_size = 0;
}

void orc_iv_set_width(orc_int_vector *v, uint8_t w) {
	v->m_width = (w > 0 && w <= 64) ? w : 64;
}

uint64_t orc_iv_capacity(const orc_int_vector *v) {
	return ((v->m_size + 63) >> 6) << 6;
}

/* memory_manager::resize (memory_management.hpp:344-375): realloc to ((bits+64)>>6)<<3 bytes when
 * the word count changes (or nothing is allocated yet), zero the bits between bit_size and
 * capacity, and zero the pad word when bits % 64 == 0. */
void orc_iv_bit_resize(orc_int_vector *v, uint64_t bits) {
	uint64_t old_bytes = ((v->m_size + 63) >> 6) << 3;
	uint64_t new_bytes = ((bits + 63) >> 6) << 3;
	int do_realloc = old_bytes != new_bytes;
	v->m_size = bits;
	if (do_realloc || v->m_data == NULL) {
		size_t alloc = (size_t)(((bits + 64) >> 6) << 3);
		uint64_t *p = (uint64_t *)realloc(v->m_data, alloc);
		if (!p && alloc) {
			abort();
		}
		/* realloc leaves grown memory undefined in the reference; the oracle zero-fills the
		 * grown part so runs are deterministic (the codec never reads slots >= count). */
		if (alloc > old_bytes) {
			uint64_t from = v->m_data == NULL ? 0 : old_bytes;
			if (v->m_data == NULL) {
				from = 0;
			}
			memset((uint8_t *)p + from, 0, alloc - from);
		}
		v->m_data = p;
		uint64_t cap = orc_iv_capacity(v);
		if (v->m_size < cap) {
			uint8_t len = (uint8_t)(cap - v->m_size);
			uint8_t in_word = (uint8_t)(v->m_size & 0x3F);
			orc_write_int(v->m_data + (v->m_size >> 6), 0, in_word, len);
		}
		if ((v->m_size % 64) == 0) {
			v->m_data[v->m_size / 64] = 0;
		}
	}
}

void orc_iv_resize(orc_int_vector *v, uint64_t n) {
	orc_iv_bit_resize(v, n * v->m_width);
}

uint64_t orc_iv_size(const orc_int_vector *v) {
	return v->m_size / v->m_width;
}

uint64_t orc_iv_get(const orc_int_vector *v, uint64_t i) {
	uint64_t bit = i * v->m_width;
	return orc_read_int(v->m_data + (bit >> 6), (uint8_t)(bit & 0x3F), v->m_width);
}

void orc_iv_set(orc_int_vector *v, uint64_t i, uint64_t x) {
	uint64_t bit = i * v->m_width;
	orc_write_int(v->m_data + (bit >> 6), x, (uint8_t)(bit & 0x3F), v->m_width);
}

/* size_in_bytes = serialize into a null stream (io.hpp:636-640): header (8-byte m_size + 1-byte
 * width for int_vector<0>, int_vector.hpp:602-610) + capacity/8 data bytes (:1546-1578). */
uint64_t orc_iv_size_in_bytes(const orc_int_vector *v) {
	return 9 + (orc_iv_capacity(v) >> 3);
}

uint64_t orc_iv_serialize(const orc_int_vector *v, uint8_t *out) {
	uint64_t data_bytes = orc_iv_capacity(v) >> 3;
	memcpy(out, &v->m_size, 8);
	out[8] = v->m_width;
	memcpy(out + 9, v->m_data, data_bytes);
	return 9 + data_bytes;
}

/* ------------------------------------------------------------------------------------------- */
/* ColumnSegment stand-in                                                                       */
/* ------------------------------------------------------------------------------------------- */

/* load element i of a T[] array as uint64_t(T): sign-extends signed T (succinct.cpp:286-287). */
static uint64_t load_as_u64(const void *values, uint64_t i, uint8_t type_size, int is_signed) {
	switch (type_size) {
	case 1:
		return is_signed ? (uint64_t)(int64_t)((const int8_t *)values)[i] : (uint64_t)((const uint8_t *)values)[i];
	case 2:
		return is_signed ? (uint64_t)(int64_t)((const int16_t *)values)[i] : (uint64_t)((const uint16_t *)values)[i];
	case 4:
		return is_signed ? (uint64_t)(int64_t)((const int32_t *)values)[i] : (uint64_t)((const uint32_t *)values)[i];
	default:
		return ((const uint64_t *)values)[i];
	}
}

/* NullValue<T>() == numeric_limits<T>::min(), converted to uint64_t
 * (src/include/duckdb/common/types/null_value.hpp:25-28). */
static uint64_t null_value_u64(uint8_t type_size, int is_signed) {
	if (!is_signed) {
		return 0;
	}
	switch (type_size) {
	case 1:
		return (uint64_t)(int64_t)INT8_MIN;
	case 2:
		return (uint64_t)(int64_t)INT16_MIN;
	case 4:
		return (uint64_t)(int64_t)INT32_MIN;
	default:
		return (uint64_t)INT64_MIN;
	}
}

static void segment_common_init(orc_segment *s, uint8_t type_size, int is_signed, uint64_t segment_size) {
	memset(s, 0, sizeof(*s));
	s->segment_size = segment_size;
	s->type_size = type_size;
	s->is_signed = (uint8_t)(is_signed != 0);
	s->min_factor = UINT64_MAX; /* column_segment.cpp:96 */
	s->max_factor = 0;
	orc_iv_init(&s->succinct_vec);
}

/* column_segment.cpp:101-105: width(8*type_size); resize(segment_size / type_size). */
void orc_segment_init_succinct(orc_segment *s, uint8_t type_size, int is_signed, uint64_t segment_size) {
	segment_common_init(s, type_size, is_signed, segment_size);
	s->function = ORC_FUNC_SUCCINCT;
	orc_iv_set_width(&s->succinct_vec, (uint8_t)(type_size * 8));
	orc_iv_resize(&s->succinct_vec, segment_size / type_size);
}

/* column_segment.cpp:60-75: adaptive mode starts from a flat buffer-managed block. */
void orc_segment_init_uncompressed(orc_segment *s, uint8_t type_size, int is_signed, uint64_t segment_size) {
	segment_common_init(s, type_size, is_signed, segment_size);
	s->function = ORC_FUNC_UNCOMPRESSED;
	s->block = (uint8_t *)calloc(segment_size ? segment_size : 1, 1);
}

void orc_segment_free(orc_segment *s) {
	orc_iv_free(&s->succinct_vec);
	free(s->block);
	s->block = NULL;
}

/* SuccinctAppend<T> (succinct.cpp:308-322) around SuccinctAppendLoop<T> (succinct.cpp:271-306). */
uint64_t orc_segment_append(orc_segment *s, const void *values, const uint32_t *sel, const uint64_t *validity,
                            uint64_t offset, uint64_t count) {
	uint64_t max_tuples = s->segment_size / s->type_size;
	uint64_t copy_count = count < max_tuples - s->count ? count : max_tuples - s->count;
	uint64_t umin = UINT64_MAX, umax = 0;
	for (uint64_t i = 0; i < copy_count; i++) {
		uint64_t src = sel ? sel[offset + i] : offset + i;
		uint64_t dst = s->count + i;
		int valid = validity ? (int)((validity[src >> 6] >> (src & 63)) & 1) : 1;
		if (valid) {
			uint64_t v = load_as_u64(values, src, s->type_size, s->is_signed);
			orc_iv_set(&s->succinct_vec, dst, v);
			if (v < umin) {
				umin = v;
			}
			if (v > umax) {
				umax = v;
			}
		} else {
			orc_iv_set(&s->succinct_vec, dst, null_value_u64(s->type_size, s->is_signed));
		}
	}
	/* UpdateMinFactor / UpdateMaxFactor: column_segment.hpp:147-155 */
	if (umin < s->min_factor) {
		s->min_factor = umin;
	}
	if (umax > s->max_factor) {
		s->max_factor = umax;
	}
	s->count += copy_count;
	s->num_elements += count; /* quirk Q6: requested, not copied (column_segment.cpp:263) */
	return copy_count;
}

uint64_t orc_segment_append_uncompressed(orc_segment *s, const void *values, uint64_t count) {
	uint64_t max_tuples = s->segment_size / s->type_size;
	uint64_t copy_count = count < max_tuples - s->count ? count : max_tuples - s->count;
	memcpy(s->block + s->count * s->type_size, values, copy_count * s->type_size);
	s->count += copy_count;
	s->num_elements += count;
	return copy_count;
}

static uint8_t pick_width(uint64_t range, int flags) {
	unsigned w = orc_bits_hi(range) + 1;
	if (flags & ORC_PADDED) {
		w = (w + 7) & ~7u; /* column_segment.cpp:356-359 */
	}
	return (uint8_t)w;
}

/* ColumnSegment::BitCompressFromSuccinct (column_segment.cpp:348-383): in-place re-pack from the
 * slot width to hi(max-min)+1 bits with min subtracted, iff the width really shrinks. */
void orc_bit_compress_from_succinct(orc_segment *s, int flags) {
	uint64_t min = s->min_factor;
	uint64_t range = s->max_factor - s->min_factor;
	uint8_t new_w = pick_width(range, flags);
	uint8_t old_w = s->succinct_vec.m_width;
	if (old_w > new_w) {
		const uint64_t *rd = s->succinct_vec.m_data;
		uint64_t *wr = s->succinct_vec.m_data;
		uint8_t rd_off = 0, wr_off = 0;
		for (uint64_t i = 0; i < s->count; i++) {
			uint64_t x = orc_read_int_and_move(&rd, &rd_off, old_w);
			if (min != UINT64_MAX) {
				x -= min;
			}
			orc_write_int_and_move(&wr, x, &wr_off, new_w);
		}
		orc_iv_bit_resize(&s->succinct_vec, s->count * new_w);
		orc_iv_set_width(&s->succinct_vec, new_w);
		s->repacked = 1;
	}
	s->compacted = 1;
}

/* Compact() uncompressed branch (column_segment.cpp:304-316) + BitCompressFromUncompressed
 * (:385-456): zero-extended min/max (Q2b), local min never stored in the reference (Q2). */
void orc_bit_compress_from_uncompressed(orc_segment *s, int flags) {
	orc_iv_set_width(&s->succinct_vec, (uint8_t)(s->type_size * 8));
	orc_iv_resize(&s->succinct_vec, s->segment_size / s->type_size);

	uint64_t min = UINT64_MAX, max = 0;
	for (uint64_t i = 0; i < s->count; i++) {
		uint64_t cur = 0;
		memcpy(&cur, s->block + i * s->type_size, s->type_size);
		if (cur < min) {
			min = cur;
		}
		if (cur > max) {
			max = cur;
		}
	}
	uint64_t local_max = max;
	if (max != 0 && min != UINT64_MAX && max > min) {
		max -= min;
	}
	uint8_t new_w = pick_width(max, flags);
	uint8_t old_w = s->succinct_vec.m_width;
	if (old_w > new_w) {
		uint64_t *wr = s->succinct_vec.m_data;
		uint8_t wr_off = 0;
		for (uint64_t i = 0; i < s->count; i++) {
			int64_t x = 0;
			memcpy(&x, s->block + i * s->type_size, s->type_size);
			if (min != UINT64_MAX) {
				x = (int64_t)((uint64_t)x - min);
			}
			orc_write_int_and_move(&wr, (uint64_t)x, &wr_off, new_w);
		}
		orc_iv_bit_resize(&s->succinct_vec, s->count * new_w);
		orc_iv_set_width(&s->succinct_vec, new_w);
		s->repacked = 1;
	}
	if (!(flags & ORC_LITERAL)) {
		/* intended semantics (divergence Q2): remember the frame of reference */
		s->min_factor = min;
		s->max_factor = local_max;
	}
	s->function = ORC_FUNC_SUCCINCT;
	s->compacted = 1;
}

static int adds_min(const orc_segment *s, int flags) {
	if (s->min_factor == UINT64_MAX) {
		return 0; /* succinct.cpp:138 guard */
	}
	if (flags & ORC_LITERAL) {
		return 1; /* Q1: the reference adds min even when the segment was not re-packed */
	}
	return s->repacked;
}

static void store_low_bytes(void *result, uint64_t idx, uint64_t x, uint8_t type_size) {
	memcpy((uint8_t *)result + idx * type_size, &x, type_size); /* little endian: low S bytes */
}

/* SuccinctScanPartial<T> (succinct.cpp:123-144). */
void orc_scan_partial(const orc_segment *s, uint64_t start, uint64_t scan_count, void *result, uint64_t result_offset,
                      int flags) {
	int add = adds_min(s, flags);
	for (uint64_t i = 0; i < scan_count; i++) {
		uint64_t x = orc_iv_get(&s->succinct_vec, start + i);
		if (add) {
			x += s->min_factor;
		}
		store_low_bytes(result, result_offset + i, x, s->type_size);
	}
}

/* ColumnSegment::UncompressSuccinct (column_segment.cpp:458-506): literal mode adds min_factor
 * unconditionally (even UINT64_MAX, Q2); intended mode follows the scan rule. */
void orc_uncompress_succinct(orc_segment *s, int flags) {
	free(s->block);
	s->block = (uint8_t *)calloc(s->segment_size ? s->segment_size : 1, 1);
	uint64_t n = orc_iv_size(&s->succinct_vec);
	int add = (flags & ORC_LITERAL) ? 1 : adds_min(s, flags);
	for (uint64_t i = 0; i < n; i++) {
		uint64_t cur = orc_iv_get(&s->succinct_vec, i);
		if (add) {
			cur += s->min_factor;
		}
		store_low_bytes(s->block, i, cur, s->type_size);
	}
	s->function = ORC_FUNC_UNCOMPRESSED;
	s->compacted = 0;
	s->repacked = 0;
	orc_iv_resize(&s->succinct_vec, 0);
}

/* intended fetch_row: result[result_idx] = T(vec[row_id] + min). */
void orc_fetch_row(const orc_segment *s, uint64_t row_id, void *result, uint64_t result_idx, int flags) {
	uint64_t x = orc_iv_get(&s->succinct_vec, row_id);
	if (adds_min(s, flags)) {
		x += s->min_factor;
	}
	store_low_bytes(result, result_idx, x, s->type_size);
}

/* ColumnSegment::GetDataSize (column_segment.cpp:204-214) for a data segment. */
uint64_t orc_segment_data_size(const orc_segment *s) {
	if (s->function == ORC_FUNC_SUCCINCT) {
		return orc_iv_size_in_bytes(&s->succinct_vec);
	}
	return s->segment_size;
}

/* ------------------------------------------------------------------------------------------- */
/* Aggregates over decoded values (what DuckDB's SUM / filter run on the scan output)           */
/* ------------------------------------------------------------------------------------------- */

#define ORC_VECTOR_SIZE 2048 /* STANDARD_VECTOR_SIZE, src/include/duckdb/common/vector_size.hpp:17 */

static uint64_t extend_T(const uint8_t *p, uint8_t type_size, int is_signed) {
	return load_as_u64(p, 0, type_size, is_signed);
}

static void sum128_add(orc_sum128 *acc, uint64_t v, int is_signed) {
	uint64_t old = acc->lo;
	acc->lo += v;
	acc->hi += (acc->lo < old);
	if (is_signed && (int64_t)v < 0) {
		acc->hi += UINT64_MAX; /* sign extension of v into the high limb */
	}
	acc->count++;
}

void orc_segment_sum(const orc_segment *s, int flags, orc_sum128 *out) {
	uint8_t buf[ORC_VECTOR_SIZE * 8];
	memset(out, 0, sizeof(*out));
	for (uint64_t start = 0; start < s->count; start += ORC_VECTOR_SIZE) {
		uint64_t n = s->count - start < ORC_VECTOR_SIZE ? s->count - start : ORC_VECTOR_SIZE;
		orc_scan_partial(s, start, n, buf, 0, flags);
		for (uint64_t i = 0; i < n; i++) {
			sum128_add(out, extend_T(buf + i * s->type_size, s->type_size, s->is_signed), s->is_signed);
		}
	}
}

static int compare_T(uint64_t a, uint64_t b, int op, int is_signed) {
	/* a, b are sign/zero-extended T values; TemplatedFilterSelection ops, column_segment.cpp:593-655 */
	int lt = is_signed ? ((int64_t)a < (int64_t)b) : (a < b);
	int eq = a == b;
	switch (op) {
	case ORC_CMP_EQ:
		return eq;
	case ORC_CMP_NE:
		return !eq;
	case ORC_CMP_LT:
		return lt;
	case ORC_CMP_GT:
		return !lt && !eq;
	case ORC_CMP_LE:
		return lt || eq;
	default:
		return !lt;
	}
}

void orc_segment_filter(const orc_segment *s, int flags, int op, uint64_t constant, orc_sum128 *out,
                        uint64_t *bitmap_words) {
	uint8_t buf[ORC_VECTOR_SIZE * 8];
	memset(out, 0, sizeof(*out));
	for (uint64_t start = 0; start < s->count; start += ORC_VECTOR_SIZE) {
		uint64_t n = s->count - start < ORC_VECTOR_SIZE ? s->count - start : ORC_VECTOR_SIZE;
		orc_scan_partial(s, start, n, buf, 0, flags);
		for (uint64_t i = 0; i < n; i++) {
			uint64_t v = extend_T(buf + i * s->type_size, s->type_size, s->is_signed);
			if (compare_T(v, constant, op, s->is_signed)) {
				sum128_add(out, v, s->is_signed);
				if (bitmap_words) {
					uint64_t row = start + i;
					bitmap_words[row >> 6] |= 1ULL << (row & 63);
				}
			}
		}
	}
}

/* ColumnSegment::FilterSelection restated over one segment, vector by vector (column_segment.cpp:685-844):
 * the selection vector `sel` is refined child by child (CONJUNCTION_AND, :717-723) or the children's results are
 * united (CONJUNCTION_OR, :688-716); CONSTANT_COMPARISON = TemplatedFilterSelection (:575-586: a row passes when
 * `(!HAS_NULL || mask.RowIsValid(idx)) && OP(vec[idx], constant)`), IS_NULL / IS_NOT_NULL = TemplatedNullSelection
 * (:660-683).  Afterwards the aggregate DuckDB runs above the scan: SUM / COUNT(col) over the selected rows that are
 * valid, COUNT(*) = approved_tuple_count. */
static uint32_t select_child(const orc_segment *s, const uint8_t *vec, const orc_pred *p, const uint32_t *sel, uint32_t n,
                             const uint64_t *validity, uint64_t vrow0, uint32_t *result_sel) {
	uint32_t m = 0;
	for (uint32_t i = 0; i < n; i++) {
		uint32_t idx = sel[i];
		int valid = 1;
		if (validity) {
			uint64_t r = vrow0 + idx;
			valid = (int)((validity[r >> 6] >> (r & 63)) & 1);
		}
		int pass;
		if (p->kind == ORC_PRED_COMPARE) {
			uint64_t v = extend_T(vec + (uint64_t)idx * s->type_size, s->type_size, s->is_signed);
			pass = valid && compare_T(v, (uint64_t)p->constant, p->op, s->is_signed);
		} else {
			pass = valid != (p->kind == ORC_PRED_IS_NULL);
		}
		if (pass) {
			result_sel[m++] = idx;
		}
	}
	return m;
}

void orc_segment_select(const orc_segment *s, int flags, const orc_pred *preds, uint32_t n_preds, int conj_or,
                        const uint64_t *validity, uint64_t validity_row0, const uint64_t *bitmap_in,
                        uint64_t *bitmap_out, uint64_t bitmap_row0, orc_sum128 *out, uint64_t *valid_count) {
	uint8_t buf[ORC_VECTOR_SIZE * 8];
	static __thread uint32_t sel[ORC_VECTOR_SIZE], tmp[ORC_VECTOR_SIZE], res[ORC_VECTOR_SIZE];
	static __thread uint8_t seen[ORC_VECTOR_SIZE];
	memset(out, 0, sizeof(*out));
	*valid_count = 0;
	for (uint64_t start = 0; start < s->count; start += ORC_VECTOR_SIZE) {
		uint32_t n = (uint32_t)(s->count - start < ORC_VECTOR_SIZE ? s->count - start : ORC_VECTOR_SIZE);
		orc_scan_partial(s, start, n, buf, 0, flags);
		/* incoming selection */
		uint32_t approved = 0;
		for (uint32_t i = 0; i < n; i++) {
			uint64_t r = bitmap_row0 + start + i;
			if (!bitmap_in || ((bitmap_in[r >> 6] >> (r & 63)) & 1)) {
				sel[approved++] = i;
			}
		}
		uint64_t vrow0 = validity_row0 + start;
		if (n_preds > 0 && !conj_or) {
			for (uint32_t p = 0; p < n_preds; p++) {
				approved = select_child(s, buf, &preds[p], sel, approved, validity, vrow0, tmp);
				memcpy(sel, tmp, approved * sizeof(uint32_t));
			}
		} else if (n_preds > 0) {
			uint32_t total = 0;
			memset(seen, 0, sizeof(seen));
			for (uint32_t p = 0; p < n_preds; p++) {
				uint32_t m = select_child(s, buf, &preds[p], sel, approved, validity, vrow0, tmp);
				for (uint32_t i = 0; i < m; i++) {
					if (!seen[tmp[i]]) { /* the reference de-duplicates with a linear search, :701-711 */
						seen[tmp[i]] = 1;
						res[total++] = tmp[i];
					}
				}
			}
			memcpy(sel, res, total * sizeof(uint32_t));
			approved = total;
		}
		for (uint32_t i = 0; i < approved; i++) {
			uint32_t idx = sel[i];
			uint64_t r = vrow0 + idx;
			int valid = !validity || ((validity[r >> 6] >> (r & 63)) & 1);
			out->count++;
			if (valid) {
				uint64_t c = out->count;
				sum128_add(out, extend_T(buf + (uint64_t)idx * s->type_size, s->type_size, s->is_signed), s->is_signed);
				out->count = c;
				(*valid_count)++;
			}
			if (bitmap_out) {
				uint64_t b = bitmap_row0 + start + idx;
				bitmap_out[b >> 6] |= 1ULL << (b & 63);
			}
		}
	}
}

/* ------------------------------------------------------------------------------------------- */
/* Timed CPU baseline: the reference's loop structure, one segment per task, T threads          */
/* ------------------------------------------------------------------------------------------- */

typedef struct {
	const uint8_t *values;
	uint64_t n, seg_capacity, n_segments;
	uint8_t type_size;
	int is_signed, threads, tid, copy_per_call, phase;
	orc_segment *segs;
	pthread_barrier_t *barrier;
	double t[5];
	orc_sum128 sum;
	uint64_t packed_bytes;
	uint8_t *sink; /* per-thread 2048-value result vector */
} bench_task;

static double now_s(void) {
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static void scan_vector_call(const orc_segment *s, uint64_t start, uint64_t n, void *result, int copy_per_call) {
	if (copy_per_call) {
		/* quirk Q4: `auto source = segment.succinct_vec;` deep-copies the payload (succinct.cpp:127) */
		orc_segment tmp = *s;
		uint64_t bytes = (orc_iv_capacity(&s->succinct_vec) >> 3) + 8;
		tmp.succinct_vec.m_data = (uint64_t *)malloc(bytes);
		memcpy(tmp.succinct_vec.m_data, s->succinct_vec.m_data, bytes);
		orc_scan_partial(&tmp, start, n, result, 0, 0);
		free(tmp.succinct_vec.m_data);
	} else {
		orc_scan_partial(s, start, n, result, 0, 0);
	}
}

static void *bench_worker(void *arg) {
	bench_task *t = (bench_task *)arg;
	double t0;
	/* phase 0: append + analyze (SuccinctAppend, 2048 values per call like ColumnData::AppendData) */
	pthread_barrier_wait(t->barrier);
	t0 = now_s();
	for (uint64_t g = (uint64_t)t->tid; g < t->n_segments; g += (uint64_t)t->threads) {
		orc_segment *s = &t->segs[g];
		uint64_t first = g * t->seg_capacity;
		uint64_t cnt = t->n - first < t->seg_capacity ? t->n - first : t->seg_capacity;
		for (uint64_t off = 0; off < cnt; off += ORC_VECTOR_SIZE) {
			uint64_t m = cnt - off < ORC_VECTOR_SIZE ? cnt - off : ORC_VECTOR_SIZE;
			orc_segment_append(s, t->values + (first + off) * t->type_size, NULL, NULL, 0, m);
		}
	}
	t->t[0] = now_s() - t0;
	/* phase 1: encode */
	pthread_barrier_wait(t->barrier);
	t0 = now_s();
	for (uint64_t g = (uint64_t)t->tid; g < t->n_segments; g += (uint64_t)t->threads) {
		orc_bit_compress_from_succinct(&t->segs[g], 0);
		t->packed_bytes += orc_iv_capacity(&t->segs[g].succinct_vec) >> 3;
	}
	t->t[1] = now_s() - t0;
	/* phase 2: full scan into 2048-value vectors */
	pthread_barrier_wait(t->barrier);
	t0 = now_s();
	for (uint64_t g = (uint64_t)t->tid; g < t->n_segments; g += (uint64_t)t->threads) {
		const orc_segment *s = &t->segs[g];
		for (uint64_t off = 0; off < s->count; off += ORC_VECTOR_SIZE) {
			uint64_t m = s->count - off < ORC_VECTOR_SIZE ? s->count - off : ORC_VECTOR_SIZE;
			scan_vector_call(s, off, m, t->sink, t->copy_per_call);
		}
	}
	t->t[2] = now_s() - t0;
	/* phase 3: scan + SUM (decode a vector, then aggregate it, as the pipeline above the scan does) */
	pthread_barrier_wait(t->barrier);
	t0 = now_s();
	for (uint64_t g = (uint64_t)t->tid; g < t->n_segments; g += (uint64_t)t->threads) {
		const orc_segment *s = &t->segs[g];
		for (uint64_t off = 0; off < s->count; off += ORC_VECTOR_SIZE) {
			uint64_t m = s->count - off < ORC_VECTOR_SIZE ? s->count - off : ORC_VECTOR_SIZE;
			scan_vector_call(s, off, m, t->sink, t->copy_per_call);
			for (uint64_t i = 0; i < m; i++) {
				sum128_add(&t->sum, extend_T(t->sink + i * t->type_size, t->type_size, t->is_signed), t->is_signed);
			}
		}
	}
	t->t[3] = now_s() - t0;
	/* phase 4: UncompressSuccinct back to a flat block */
	pthread_barrier_wait(t->barrier);
	t0 = now_s();
	for (uint64_t g = (uint64_t)t->tid; g < t->n_segments; g += (uint64_t)t->threads) {
		orc_uncompress_succinct(&t->segs[g], 0);
	}
	t->t[4] = now_s() - t0;
	return NULL;
}

int orc_bench_column(const void *values, uint64_t n, uint8_t type_size, int is_signed, uint64_t seg_capacity,
                     int threads, int copy_per_call, orc_bench_result *out) {
	if (threads < 1 || seg_capacity == 0 || n == 0) {
		return -1;
	}
	uint64_t n_segments = (n + seg_capacity - 1) / seg_capacity;
	orc_segment *segs = (orc_segment *)malloc(sizeof(orc_segment) * n_segments);
	for (uint64_t g = 0; g < n_segments; g++) {
		orc_segment_init_succinct(&segs[g], type_size, is_signed, seg_capacity * type_size);
	}
	pthread_barrier_t barrier;
	pthread_barrier_init(&barrier, NULL, (unsigned)threads);
	bench_task *tasks = (bench_task *)calloc((size_t)threads, sizeof(bench_task));
	pthread_t *tids = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
	for (int i = 0; i < threads; i++) {
		bench_task *t = &tasks[i];
		t->values = (const uint8_t *)values;
		t->n = n;
		t->seg_capacity = seg_capacity;
		t->n_segments = n_segments;
		t->type_size = type_size;
		t->is_signed = is_signed;
		t->threads = threads;
		t->tid = i;
		t->copy_per_call = copy_per_call;
		t->segs = segs;
		t->barrier = &barrier;
		t->sink = (uint8_t *)malloc(ORC_VECTOR_SIZE * 8);
		pthread_create(&tids[i], NULL, bench_worker, t);
	}
	memset(out, 0, sizeof(*out));
	for (int i = 0; i < threads; i++) {
		pthread_join(tids[i], NULL);
		bench_task *t = &tasks[i];
		if (t->t[0] > out->analyze_s) out->analyze_s = t->t[0];
		if (t->t[1] > out->encode_s) out->encode_s = t->t[1];
		if (t->t[2] > out->scan_s) out->scan_s = t->t[2];
		if (t->t[3] > out->sum_s) out->sum_s = t->t[3];
		if (t->t[4] > out->uncompact_s) out->uncompact_s = t->t[4];
		uint64_t old = out->sum.lo;
		out->sum.lo += t->sum.lo;
		out->sum.hi += t->sum.hi + (out->sum.lo < old);
		out->sum.count += t->sum.count;
		out->packed_bytes += t->packed_bytes;
		free(t->sink);
	}
	out->values = n;
	for (uint64_t g = 0; g < n_segments; g++) {
		orc_segment_free(&segs[g]);
	}
	pthread_barrier_destroy(&barrier);
	free(segs);
	free(tasks);
	free(tids);
	return 0;
}

/* ------------------------------------------------------------------------------------------- */
/* Opaque-handle mirror of the calls above, same shape as oracle/_ref's ref_* entry points so    */
/* that the Python test glue can drive L0 and L1 through one code path.                          */
/* ------------------------------------------------------------------------------------------- */

void *orc_segment_new_succinct(uint8_t type_size, int is_signed, uint64_t segment_size) {
	orc_segment *s = (orc_segment *)malloc(sizeof(orc_segment));
	orc_segment_init_succinct(s, type_size, is_signed, segment_size);
	return s;
}

void *orc_segment_new_uncompressed(uint8_t type_size, int is_signed, uint64_t segment_size) {
	orc_segment *s = (orc_segment *)malloc(sizeof(orc_segment));
	orc_segment_init_uncompressed(s, type_size, is_signed, segment_size);
	return s;
}

void orc_segment_delete(void *h) {
	if (h) {
		orc_segment_free((orc_segment *)h);
		free(h);
	}
}

uint64_t orc_segment_get(void *h, int field) {
	const orc_segment *s = (const orc_segment *)h;
	switch (field) {
	case 0: return s->count;
	case 1: return s->succinct_vec.m_width;
	case 2: return s->succinct_vec.m_size;
	case 3: return s->min_factor;
	case 4: return s->max_factor;
	case 5: return s->compacted;
	case 6: return s->repacked;
	case 7: return (uint64_t)s->function;
	case 8: return orc_iv_capacity(&s->succinct_vec) >> 6;
	case 9: return s->num_elements;
	default: return 0;
	}
}

const uint64_t *orc_segment_words(void *h) {
	return ((const orc_segment *)h)->succinct_vec.m_data;
}

const uint8_t *orc_segment_block(void *h) {
	return ((const orc_segment *)h)->block;
}

uint64_t orc_segment_serialize(void *h, uint8_t *out, uint64_t out_capacity) {
	const orc_segment *s = (const orc_segment *)h;
	uint64_t need = orc_iv_size_in_bytes(&s->succinct_vec);
	if (need <= out_capacity) {
		orc_iv_serialize(&s->succinct_vec, out);
	}
	return need;
}
