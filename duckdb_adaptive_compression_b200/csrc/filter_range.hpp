// filter_range.hpp -- host-side canonical form of a single-column table filter (plain C++, no CUDA: the unit test
// csrc/tests/filter_range_test.cpp compiles it with g++ and checks it exhaustively over an 8-bit domain).
//
// ColumnSegment::FilterSelection (src/storage/table/column_segment.cpp:685-844) evaluates a conjunction of
// CONSTANT_COMPARISON / IS [NOT] NULL children row by row.  Whenever that conjunction amounts to "the value lies in ONE
// range [lo, hi] of T, or outside it" over the valid rows, the field-domain filter kernel can serve it
// (kernels_filter.cu: filter_words_kernel); canonical_filter_range decides that and produces the range.
#pragma once

#include <cstdint>

namespace adacom {

// A single-column table filter in canonical form: the decoded value (sign- or zero-extended T) lies in [lo, hi] (compared
// as T: signed or unsigned), or outside it (complement).  Every comparison of TemplatedFilterSelection, a BETWEEN (AND of
// two) and an OR of two half-lines fit; `empty`: the range holds no value at all.
struct FilterRange {
	uint64_t lo = 0, hi = 0;
	uint32_t empty = 0, complement = 0;
};

constexpr uint32_t kSelectMaxPreds = 8;
struct SelectPred {
	uint32_t kind;       // 0 comparison, 1 IS NULL, 2 IS NOT NULL
	uint32_t op_mask;    // cmp_mask_for_op: truth table packed as bits {lt, eq, gt}
	uint64_t constant;   // sign/zero-extended T value
};

#ifdef __CUDACC__
#define ADACOM_FR_HD __host__ __device__ __forceinline__
#else
#define ADACOM_FR_HD inline
#endif

// The canonical range against ONE segment whose decoded values are extend(min) + field, field in [0, mask] without
// wrap-around (a "linear" segment: [emin, etop] is an interval of T).  The range is clipped to the interval and moved to
// the FIELD domain: a row passes iff kind == 1 and (field - flo) <= span, or kind == 2 and not (field - flo) <= span;
// kind == 0: no row of the segment passes (zone map).  Used by filter_words_kernel per segment, and by the CPU test.
struct FieldRange {
	uint64_t flo, span;
	int kind;
};
template <bool SIGNED>
ADACOM_FR_HD bool filter_value_less(uint64_t a, uint64_t b) {
	return SIGNED ? (int64_t)a < (int64_t)b : a < b;
}
template <bool SIGNED>
ADACOM_FR_HD FieldRange clip_range_to_segment(const FilterRange &fr, uint64_t emin, uint64_t etop, uint64_t mask64) {
	bool none = fr.empty != 0u;
	uint64_t l = fr.lo, h = fr.hi;
	if (!none) {
		if (filter_value_less<SIGNED>(h, emin) || filter_value_less<SIGNED>(etop, l)) {
			none = true;  // the segment's interval and the range do not meet (zone map)
		} else {
			l = filter_value_less<SIGNED>(l, emin) ? emin : l;
			h = filter_value_less<SIGNED>(etop, h) ? etop : h;
		}
	}
	FieldRange r;
	r.flo = 0;
	r.span = mask64;
	if (none) {
		r.kind = fr.complement ? 1 : 0;  // complement of nothing: every row of the segment
	} else {
		r.flo = l - emin;
		r.span = h - l;
		const bool whole = r.flo == 0 && r.span == mask64;
		r.kind = !fr.complement ? 1 : (whole ? 0 : 2);
	}
	return r;
}

// Value ranges in KEY space: key(x) = x with the sign bit flipped for signed types, so unsigned order on keys is the
// type's order on 64-bit extended values (the constants of a comparison are taken as 64-bit, like cmp_passes does).
struct KeyRange {
	uint64_t lo = 0, hi = ~0ull;
	bool empty = false;
	bool holds(uint64_t k) const {
		return !empty && k >= lo && k <= hi;
	}
};

inline KeyRange key_range_for(uint32_t op_mask, uint64_t k) {
	KeyRange r;
	switch (op_mask & 7u) {
	case 1: r.empty = k == 0; r.hi = k - 1; break;        // <
	case 3: r.hi = k; break;                              // <=
	case 4: r.empty = k == ~0ull; r.lo = k + 1; break;    // >
	case 6: r.lo = k; break;                              // >=
	case 2: r.lo = r.hi = k; break;                       // =
	default: break;                                       // (<> is handled by the callers: complement of [k, k])
	}
	return r;
}

inline FilterRange filter_range_from(const KeyRange &r, bool complement, bool is_signed) {
	const uint64_t flip = is_signed ? (1ull << 63) : 0ull;
	FilterRange f;
	f.lo = r.lo ^ flip;
	f.hi = r.hi ^ flip;
	f.empty = r.empty ? 1u : 0u;
	f.complement = complement ? 1u : 0u;
	return f;
}

inline FilterRange filter_range_for_cmp(uint32_t op_mask, uint64_t constant, bool is_signed) {
	const uint64_t k = constant ^ (is_signed ? (1ull << 63) : 0ull);
	if ((op_mask & 7u) == 5u) {
		KeyRange r;
		r.lo = r.hi = k;
		return filter_range_from(r, true, is_signed);
	}
	return filter_range_from(key_range_for(op_mask, k), false, is_signed);
}

// The conjunction of a FilterSelection as ONE value range (or the complement of one) + "NULL rows never pass".
// Returns false when the conjunction has no such form (IS NULL children, two separate ranges ...): select_kernel's job.
inline bool canonical_filter_range(const SelectPred *preds, uint32_t n_preds, bool conj_or, bool is_signed,
                                   FilterRange *out) {
	const uint64_t flip = is_signed ? (1ull << 63) : 0ull;
	KeyRange acc;
	uint64_t ne[kSelectMaxPreds];
	uint32_t n_ne = 0;
	if (!conj_or) {
		for (uint32_t p = 0; p < n_preds; p++) {
			const SelectPred &pr = preds[p];
			if (pr.kind == 1) {
				return false;  // IS NULL
			}
			if (pr.kind != 0) {
				continue;  // IS NOT NULL: the validity mask already says so
			}
			const uint64_t k = pr.constant ^ flip;
			if ((pr.op_mask & 7u) == 5u) {
				ne[n_ne++] = k;
				continue;
			}
			const KeyRange r = key_range_for(pr.op_mask, k);
			acc.empty = acc.empty || r.empty;
			acc.lo = r.lo > acc.lo ? r.lo : acc.lo;
			acc.hi = r.hi < acc.hi ? r.hi : acc.hi;
			acc.empty = acc.empty || acc.lo > acc.hi;
		}
		// <> children: outside the range they say nothing, at an end they shorten it, inside they split it
		bool changed = true;
		while (changed && !acc.empty) {
			changed = false;
			for (uint32_t i = 0; i < n_ne && !acc.empty; i++) {
				if (!acc.holds(ne[i])) {
					continue;
				}
				if (acc.lo == acc.hi) {
					acc.empty = true;
				} else if (ne[i] == acc.lo) {
					acc.lo++;
					changed = true;
				} else if (ne[i] == acc.hi) {
					acc.hi--;
					changed = true;
				}
			}
		}
		uint32_t inside = 0;
		uint64_t hole = 0;
		for (uint32_t i = 0; i < n_ne; i++) {
			if (acc.holds(ne[i]) && !(inside && ne[i] == hole)) {
				inside++;
				hole = ne[i];
			}
		}
		if (inside == 0) {
			*out = filter_range_from(acc, false, is_signed);
			return true;
		}
		if (inside == 1 && acc.lo == 0 && acc.hi == ~0ull) {
			KeyRange r;
			r.lo = r.hi = hole;
			*out = filter_range_from(r, true, is_signed);
			return true;
		}
		return false;
	}
	// OR: zero children pass everything (like select_kernel); IS NOT NULL passes every valid row
	if (n_preds == 0) {
		*out = filter_range_from(acc, false, is_signed);
		return true;
	}
	KeyRange rs[kSelectMaxPreds];
	uint32_t n_r = 0;
	for (uint32_t p = 0; p < n_preds; p++) {
		const SelectPred &pr = preds[p];
		if (pr.kind == 1) {
			return false;
		}
		if (pr.kind != 0) {
			*out = filter_range_from(KeyRange(), false, is_signed);
			return true;
		}
		const uint64_t k = pr.constant ^ flip;
		if ((pr.op_mask & 7u) == 5u) {
			ne[n_ne++] = k;
			continue;
		}
		const KeyRange r = key_range_for(pr.op_mask, k);
		if (!r.empty) {
			rs[n_r++] = r;
		}
	}
	if (n_ne) {
		// (x <> c) OR X: everything when X holds c (or another <> does), else x <> c
		bool all = false;
		for (uint32_t i = 1; i < n_ne; i++) {
			all = all || ne[i] != ne[0];
		}
		for (uint32_t i = 0; i < n_r; i++) {
			all = all || rs[i].holds(ne[0]);
		}
		KeyRange r;
		if (!all) {
			r.lo = r.hi = ne[0];
		}
		*out = filter_range_from(r, !all, is_signed);
		return true;
	}
	if (n_r == 0) {
		acc.empty = true;
		*out = filter_range_from(acc, false, is_signed);
		return true;
	}
	// merge the ranges in order of their lower ends
	for (uint32_t i = 1; i < n_r; i++) {
		for (uint32_t j = i; j > 0 && rs[j].lo < rs[j - 1].lo; j--) {
			const KeyRange t = rs[j];
			rs[j] = rs[j - 1];
			rs[j - 1] = t;
		}
	}
	KeyRange merged[kSelectMaxPreds];
	uint32_t n_m = 0;
	for (uint32_t i = 0; i < n_r; i++) {
		if (n_m && (merged[n_m - 1].hi == ~0ull || rs[i].lo <= merged[n_m - 1].hi + 1)) {
			merged[n_m - 1].hi = rs[i].hi > merged[n_m - 1].hi ? rs[i].hi : merged[n_m - 1].hi;
		} else {
			merged[n_m++] = rs[i];
		}
	}
	if (n_m == 1) {
		*out = filter_range_from(merged[0], false, is_signed);
		return true;
	}
	if (n_m == 2 && merged[0].lo == 0 && merged[1].hi == ~0ull) {
		KeyRange gap;
		gap.lo = merged[0].hi + 1;
		gap.hi = merged[1].lo - 1;
		*out = filter_range_from(gap, true, is_signed);
		return true;
	}
	return false;
}

}  // namespace adacom
