// filter_range_test.cpp -- exhaustive check of canonical_filter_range (csrc/filter_range.hpp) on an 8-bit domain.
// For every conjunction of 1..3 children (6 comparisons x 12 constants inside, at the ends of and outside the type's
// range, IS NOT NULL) under AND and OR, for a signed and an unsigned 8-bit type: whenever the function says "one range",
// that range (or its complement) must select exactly the values TemplatedFilterSelection would
// (src/storage/table/column_segment.cpp:575-586: `mask.RowIsValid(idx) && OP(vec[idx], constant)`); IS NULL children
// must be refused; the shapes the fast kernel is meant for must be accepted.  Every accepted range of up to two children
// is then clipped to every linear 8-bit segment the way filter_words_kernel does it (clip_range_to_segment) and the
// field-domain test is compared with the value test on every field.  Plain C++: g++ -std=c++17, no GPU.
#include "../filter_range.hpp"

#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace adacom;

static bool t_less(bool is_signed, uint64_t a, uint64_t b) {
	return is_signed ? (int64_t)a < (int64_t)b : a < b;
}

static bool cmp_passes(bool is_signed, uint64_t v, uint64_t c, uint32_t op_mask) {
	const bool lt = t_less(is_signed, v, c);
	const uint32_t sel = lt ? 0u : (v == c ? 1u : 2u);
	return (op_mask >> sel) & 1u;
}

// clip_range_to_segment (what filter_words_kernel runs per segment): every linear 8-bit segment -- extend(min) + field,
// field < 2^w, the interval inside the type -- and every field of it: the field-domain test must agree with the value test
static unsigned long long clipped = 0;
static unsigned long long check_segments(bool is_signed, const FilterRange &fr) {
	unsigned long long n_checked = 0;
	const int lo = is_signed ? -128 : 0, hi = is_signed ? 127 : 255;
	for (int mn = lo; mn <= hi; mn += 3) {
		for (int w = 1; w <= 8; w++) {
			const int mask = (1 << w) - 1;
			if (mn + mask > hi) {
				break;
			}
			const uint64_t emin = (uint64_t)(int64_t)mn, etop = (uint64_t)(int64_t)(mn + mask);
			const FieldRange r = is_signed ? clip_range_to_segment<true>(fr, emin, etop, (uint64_t)mask)
			                               : clip_range_to_segment<false>(fr, emin, etop, (uint64_t)mask);
			for (int f = 0; f <= mask; f++) {
				const uint64_t v = (uint64_t)(int64_t)(mn + f);
				const bool inr = !fr.empty && !t_less(is_signed, v, fr.lo) && !t_less(is_signed, fr.hi, v);
				const bool want = inr != (fr.complement != 0u);
				const bool in_field = ((uint64_t)f - r.flo) <= r.span;
				const bool got = r.kind == 0 ? false : (in_field != (r.kind == 2));
				if (got != want || r.kind < 0 || r.kind > 2) {
					printf("CLIP MISMATCH signed=%d min=%d w=%d f=%d want=%d got=%d kind=%d flo=%llu span=%llu\n", (int)is_signed, mn,
					       w, f, want, got, r.kind, (unsigned long long)r.flo, (unsigned long long)r.span);
					exit(1);
				}
				n_checked++;
			}
		}
	}
	return n_checked;
}

int main() {
	const uint32_t masks[6] = {0x2u, 0x5u, 0x1u, 0x4u, 0x3u, 0x6u};  // EQ NE LT GT LE GE (cmp_mask_for_op)
	const int64_t consts[12] = {-300, -129, -128, -127, -1, 0, 1, 126, 127, 128, 255, 256};
	std::vector<SelectPred> leaves;
	for (uint32_t m : masks) {
		for (int64_t c : consts) {
			leaves.push_back(SelectPred {0u, m, (uint64_t)c});
		}
	}
	leaves.push_back(SelectPred {2u, 0u, 0ull});  // IS NOT NULL
	const size_t L = leaves.size();
	unsigned long long checked = 0, accepted = 0, refused = 0;
	for (int is_signed = 0; is_signed < 2; is_signed++) {
		uint64_t domain[256];
		for (int i = 0; i < 256; i++) {
			domain[i] = is_signed ? (uint64_t)(int64_t)(int8_t)i : (uint64_t)(uint8_t)i;
		}
		for (int conj_or = 0; conj_or < 2; conj_or++) {
			for (uint32_t n = 0; n <= 3; n++) {
				size_t total = 1;
				for (uint32_t k = 0; k < n; k++) {
					total *= L;
				}
				for (size_t code = 0; code < total; code++) {
					SelectPred preds[3];
					size_t x = code;
					for (uint32_t k = 0; k < n; k++) {
						preds[k] = leaves[x % L];
						x /= L;
					}
					FilterRange fr;
					checked++;
					if (!canonical_filter_range(preds, n, conj_or != 0, is_signed != 0, &fr)) {
						refused++;
						continue;
					}
					accepted++;
					if (n <= 2) {
						clipped += check_segments(is_signed != 0, fr);
					}
					for (int i = 0; i < 256; i++) {
						const uint64_t v = domain[i];
						bool any = false, all = true;
						for (uint32_t k = 0; k < n; k++) {
							const bool r = preds[k].kind == 0 ? cmp_passes(is_signed, v, preds[k].constant, preds[k].op_mask) : true;
							any |= r;
							all &= r;
						}
						const bool want = conj_or ? (n == 0 || any) : all;
						const bool inr = !fr.empty && !t_less(is_signed, v, fr.lo) && !t_less(is_signed, fr.hi, v);
						const bool got = inr != (fr.complement != 0u);
						if (got != want) {
							printf("MISMATCH signed=%d or=%d n=%u code=%zu v=%lld want=%d got=%d range=[%lld,%lld] empty=%u compl=%u\n",
							       is_signed, conj_or, n, code, (long long)v, want, got, (long long)fr.lo, (long long)fr.hi, fr.empty,
							       fr.complement);
							return 1;
						}
					}
				}
			}
			// the shapes the fast kernel is there for are accepted
			FilterRange fr;
			for (uint32_t m : masks) {
				SelectPred one[1] = {{0u, m, 5ull}};
				if (!canonical_filter_range(one, 1, conj_or != 0, is_signed != 0, &fr)) {
					printf("single comparison refused\n");
					return 1;
				}
			}
			SelectPred between[2] = {{0u, 0x6u, 3ull}, {0u, 0x3u, 90ull}};
			SelectPred outside[2] = {{0u, 0x1u, 3ull}, {0u, 0x4u, 90ull}};
			if (!canonical_filter_range(between, 2, false, is_signed != 0, &fr) ||
			    !canonical_filter_range(outside, 2, true, is_signed != 0, &fr) || !fr.complement) {
				printf("BETWEEN / outside refused\n");
				return 1;
			}
			// IS NULL children are never a value range
			SelectPred with_null[2] = {{0u, 0x2u, 3ull}, {1u, 0u, 0ull}};
			if (canonical_filter_range(with_null, 2, conj_or != 0, is_signed != 0, &fr) ||
			    canonical_filter_range(with_null + 1, 1, conj_or != 0, is_signed != 0, &fr)) {
				printf("IS NULL accepted\n");
				return 1;
			}
		}
	}
	printf("FILTER_RANGE_OK checked=%llu accepted=%llu refused=%llu clipped=%llu\n", checked, accepted, refused, clipped);
	return accepted > refused / 4 ? 0 : 1;
}
