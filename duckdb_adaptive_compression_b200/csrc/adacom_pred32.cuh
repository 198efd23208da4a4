// adacom_pred32.cuh -- range predicate (and optional SUM) over the 32 consecutive rows a lane owns, with the field
// positions known at compile time.  Shared by the fused multi-column scan (kernels_query.cu) and the single-column
// filter (kernels_filter.cu).
#pragma once

#include "adacom_device.cuh"

namespace adacom {
namespace {

// ---- predicate on 32 consecutive rows of one lane, width known at compile time -----------------------------------
// 32 fields of W bits are exactly W 32-bit words.  The lane loads its W (+1) words, realigns them by the piece's bit
// offset with W funnel shifts, and every field then sits at a compile-time position: a row costs a shift, a mask, a
// subtract, a compare and a select (the lane = row formulation pays two shared loads, an address computation and a
// ballot share on top).  Consecutive lanes own consecutive groups (stride W words: conflict-free for odd W).
__host__ __device__ constexpr int gcd32(int w) {
	return (w % 32 == 0) ? 32 : ((w % 16 == 0) ? 16 : ((w % 8 == 0) ? 8 : ((w % 4 == 0) ? 4 : ((w % 2 == 0) ? 2 : 1))));
}

// The field positions repeat every 32 / gcd(W, 32) fields (= W / gcd words), so the 32 rows are walked as gcd(W, 32)
// iterations of ONE short loop body with compile-time positions: the straight-line version (230 instructions per
// width, three widths per query) made instruction fetch the top stall of the kernel (ncu: no_instruction).
template <int W>
__device__ __forceinline__ uint32_t pred32(const uint32_t *p, uint32_t s0, uint32_t flo, uint32_t span) {
	constexpr int G = gcd32(W);   // iterations
	constexpr int F = 32 / G;     // fields per iteration
	constexpr int PW = W / G;     // words per iteration
	constexpr uint32_t mask = W >= 32 ? 0xffffffffu : ((1u << (W & 31)) - 1u);
	uint32_t word = 0;
	// lanes start at different periods of their group and wrap around: for the widths that share factors with the 32
	// banks (8, 16, 32 ...) the lanes of a warp would otherwise hit the same bank (W = 32: a 32-way conflict)
	const uint32_t rot = threadIdx.x & (G - 1);
#pragma unroll 1
	for (int step = 0; step < G; step++) {
		const uint32_t it = (step + rot) & (G - 1);
		const uint32_t *q = p + it * PW;
		uint32_t r[PW + 1];
#pragma unroll
		for (int t = 0; t <= PW; t++) {
			r[t] = q[t];
		}
#pragma unroll
		for (int t = 0; t < PW; t++) {
			r[t] = __funnelshift_r(r[t], r[t + 1], s0);  // s0 == 0 leaves r[t]
		}
		uint32_t part[2] = {0u, 0u};
#pragma unroll
		for (int k = 0; k < F; k++) {
			const int bit = k * W, wi = bit >> 5, sh = bit & 31;
			uint32_t f;
			if (sh + W <= 32) {
				f = r[wi] >> sh;
				if (sh + W < 32) {
					f &= mask;
				}
			} else {
				f = __funnelshift_r(r[wi], r[wi + 1 < PW ? wi + 1 : wi], sh) & mask;
			}
			part[k & 1] |= (f - flo) <= span ? (1u << k) : 0u;
		}
		word |= (part[0] | part[1]) << (it * F);
	}
	return word;
}

// one copy of every width per kernel (not inlined at each call site): instruction fetch was the top stall
__device__ __noinline__ uint32_t pred32_any(uint32_t w, const uint32_t *p, uint32_t s0, uint32_t flo, uint32_t span) {
	switch (w) {
#define ADACOM_P32(W) case W: return pred32<W>(p, s0, flo, span);
	ADACOM_P32(1) ADACOM_P32(2) ADACOM_P32(3) ADACOM_P32(4) ADACOM_P32(5) ADACOM_P32(6) ADACOM_P32(7) ADACOM_P32(8)
	ADACOM_P32(9) ADACOM_P32(10) ADACOM_P32(11) ADACOM_P32(12) ADACOM_P32(13) ADACOM_P32(14) ADACOM_P32(15) ADACOM_P32(16)
	ADACOM_P32(17) ADACOM_P32(18) ADACOM_P32(19) ADACOM_P32(20) ADACOM_P32(21) ADACOM_P32(22) ADACOM_P32(23) ADACOM_P32(24)
	ADACOM_P32(25) ADACOM_P32(26) ADACOM_P32(27) ADACOM_P32(28) ADACOM_P32(29) ADACOM_P32(30) ADACOM_P32(31) ADACOM_P32(32)
#undef ADACOM_P32
	default: return 0xffffffffu;
	}
}


// the same with the sum of the passing fields (SUM(value) = sum(f) + n * extend(min) on a linear frame of reference)
template <int W>
__device__ __forceinline__ uint32_t pred32_sum(const uint32_t *p, uint32_t s0, uint32_t flo, uint32_t span, uint64_t &fsum) {
	constexpr int G = gcd32(W);
	constexpr int F = 32 / G;
	constexpr int PW = W / G;
	constexpr uint32_t mask = W >= 32 ? 0xffffffffu : ((1u << (W & 31)) - 1u);
	uint32_t word = 0;
	const uint32_t rot = threadIdx.x & (G - 1);  // see pred32
#pragma unroll 1
	for (int step = 0; step < G; step++) {
		const uint32_t it = (step + rot) & (G - 1);
		const uint32_t *q = p + it * PW;
		uint32_t r[PW + 1];
#pragma unroll
		for (int t = 0; t <= PW; t++) {
			r[t] = q[t];
		}
#pragma unroll
		for (int t = 0; t < PW; t++) {
			r[t] = __funnelshift_r(r[t], r[t + 1], s0);
		}
		uint32_t part[2] = {0u, 0u};
		uint32_t s32[2] = {0u, 0u};  // <= 16 fields below 2^26 each per chain, else 64-bit below
		uint64_t s64 = 0;
#pragma unroll
		for (int k = 0; k < F; k++) {
			const int bit = k * W, wi = bit >> 5, sh = bit & 31;
			uint32_t f;
			if (sh + W <= 32) {
				f = r[wi] >> sh;
				if (sh + W < 32) {
					f &= mask;
				}
			} else {
				f = __funnelshift_r(r[wi], r[wi + 1 < PW ? wi + 1 : wi], sh) & mask;
			}
			const bool pass = (f - flo) <= span;
			part[k & 1] |= pass ? (1u << k) : 0u;
			if (W <= 26) {
				s32[k & 1] += pass ? f : 0u;
			} else {
				s64 += pass ? (uint64_t)f : 0ull;
			}
		}
		word |= (part[0] | part[1]) << (it * F);
		fsum += W <= 26 ? (uint64_t)s32[0] + (uint64_t)s32[1] : s64;
	}
	return word;
}

__device__ __noinline__ uint32_t pred32_sum_any(uint32_t w, const uint32_t *p, uint32_t s0, uint32_t flo, uint32_t span,
                                                uint64_t &fsum) {
	switch (w) {
#define ADACOM_P32(W) case W: return pred32_sum<W>(p, s0, flo, span, fsum);
	ADACOM_P32(1) ADACOM_P32(2) ADACOM_P32(3) ADACOM_P32(4) ADACOM_P32(5) ADACOM_P32(6) ADACOM_P32(7) ADACOM_P32(8)
	ADACOM_P32(9) ADACOM_P32(10) ADACOM_P32(11) ADACOM_P32(12) ADACOM_P32(13) ADACOM_P32(14) ADACOM_P32(15) ADACOM_P32(16)
	ADACOM_P32(17) ADACOM_P32(18) ADACOM_P32(19) ADACOM_P32(20) ADACOM_P32(21) ADACOM_P32(22) ADACOM_P32(23) ADACOM_P32(24)
	ADACOM_P32(25) ADACOM_P32(26) ADACOM_P32(27) ADACOM_P32(28) ADACOM_P32(29) ADACOM_P32(30) ADACOM_P32(31) ADACOM_P32(32)
#undef ADACOM_P32
	default: return 0u;
	}
}

}  // namespace
}  // namespace adacom
