// kernels_encode.cu -- analyze / layout / encode side of the succinct codec for sm_100a.
//
//   analyze  <- SuccinctAppendLoop<T>'s min/max tracking (src/storage/compression/succinct.cpp:271-306):
//               per-segment unsigned min/max of uint64_t(T) (sign-extending) or of the zero-extended raw
//               bytes (adaptive encoder, src/storage/table/column_segment.cpp:390-400); NULL rows are
//               skipped and their slot is overwritten with NullValue<T>() (succinct.cpp:288-292).
//   layout   <- the width rule of BitCompressFromSuccinct / BitCompressFromUncompressed
//               (column_segment.cpp:351-363, 408-419): w = hi(max-min)+1, optional pad to bytes, re-pack
//               iff 8*S > w; plus an exclusive scan that places every payload on a 128-byte boundary.
//   encode   <- the re-pack loop (column_segment.cpp:365-378, 427-446): field = (slot - min) & lo_set[w],
//               written LSB-first at bit i*w; tail bits of the last word are zero
//               (third_party/sdsl/include/sdsl/memory_management.hpp:360-365).
//
// The reference packs in place with a scalar read/modify/write cursor; a parallel encoder cannot, so raw
// tiles are staged through shared memory by the TMA engine and every thread assembles whole output words
// (gather formulation: no atomics, no inter-thread carries) that leave as coalesced stores.
#include "adacom_kernels.cuh"

namespace adacom {

namespace {

// ---------------------------------------------------------------------------------------------------
// analyze
// ---------------------------------------------------------------------------------------------------

constexpr int kAnalyzeThreads = 256;

template <int S>
__device__ __forceinline__ void minmax_word(uint32_t wd, uint32_t &mn, uint32_t &mx) {
	// unsigned min/max of the S-byte lanes packed in one 32-bit word
	if (S == 4) {
		mn = min(mn, wd);
		mx = max(mx, wd);
	} else if (S == 2) {
		uint32_t a = wd & 0xffffu, b = wd >> 16;
		mn = min(mn, min(a, b));
		mx = max(mx, max(a, b));
	} else {
		uint32_t a = wd & 0xffu, b = (wd >> 8) & 0xffu, c = (wd >> 16) & 0xffu, d = wd >> 24;
		mn = min(mn, min(min(a, b), min(c, d)));
		mx = max(mx, max(max(a, b), max(c, d)));
	}
}

template <int S>
__device__ __forceinline__ uint64_t null_pattern() {
	// NullValue<T>() for signed T = numeric_limits<T>::min(); unsigned T = 0
	return 1ull << (8 * S - 1);
}

// One CTA per segment.  The unsigned order of the raw S-byte patterns equals the order of both
// uint64_t(sign_extend(v)) and zero_extend(v), so min/max are taken on the patterns and extended once.
template <int S, bool SIGNED>
__global__ void __launch_bounds__(kAnalyzeThreads) analyze_kernel(SegDesc *desc, uint32_t seg_begin, uint8_t *raw,
                                                                  const uint64_t *validity, int zero_ext,
                                                                  const uint32_t *seg_list) {
	using T = typename UIntOf<S>::type;
	SegDesc *d = desc + (seg_list ? __ldg(seg_list + blockIdx.x) : seg_begin + blockIdx.x);
	const uint64_t row_start = d->row_start;
	const uint32_t count = d->count;
	T *vals = (T *)raw + row_start;
	const uint32_t tid = threadIdx.x;

	uint64_t mn = ~0ull, mx = 0ull;  // over zero-extended patterns
	bool any = false;

	if (validity == nullptr) {
		// head elements up to the first 16-byte boundary, 128-bit body, tail elements
		uint64_t addr = (uint64_t)vals;
		uint32_t head = (uint32_t)(((16u - (addr & 15u)) & 15u) / S);
		if (head > count) {
			head = count;
		}
		uint32_t body_vecs = (count - head) / (16 / S);
		uint32_t tail_first = head + body_vecs * (16 / S);
		if (tid < head) {
			uint64_t v = vals[tid];
			mn = min(mn, v);
			mx = max(mx, v);
		}
		if (tail_first + tid < count) {  // fewer than 16/S <= 16 tail elements
			uint64_t v = vals[tail_first + tid];
			mn = min(mn, v);
			mx = max(mx, v);
		}
		const uint4 *body = (const uint4 *)(vals + head);
		if (S == 8) {
			uint64_t m0 = ~0ull, m1 = ~0ull, x0 = 0, x1 = 0;
#pragma unroll 4
			for (uint32_t i = tid; i < body_vecs; i += kAnalyzeThreads) {
				uint4 q = __ldg(body + i);
				uint64_t a = ((uint64_t)q.y << 32) | q.x, b = ((uint64_t)q.w << 32) | q.z;
				m0 = min(m0, a);
				x0 = max(x0, a);
				m1 = min(m1, b);
				x1 = max(x1, b);
			}
			mn = min(mn, min(m0, m1));
			mx = max(mx, max(x0, x1));
		} else {
			uint32_t m32 = 0xffffffffu, x32 = 0;
#pragma unroll 4
			for (uint32_t i = tid; i < body_vecs; i += kAnalyzeThreads) {
				uint4 q = __ldg(body + i);
				minmax_word<S>(q.x, m32, x32);
				minmax_word<S>(q.y, m32, x32);
				minmax_word<S>(q.z, m32, x32);
				minmax_word<S>(q.w, m32, x32);
			}
			if (body_vecs > tid) {
				mn = min(mn, (uint64_t)m32);
				mx = max(mx, (uint64_t)x32);
			}
		}
		any = count > 0;
	} else {
		for (uint32_t i = tid; i < count; i += kAnalyzeThreads) {
			uint64_t row = row_start + i;
			bool ok = (validity[row >> 6] >> (row & 63)) & 1ull;
			if (ok) {
				uint64_t v = vals[i];
				mn = min(mn, v);
				mx = max(mx, v);
				any = true;
			} else {
				vals[i] = SIGNED ? (T)null_pattern<S>() : (T)0;
			}
		}
	}

	// block reduction
	__shared__ uint64_t s_mn[kAnalyzeThreads / 32], s_mx[kAnalyzeThreads / 32];
	__shared__ int s_any[kAnalyzeThreads / 32];
	const unsigned full = 0xffffffffu;
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		mn = min(mn, __shfl_down_sync(full, mn, o));
		mx = max(mx, __shfl_down_sync(full, mx, o));
	}
	int warp_any = __any_sync(full, any);
	if ((tid & 31) == 0) {
		s_mn[tid >> 5] = mn;
		s_mx[tid >> 5] = mx;
		s_any[tid >> 5] = warp_any;
	}
	__syncthreads();
	if (tid == 0) {
		int a = 0;
		for (int wv = 0; wv < kAnalyzeThreads / 32; wv++) {
			mn = min(mn, s_mn[wv]);
			mx = max(mx, s_mx[wv]);
			a |= s_any[wv];
		}
		uint64_t umin = ~0ull, umax = 0ull;
		if (a) {
			if (SIGNED && !zero_ext) {
				umin = extend_low<S, true>(mn);
				umax = extend_low<S, true>(mx);
			} else {
				umin = mn;
				umax = mx;
			}
		}
		d->min_factor = umin;
		d->max_factor = umax;
		d->state |= kSegAnalyzed;
	}
}

// ---------------------------------------------------------------------------------------------------
// layout: width + repack decision + exclusive scan of 128-byte aligned payload sizes (single CTA)
// ---------------------------------------------------------------------------------------------------

constexpr int kLayoutThreads = 256;
constexpr int kLayoutMaxCtas = 128;  // <= SM count: every CTA is resident, a CTA only ever waits for lower-indexed ones
// scratch: two words per CTA, {scan chunks (<< 43) | words (43 bits), ticket of the call that wrote them}: a slot is
// trusted only while its ticket word equals the call's full 64-bit ticket (tickets never repeat on a context)

struct LayoutItem {
	unsigned long long aligned;  // payload words rounded up to 128 bytes; ~0: descriptor left untouched
	uint32_t wst;                // new width | state << 8 | pad
	uint32_t parts;              // scan chunks the segment needs at its (new) width
};

// the width rule + re-pack decision of one segment (column_segment.cpp:351-363, 408-419)
__device__ __forceinline__ LayoutItem layout_item(const SegDesc *d, bool padded, bool zero_ext, uint32_t slot_w) {
	LayoutItem it;
	const uint64_t mn = d->min_factor, mx = d->max_factor;
	const uint32_t count = d->count;
	uint32_t wst = ((const uint32_t *)d)[14];
	const uint32_t state = (wst >> 8) & 0xffu;
	uint32_t w_now = wst & 0xffu;
	it.aligned = ~0ull;
	// Compact() returns early for empty or already compacted segments (column_segment.cpp:278)
	if (count != 0 && !(state & kSegCompacted)) {
		uint64_t range;
		if (zero_ext) {
			// BitCompressFromUncompressed only subtracts when max > min (column_segment.cpp:408-411)
			range = (mx != 0 && mn != ~0ull && mx > mn) ? mx - mn : mx;
		} else {
			range = mx - mn;
		}
		const uint32_t w = width_for_range(range, padded);
		const bool repack = slot_w > w;
		const uint64_t words = repack ? ((uint64_t)count * w + 63) >> 6 : 0;
		it.aligned = (words + kSegAlignWords - 1) & ~(uint64_t)(kSegAlignWords - 1);
		const uint32_t new_state = (state & kSegAnalyzed) | kSegCompacted | (repack ? kSegRepacked : 0);
		w_now = repack ? w : slot_w;
		wst = (wst & 0xffff0000u) | (new_state << 8) | w_now;
	}
	it.wst = wst;
	it.parts = 1;
	if (count != 0 && w_now != 0) {
		const uint32_t gran = scan_chunk_values(w_now);
		it.parts = (count + gran - 1) / gran;
	}
	return it;
}

// Up to 128 resident CTAs, thread t of CTA b owns `per` consecutive segments.  Pass 1 sizes them, a block scan and
// a chained scan over the CTAs (each CTA publishes {words, scan chunks} tagged with the call's ticket and sums
// the words of the CTAs before it -- they were dispatched earlier, so the wait cannot deadlock) give every
// segment its 128-byte aligned word offset; pass 2 writes the descriptors back.  The last CTA publishes the arena
// size into host-mapped pinned memory {words, ticket}: the host spins on the ticket instead of paying a stream
// synchronisation.
__global__ void __launch_bounds__(kLayoutThreads) layout_kernel(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs,
                                                                uint32_t per, int type_size, int flags, uint32_t extent,
                                                                volatile unsigned long long *totals,
                                                                unsigned long long ticket, uint32_t *parts,
                                                                int whole_column, volatile unsigned long long *scratch,
                                                                const uint32_t *seg_list) {
	__shared__ unsigned long long warp_tot[kLayoutThreads / 32];
	__shared__ unsigned long long base_s;
	__shared__ uint32_t parts_s;
	const uint32_t tid = threadIdx.x;
	const uint32_t lane = tid & 31u, warp = tid >> 5;
	const bool padded = flags & kFlagPadded;
	const bool zero_ext = flags & kFlagZeroExt;
	const uint32_t slot_w = 8u * (uint32_t)type_size;
	if (tid == 0) {
		base_s = 0;
		parts_s = 1;
	}
	const uint32_t g0 = (blockIdx.x * kLayoutThreads + tid) * per;

	// ---- pass 1: sizes ----
	unsigned long long mine = 0;
	uint32_t my_parts = 1;
	LayoutItem first;  // per == 1 (up to 32768 segments): nothing is computed twice
	for (uint32_t i = 0; i < per && g0 + i < n_segs; i++) {
		const LayoutItem it = layout_item(desc + (seg_list ? __ldg(seg_list + g0 + i) : seg_begin + g0 + i), padded, zero_ext, slot_w);
		if (i == 0) {
			first = it;
		}
		if (it.aligned != ~0ull) {
			mine += it.aligned;
		}
		my_parts = max(my_parts, it.parts);
	}
	// ---- exclusive scan inside the CTA ----
	unsigned long long incl = mine;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		unsigned long long v = __shfl_up_sync(0xffffffffu, incl, d);
		if ((int)lane >= d) {
			incl += v;
		}
	}
	my_parts = __reduce_max_sync(0xffffffffu, my_parts);
	if (lane == 31) {
		warp_tot[warp] = incl;
	}
	__syncthreads();
	if (lane == 0) {
		atomicMax(&parts_s, my_parts);
	}
	unsigned long long warp_base = 0, cta_total = 0;
#pragma unroll
	for (int wv = 0; wv < kLayoutThreads / 32; wv++) {
		const unsigned long long t = warp_tot[wv];
		warp_base += wv < (int)warp ? t : 0ull;
		cta_total += t;
	}
	__syncthreads();
	// ---- chained scan over the CTAs ----
	if (tid == 0) {
		scratch[2u * blockIdx.x] = ((unsigned long long)parts_s << 43) | cta_total;
		__threadfence();
		scratch[2u * blockIdx.x + 1u] = ticket;
	}
	unsigned long long before = 0;
	uint32_t parts_before = 1;
	for (uint32_t j = tid; j < blockIdx.x; j += kLayoutThreads) {
		while (scratch[2u * j + 1u] != ticket) {
		}
		__threadfence();
		const unsigned long long v = scratch[2u * j];
		before += v & ((1ull << 43) - 1ull);
		parts_before = max(parts_before, (uint32_t)((v >> 43) & 31u));
	}
	if (blockIdx.x != 0) {
#pragma unroll
		for (int d = 16; d > 0; d >>= 1) {
			before += __shfl_down_sync(0xffffffffu, before, d);
		}
		parts_before = __reduce_max_sync(0xffffffffu, parts_before);
		if (lane == 0) {
			atomicAdd((unsigned long long *)&base_s, before);
			atomicMax(&parts_s, parts_before);
		}
	}
	__syncthreads();
	// ---- pass 2: write back ----
	unsigned long long off = base_s + warp_base + incl - mine;
	for (uint32_t i = 0; i < per && g0 + i < n_segs; i++) {
		SegDesc *d = desc + (seg_list ? __ldg(seg_list + g0 + i) : seg_begin + g0 + i);
		const LayoutItem it = i == 0 ? first : layout_item(d, padded, zero_ext, slot_w);
		if (it.aligned != ~0ull) {
			d->packed_ptr = 0;
			d->word_offset = off;
			d->extent = extent;
			d->prev_wst = ((const uint32_t *)d)[14] & 0xffffu;
			((uint32_t *)d)[14] = it.wst;
			off += it.aligned;
		}
	}
	if (blockIdx.x == gridDim.x - 1 && tid == 0) {
		if (whole_column) {
			*parts = parts_s;
		} else {
			atomicMax(parts, parts_s);
		}
		totals[0] = base_s + cta_total;
		__threadfence_system();
		totals[1] = ticket;
	}
}

__global__ void bind_extent_kernel(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint64_t *base,
                                   uint32_t extent, const uint32_t *seg_list) {
	uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g < n_segs) {
		SegDesc *d = desc + (seg_list ? __ldg(seg_list + g) : seg_begin + g);
		if ((d->state & kSegRepacked) && d->extent == extent && d->packed_ptr == 0) {
			d->packed_ptr = (uint64_t)(base + d->word_offset);
		}
	}
}

// Rollback of a layout whose encode could not be completed (extent allocation failed, a later launch failed): every
// descriptor the layout of extent id `extent` rewrote and that was never bound to a payload gets the width / state it
// had before; its extent id is cleared so that the next encode -- which reuses the id -- cannot bind it by mistake.
__global__ void revert_layout_kernel(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint32_t extent, uint32_t *parts,
                                     uint32_t raw_parts, const uint32_t *seg_list) {
	uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g == 0) {
		atomicMax(parts, raw_parts);  // slot-width segments need the most scan chunks
	}
	if (g < n_segs) {
		SegDesc *d = desc + (seg_list ? __ldg(seg_list + g) : seg_begin + g);
		if ((d->state & kSegCompacted) && d->extent == extent && d->packed_ptr == 0) {
			const uint32_t prev = d->prev_wst;
			d->width = (uint8_t)(prev & 0xffu);
			d->state = (uint8_t)((prev >> 8) & 0xffu);
			d->word_offset = 0;
			d->extent = ~0u;
		}
	}
}

// ColumnSegment::UncompressSuccinct's bookkeeping (column_segment.cpp:494-503): function -> UNCOMPRESSED,
// compacted = false; the payload was already decoded into the raw arena by the decode kernel.
__global__ void reset_uncompacted_kernel(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int type_size,
                                         uint32_t *parts, uint32_t raw_parts, const uint32_t *seg_list) {
	uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g == 0) {
		atomicMax(parts, raw_parts);  // slot-width segments need the most scan chunks
	}
	if (g < n_segs) {
		SegDesc *d = desc + (seg_list ? __ldg(seg_list + g) : seg_begin + g);
		if (d->state & kSegCompacted) {
			d->state = (uint8_t)((d->state & kSegAnalyzed) | kSegUncompressedFn);
			d->width = (uint8_t)(8 * type_size);
			d->packed_ptr = 0;
			d->word_offset = 0;
		}
	}
}

template <int S>
cudaError_t launch_analyze_s(bool is_signed, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint8_t *raw,
                             const uint64_t *validity, int zero_ext, cudaStream_t stream, const uint32_t *seg_list) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	if (is_signed) {
		analyze_kernel<S, true><<<n_segs, kAnalyzeThreads, 0, stream>>>(desc, seg_begin, raw, validity, zero_ext, seg_list);
	} else {
		analyze_kernel<S, false><<<n_segs, kAnalyzeThreads, 0, stream>>>(desc, seg_begin, raw, validity, zero_ext, seg_list);
	}
	return cudaGetLastError();
}

}  // namespace

cudaError_t launch_analyze(TypeInfo t, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint8_t *raw,
                           const uint64_t *validity, int zero_ext, cudaStream_t stream, const uint32_t *seg_list) {
	switch (t.size) {
	case 1:
		return launch_analyze_s<1>(t.is_signed, desc, seg_begin, n_segs, raw, validity, zero_ext, stream, seg_list);
	case 2:
		return launch_analyze_s<2>(t.is_signed, desc, seg_begin, n_segs, raw, validity, zero_ext, stream, seg_list);
	case 4:
		return launch_analyze_s<4>(t.is_signed, desc, seg_begin, n_segs, raw, validity, zero_ext, stream, seg_list);
	case 8:
		return launch_analyze_s<8>(t.is_signed, desc, seg_begin, n_segs, raw, validity, zero_ext, stream, seg_list);
	default:
		return cudaErrorInvalidValue;
	}
}

cudaError_t launch_layout(TypeInfo t, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int flags, uint32_t extent,
                          unsigned long long *totals, unsigned long long ticket, uint32_t *parts, int whole_column,
                          unsigned long long *scratch, cudaStream_t stream, const uint32_t *seg_list) {
	uint32_t grid = (n_segs + kLayoutThreads - 1) / kLayoutThreads;
	if (grid > (uint32_t)kLayoutMaxCtas) {
		grid = kLayoutMaxCtas;
	}
	if (grid == 0) {
		grid = 1;
	}
	const uint32_t per = (n_segs + grid * kLayoutThreads - 1) / (grid * kLayoutThreads);
	layout_kernel<<<grid, kLayoutThreads, 0, stream>>>(desc, seg_begin, n_segs, per ? per : 1, t.size, flags, extent, totals,
	                                                   ticket, parts, whole_column, scratch, seg_list);
	return cudaGetLastError();
}

cudaError_t launch_bind_extent(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint64_t *base, uint32_t extent,
                               cudaStream_t stream, const uint32_t *seg_list) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	bind_extent_kernel<<<(n_segs + 255) / 256, 256, 0, stream>>>(desc, seg_begin, n_segs, base, extent, seg_list);
	return cudaGetLastError();
}

cudaError_t launch_revert_layout(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint32_t extent, uint32_t *parts,
                                 uint32_t raw_parts, cudaStream_t stream, const uint32_t *seg_list) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	revert_layout_kernel<<<(n_segs + 255) / 256, 256, 0, stream>>>(desc, seg_begin, n_segs, extent, parts, raw_parts, seg_list);
	return cudaGetLastError();
}

cudaError_t launch_reset_uncompacted(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int type_size,
                                     uint32_t *parts, uint32_t raw_parts, cudaStream_t stream,
                                     const uint32_t *seg_list) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	reset_uncompacted_kernel<<<(n_segs + 255) / 256, 256, 0, stream>>>(desc, seg_begin, n_segs, type_size, parts, raw_parts,
	                                                                   seg_list);
	return cudaGetLastError();
}

}  // namespace adacom
