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
                                                                  const uint64_t *validity, int zero_ext) {
	using T = typename UIntOf<S>::type;
	SegDesc *d = desc + seg_begin + blockIdx.x;
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

constexpr int kLayoutThreads = 1024;

__global__ void __launch_bounds__(kLayoutThreads) layout_kernel(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs,
                                                                int type_size, int flags, uint32_t extent,
                                                                unsigned long long *totals) {
	__shared__ unsigned long long partial[kLayoutThreads];
	const uint32_t tid = threadIdx.x;
	const uint32_t per = (n_segs + kLayoutThreads - 1) / kLayoutThreads;
	const uint32_t lo = min(n_segs, tid * per), hi = min(n_segs, lo + per);
	const bool padded = flags & kFlagPadded;
	const bool zero_ext = flags & kFlagZeroExt;
	const uint32_t slot_w = 8u * (uint32_t)type_size;

	unsigned long long mine = 0;
	for (uint32_t g = lo; g < hi; g++) {
		SegDesc *d = desc + seg_begin + g;
		uint32_t count = d->count;
		if (count == 0 || (d->state & kSegCompacted)) {
			continue;  // Compact() returns early for empty or already compacted segments (column_segment.cpp:278)
		}
		uint64_t mn = d->min_factor, mx = d->max_factor;
		uint64_t range;
		if (zero_ext) {
			// BitCompressFromUncompressed only subtracts when max > min (column_segment.cpp:408-411)
			range = (mx != 0 && mn != ~0ull && mx > mn) ? mx - mn : mx;
		} else {
			range = mx - mn;
		}
		uint32_t w = width_for_range(range, padded);
		bool repack = slot_w > w;
		uint64_t words = repack ? ((uint64_t)count * w + 63) >> 6 : 0;
		uint64_t aligned = (words + kSegAlignWords - 1) & ~(uint64_t)(kSegAlignWords - 1);
		d->width = (uint8_t)(repack ? w : slot_w);
		d->state = (uint8_t)((d->state & kSegAnalyzed) | kSegCompacted | (repack ? kSegRepacked : 0));
		d->word_offset = mine;  // local offset, rebased below
		d->extent = extent;
		d->packed_ptr = 0;
		mine += aligned;
	}
	partial[tid] = mine;
	__syncthreads();
	// inclusive Hillis-Steele scan over the 1024 partials
	for (uint32_t off = 1; off < kLayoutThreads; off <<= 1) {
		unsigned long long v = tid >= off ? partial[tid - off] : 0;
		__syncthreads();
		partial[tid] += v;
		__syncthreads();
	}
	unsigned long long base = partial[tid] - mine;
	for (uint32_t g = lo; g < hi; g++) {
		SegDesc *d = desc + seg_begin + g;
		if (d->count != 0 && d->extent == extent && d->packed_ptr == 0 && (d->state & kSegRepacked)) {
			d->word_offset += base;
		}
	}
	if (tid == kLayoutThreads - 1) {
		totals[0] = partial[tid];
	}
}

__global__ void bind_extent_kernel(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint64_t *base,
                                   uint32_t extent) {
	uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g < n_segs) {
		SegDesc *d = desc + seg_begin + g;
		if ((d->state & kSegRepacked) && d->extent == extent && d->packed_ptr == 0) {
			d->packed_ptr = (uint64_t)(base + d->word_offset);
		}
	}
}

// ColumnSegment::UncompressSuccinct's bookkeeping (column_segment.cpp:494-503): function -> UNCOMPRESSED,
// compacted = false; the payload was already decoded into the raw arena by the decode kernel.
__global__ void reset_uncompacted_kernel(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int type_size) {
	uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g < n_segs) {
		SegDesc *d = desc + seg_begin + g;
		if (d->state & kSegCompacted) {
			d->state = (uint8_t)((d->state & kSegAnalyzed) | kSegUncompressedFn);
			d->width = (uint8_t)(8 * type_size);
			d->packed_ptr = 0;
			d->word_offset = 0;
		}
	}
}

// ---------------------------------------------------------------------------------------------------
// encode
// ---------------------------------------------------------------------------------------------------

constexpr int kEncodeConsumerWarps = 8;
constexpr int kEncodeConsumers = kEncodeConsumerWarps * 32;
constexpr int kEncodeThreads = kEncodeConsumers + 32;  // + one producer warp feeding the TMA ring
constexpr int kEncodeStages = 3;
template <int S> struct EncodeStage { static constexpr uint32_t kBytes = kVectorSize * S + 32; };

struct EncodeInfo {
	uint64_t sub;        // value subtracted from every slot (0 when min_factor is unset)
	uint64_t *out;       // first output word of this tile
	uint32_t byte_base;  // offset of the tile's first value inside the staged bytes (0..15)
	uint32_t cnt;        // values in the tile (0: segment not repacked, nothing to do)
	uint32_t w;
	uint32_t gap;        // last tile of a segment: zero words up to the next 128-byte boundary
	uint32_t magic;      // ceil(2^24 / w): n / w == (n * magic) >> 24 for n < 2^17, w <= 64
	uint32_t pad;
};

struct EncodeResolved {
	EncodeInfo info;
	uint64_t src16;
	uint32_t bytes;
};

template <int S>
__device__ __forceinline__ EncodeResolved resolve_encode_tile(const SegDesc *desc, const Job *tiles, uint32_t tile_idx,
                                                              const uint8_t *raw, uint32_t extent) {
	EncodeResolved r;
	Job j = tiles[tile_idx];
	const SegDesc *d = desc + j.seg;
	uint32_t count = d->count;
	uint32_t w = d->width;
	uint32_t first = j.k * kVectorSize;
	bool repacked = (d->state & kSegRepacked) && d->extent == extent && d->packed_ptr != 0;
	uint32_t cnt = repacked ? min(kVectorSize, count - first) : 0u;
	uint64_t src = (uint64_t)raw + (d->row_start + first) * (uint64_t)S;
	r.src16 = src & ~15ull;
	uint32_t byte_base = (uint32_t)(src - r.src16);
	r.bytes = cnt ? ((byte_base + cnt * S + 15u) & ~15u) : 0u;
	uint64_t mn = d->min_factor;
	r.info.sub = mn == ~0ull ? 0ull : mn;
	r.info.out = (uint64_t *)d->packed_ptr + (uint64_t)j.k * (kVectorSize / 64) * w;
	r.info.byte_base = byte_base;
	r.info.cnt = cnt;
	r.info.w = w;
	uint32_t gap = 0;
	if (cnt && first + cnt == count) {
		uint64_t total = ((uint64_t)count * w + 63) >> 6;
		gap = (uint32_t)(((total + kSegAlignWords - 1) & ~(uint64_t)(kSegAlignWords - 1)) - total);
	}
	r.info.gap = gap;
	r.info.magic = ((1u << 24) + w - 1u) / w;
	r.info.pad = 0;
	return r;
}

template <int S>
__device__ __forceinline__ uint64_t load_slot(const uint8_t *base, uint32_t i) {
	// the slot-width int_vector read: zero-extended S raw bytes (bits::read_int at width 8*S)
	using T = typename UIntOf<S>::type;
	return (uint64_t)((const T *)base)[i];
}

__device__ __forceinline__ uint32_t div_magic(uint32_t n, uint32_t magic) {
	return (uint32_t)(((uint64_t)n * magic) >> 24);
}

template <int S>
__global__ void __launch_bounds__(kEncodeThreads) encode_kernel(const SegDesc *desc, const Job *tiles, uint32_t n_tiles,
                                                                const uint8_t *raw, uint32_t extent) {
	extern __shared__ __align__(128) uint8_t stage_mem[];
	__shared__ __align__(8) uint64_t full_bar[kEncodeStages];
	__shared__ __align__(8) uint64_t empty_bar[kEncodeStages];
	__shared__ EncodeInfo infos[kEncodeStages];
	constexpr uint32_t kStage = EncodeStage<S>::kBytes;
	const uint32_t tid = threadIdx.x;
	const uint32_t lane = tid & 31u;
	const uint32_t warp = tid >> 5;

	if (tid == 0) {
		for (int s = 0; s < kEncodeStages; s++) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&empty_bar[s], 1);
		}
		mbar_fence_init();
	}
	__syncthreads();

	if (warp == kEncodeConsumerWarps) {
		// ---------------- producer warp: lane s feeds stage s ----------------
		for (uint32_t it0 = 0;; it0 += kEncodeStages) {
			if (blockIdx.x + it0 * gridDim.x >= n_tiles) {
				break;
			}
			const uint32_t my_it = it0 + lane;
			const uint32_t t = blockIdx.x + my_it * gridDim.x;
			const bool have = lane < kEncodeStages && t < n_tiles;
			EncodeResolved r;
			if (have) {
				r = resolve_encode_tile<S>(desc, tiles, t, raw, extent);
			}
#pragma unroll
			for (int s = 0; s < kEncodeStages; s++) {
				if ((int)lane == s && have) {
					mbar_wait(&empty_bar[s], ((my_it / kEncodeStages) & 1u) ^ 1u);
					infos[s] = r.info;
					mbar_arrive_expect_tx(&full_bar[s], r.bytes);
					if (r.bytes) {
						tma_load_1d(stage_mem + s * kStage, (const void *)r.src16, r.bytes, &full_bar[s]);
					}
				}
				__syncwarp();
			}
		}
		return;
	}

	// ---------------- consumer warps ----------------
	for (uint32_t it = 0;; it++) {
		uint32_t t = blockIdx.x + it * gridDim.x;
		if (t >= n_tiles) {
			break;
		}
		const uint32_t s = it % kEncodeStages;
		mbar_wait(&full_bar[s], (it / kEncodeStages) & 1u);
		const EncodeInfo in = infos[s];
		const uint8_t *vals = stage_mem + s * kStage + in.byte_base;
		const uint32_t w = in.w;
		const uint32_t cnt = in.cnt;

		if (cnt) {
			if (w <= 32u) {
				// 32-bit output units: unit j holds bits [32j, 32j+32) of the tile's stream
				const uint32_t n_units = (((cnt * w + 63u) >> 6) << 1);
				const uint32_t mask = (uint32_t)lo_set64(w);
				const uint32_t sub = (uint32_t)in.sub;
				uint32_t *out32 = (uint32_t *)in.out;
				for (uint32_t j = tid; j < n_units; j += kEncodeConsumers) {
					const uint32_t bit0 = j * 32u;
					const uint32_t i0 = div_magic(bit0, in.magic);
					const uint32_t i1 = min(cnt - 1u, div_magic(bit0 + 31u, in.magic));
					uint32_t acc = 0;
					if (i0 < cnt) {
						const uint32_t c = i1 - i0 + 1u;
						uint32_t r = (lane * c) >> 5;  // rotate the start so the lanes of a warp hit different banks
						for (uint32_t n = 0; n < c; n++) {
							const uint32_t i = i0 + r;
							const uint32_t f = ((uint32_t)load_slot<S>(vals, i) - sub) & mask;
							const int p = (int)(i * w) - (int)bit0;
							acc |= p >= 0 ? (f << p) : (f >> (-p));
							r = (r + 1u == c) ? 0u : r + 1u;
						}
					}
					out32[j] = acc;
				}
			} else {
				const uint32_t n_words = (cnt * w + 63u) >> 6;
				const uint64_t mask = lo_set64(w);
				for (uint32_t j = tid; j < n_words; j += kEncodeConsumers) {
					const uint32_t bit0 = j * 64u;
					const uint32_t i0 = div_magic(bit0, in.magic);
					const uint32_t i1 = min(cnt - 1u, div_magic(bit0 + 63u, in.magic));
					uint64_t acc = 0;
					for (uint32_t i = i0; i <= i1; i++) {  // at most 3 values touch one word when w > 32
						const uint64_t f = (load_slot<S>(vals, i) - in.sub) & mask;
						const int p = (int)(i * w) - (int)bit0;
						acc |= p >= 0 ? (f << p) : (f >> (-p));
					}
					in.out[j] = acc;
				}
			}
			if (tid < in.gap) {  // keep the alignment gap behind the payload deterministic (zero)
				in.out[((cnt * w + 63u) >> 6) + tid] = 0ull;
			}
		}
		// consumers in lockstep (named barrier 1), then one of them frees the stage for the producer
		asm volatile("bar.sync 1, %0;" ::"n"(kEncodeConsumers) : "memory");
		if (tid == 0) {
			mbar_arrive(&empty_bar[s]);
		}
	}
}

template <int S>
cudaError_t launch_encode_s(const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw,
                            uint32_t extent, int sm_count, cudaStream_t stream) {
	static bool configured[64] = {};
	const size_t smem = (size_t)kEncodeStages * EncodeStage<S>::kBytes;
	auto kernel = encode_kernel<S>;
	int dev = 0;
	cudaGetDevice(&dev);
	if (!configured[dev & 63]) {
		cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		configured[dev & 63] = true;
	}
	if (n_tiles == 0) {
		return cudaSuccess;
	}
	uint32_t per_sm = S == 8 ? 4u : (S == 4 ? 6u : 8u);
	uint32_t grid = (uint32_t)sm_count * per_sm;
	if (grid > n_tiles) {
		grid = n_tiles;
	}
	kernel<<<grid, kEncodeThreads, smem, stream>>>(desc, tiles, n_tiles, raw, extent);
	return cudaGetLastError();
}

template <int S>
cudaError_t launch_analyze_s(bool is_signed, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint8_t *raw,
                             const uint64_t *validity, int zero_ext, cudaStream_t stream) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	if (is_signed) {
		analyze_kernel<S, true><<<n_segs, kAnalyzeThreads, 0, stream>>>(desc, seg_begin, raw, validity, zero_ext);
	} else {
		analyze_kernel<S, false><<<n_segs, kAnalyzeThreads, 0, stream>>>(desc, seg_begin, raw, validity, zero_ext);
	}
	return cudaGetLastError();
}

}  // namespace

cudaError_t launch_analyze(TypeInfo t, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint8_t *raw,
                           const uint64_t *validity, int zero_ext, cudaStream_t stream) {
	switch (t.size) {
	case 1:
		return launch_analyze_s<1>(t.is_signed, desc, seg_begin, n_segs, raw, validity, zero_ext, stream);
	case 2:
		return launch_analyze_s<2>(t.is_signed, desc, seg_begin, n_segs, raw, validity, zero_ext, stream);
	case 4:
		return launch_analyze_s<4>(t.is_signed, desc, seg_begin, n_segs, raw, validity, zero_ext, stream);
	case 8:
		return launch_analyze_s<8>(t.is_signed, desc, seg_begin, n_segs, raw, validity, zero_ext, stream);
	default:
		return cudaErrorInvalidValue;
	}
}

cudaError_t launch_layout(TypeInfo t, SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int flags, uint32_t extent,
                          unsigned long long *totals, cudaStream_t stream) {
	layout_kernel<<<1, kLayoutThreads, 0, stream>>>(desc, seg_begin, n_segs, t.size, flags, extent, totals);
	return cudaGetLastError();
}

cudaError_t launch_bind_extent(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, uint64_t *base, uint32_t extent,
                               cudaStream_t stream) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	bind_extent_kernel<<<(n_segs + 255) / 256, 256, 0, stream>>>(desc, seg_begin, n_segs, base, extent);
	return cudaGetLastError();
}

cudaError_t launch_reset_uncompacted(SegDesc *desc, uint32_t seg_begin, uint32_t n_segs, int type_size,
                                     cudaStream_t stream) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	reset_uncompacted_kernel<<<(n_segs + 255) / 256, 256, 0, stream>>>(desc, seg_begin, n_segs, type_size);
	return cudaGetLastError();
}

cudaError_t launch_encode(TypeInfo t, const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw,
                          uint32_t extent, int sm_count, cudaStream_t stream) {
	switch (t.size) {
	case 1:
		return launch_encode_s<1>(desc, tiles, n_tiles, raw, extent, sm_count, stream);
	case 2:
		return launch_encode_s<2>(desc, tiles, n_tiles, raw, extent, sm_count, stream);
	case 4:
		return launch_encode_s<4>(desc, tiles, n_tiles, raw, extent, sm_count, stream);
	case 8:
		return launch_encode_s<8>(desc, tiles, n_tiles, raw, extent, sm_count, stream);
	default:
		return cudaErrorInvalidValue;
	}
}

}  // namespace adacom
