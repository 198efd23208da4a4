// kernels_codecs.cu -- SURVEY 8(f) row 4: the neighbour codecs of the succinct function, decoded on the GPU from the
// segment images DuckDB itself writes.
//
//   bitpacking  src/storage/compression/bitpacking.cpp: a segment = [u64 offset of the metadata end | group payloads growing
//               up | u32 metadata entries growing down (mode << 24 | payload offset), :60-73, :456-513]; every 2048 values
//               (BITPACKING_METADATA_GROUP_SIZE) form a group in one of four modes (:576-632):
//                 CONSTANT        [T constant]
//                 CONSTANT_DELTA  [T frame_of_reference][T delta]                  value_i = i * delta + for
//                 FOR             [T for][T width][packed]                         value_i = field_i + for
//                 DELTA_FOR       [T for][T width][T delta_offset][packed]         value_i = delta_offset + sum_{j<=i}(field_j + for)
//               packed = 32-value blocks of `width` bits, LSB first, in little-endian words (third_party/fastpforlib):
//               one contiguous bit string per group, field i at bits [i*w, (i+1)*w) -- the same layout as the succinct
//               codec's int_vector, only byte- instead of word-addressed (the payload is merely sizeof(T)-aligned).
//               Scan / fetch semantics: BitpackingScanPartial :729-811, BitpackingFetchRow :823-866.
//   RLE         src/storage/compression/rle.cpp: [u64 offset of the counts | T values[runs] | u16 counts[runs]] (:170-215);
//               RLEScanPartial / RLEFetchRow :298-331.
//
// One CTA works on one metadata group (bitpacking) or one 2048-row tile (RLE) at a time, starting from the group's
// 64-byte CodecGroup record (resolved once per column by codec_groups_kernel, so a scan has no dependent metadata
// loads).  Bitpacking: the payload is staged in double-buffered shared memory with aligned 4-byte loads (the
// misalignment becomes a bit offset; one barrier per group), lanes take consecutive values, so the stores are
// coalesced; DELTA_FOR runs a block-wide prefix sum.  RLE: the heads of the runs that start inside the tile are
// scattered into a 2048-entry shared array and a block-wide max-scan turns them into the run of every row -- no
// per-row binary search.  Fused SUM never materialises the values.
#include "adacom_kernels.cuh"

namespace adacom {

namespace {

constexpr int kCodecThreads = 256;
constexpr uint32_t kGroupValues = 2048;
enum : uint32_t { kBpConstant = 1, kBpConstantDelta = 2, kBpDeltaFor = 3, kBpFor = 4 };

__device__ __forceinline__ uint64_t load_bytes(const uint8_t *p, int n) {
	uint64_t v = 0;
	for (int b = 0; b < n; b++) {
		v |= (uint64_t)__ldg(p + b) << (8 * b);
	}
	return v;
}

// header of one bitpacking group, resolved by one thread and broadcast through shared memory
struct GroupHeader {
	uint64_t frame, constant, delta_offset;
	uint32_t mode, width, data_off, count;
};

template <int S>
__device__ __forceinline__ GroupHeader read_group_header(const uint8_t *img, uint32_t seg_count, uint32_t g) {
	GroupHeader h;
	const uint64_t meta_end = *(const uint64_t *)img;  // images start 16-byte aligned
	const uint32_t entry = (uint32_t)load_bytes(img + meta_end - 4ull * (g + 1), 4);
	h.mode = entry >> 24;
	uint32_t off = entry & 0x00FFFFFFu;
	h.frame = h.constant = h.delta_offset = 0;
	h.width = 0;
	h.count = min(kGroupValues, seg_count - g * kGroupValues);
	if (h.mode == kBpConstant) {
		h.constant = load_bytes(img + off, S);
		off += S;
	} else {
		h.frame = load_bytes(img + off, S);
		off += S;
		if (h.mode == kBpConstantDelta) {
			h.constant = load_bytes(img + off, S);
			off += S;
		} else {
			h.width = (uint32_t)load_bytes(img + off, S) & 0xffu;
			off += S;
			if (h.mode == kBpDeltaFor) {
				h.delta_offset = load_bytes(img + off, S);
				off += S;
			}
		}
	}
	h.data_off = off;
	return h;
}

// w-bit field at bit position `bit` of a staged 32-bit word array (w <= 64; two pad words behind the data)
__device__ __forceinline__ uint64_t field_at(const uint32_t *words, uint32_t bit, uint32_t w) {
	const uint32_t idx = bit >> 5, sh = bit & 31u;
	const uint32_t a = words[idx], b = words[idx + 1], c = words[idx + 2];
	const uint64_t v = ((uint64_t)__funnelshift_r(b, c, sh) << 32) | __funnelshift_r(a, b, sh);
	return v & lo_set64(w);
}

template <int S>
__device__ __forceinline__ uint64_t trunc_T(uint64_t v) {
	return S == 8 ? v : (v & ((1ull << (8 * (S & 7))) - 1ull));
}

// records of the jobs (one thread per group): bitpacking header fields / RLE run range of the tile
template <int S>
__global__ void __launch_bounds__(kCodecThreads) codec_groups_kernel(CodecArgs a, CodecGroup *groups, int codec) {
	const uint32_t j = blockIdx.x * kCodecThreads + threadIdx.x;
	if (j >= a.n_jobs) {
		return;
	}
	const Job job = a.jobs[j];
	const CodecSeg seg = a.segs[job.seg];
	const uint8_t *img = a.images + seg.image_offset;
	CodecGroup g;
	g.frame = g.constant = g.delta_offset = 0;
	g.mode = g.width = g.run_lo = g.run_hi = 0;
	g.first = job.k * kGroupValues;
	g.out_row = seg.row_start + (uint64_t)job.k * kGroupValues;
	g.count = g.first < seg.count ? min(kGroupValues, seg.count - g.first) : 0u;
	if (codec == 0) {
		const GroupHeader h = read_group_header<S>(img, seg.count, job.k);
		g.payload = seg.image_offset + h.data_off;
		g.frame = h.frame;
		g.constant = h.constant;
		g.delta_offset = h.delta_offset;
		g.mode = h.mode;
		g.width = h.width;
	} else {
		const uint64_t counts_off = *(const uint64_t *)img;
		const uint32_t n_runs = (seg.bytes - (uint32_t)counts_off) / 2u;
		const uint32_t *starts = a.run_start + seg.first_run;
		g.payload = seg.image_offset + 8u;
		auto find = [&](uint32_t row) {
			uint32_t lo = 0, hi = n_runs;  // last run whose start <= row
			while (hi - lo > 1) {
				const uint32_t mid = (lo + hi) >> 1;
				if (starts[mid] <= row) {
					lo = mid;
				} else {
					hi = mid;
				}
			}
			return lo;
		};
		if (g.count) {
			g.run_lo = seg.first_run + find(g.first);
			g.run_hi = seg.first_run + find(g.first + g.count - 1u);
		}
	}
	groups[j] = g;
}

// MODE 0: decode into out[(row_start + i) - out_row0]; MODE 1: fused SUM/COUNT
template <int S, bool SIGNED, int MODE>
__global__ void __launch_bounds__(kCodecThreads) bitpacking_groups_kernel(CodecArgs a) {
	using T = typename UIntOf<S>::type;
	constexpr uint32_t kStageWords = kGroupValues * 8 * S / 32 + 8;  // widest payload of the type + the funnel's look-ahead
	__shared__ __align__(16) uint32_t stage2[2][kStageWords];
	__shared__ uint64_t warp_tot[kCodecThreads / 32];
	__shared__ uint64_t red[96];
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	Acc96 acc;
	uint32_t n_rows = 0;
	uint32_t parity = 0;
	// Software pipeline over the CTA's groups: the payload of group j + grid is copied (cp.async, 4 bytes per request:
	// the source is only sizeof(T)-aligned) into the other stage buffer while group j is decoded, so the barrier at the
	// top of an iteration finds its data already in shared memory (the kernel was latency-bound: 35 % of the DRAM
	// throughput at 39 % of the issue slots, profiles/r03q_codecs_ncu_summary.txt).
	auto stage_group = [&](uint32_t j, uint32_t *stage) -> uint32_t {
		const CodecGroup *gp = a.groups + j;
		const uint32_t mode = __ldg(&gp->mode);
		if (mode != kBpFor && mode != kBpDeltaFor) {
			return 0u;
		}
		const uint32_t w = __ldg(&gp->width), cnt = __ldg(&gp->count);
		const uint8_t *src = a.images + __ldg(&gp->payload);
		const uint32_t mis = (uint32_t)((uintptr_t)src & 3u);
		const uint32_t *src4 = (const uint32_t *)(src - mis);
		const uint32_t n_words = (mis + ((cnt + 31u) / 32u) * 4u * w + 3u) / 4u;
		for (uint32_t t = tid; t < n_words; t += kCodecThreads) {
			asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr(stage + t)), "l"(src4 + t) : "memory");
		}
		if (tid < 4) {
			stage[n_words + tid] = 0u;  // the funnel's look-ahead words
		}
		return mis * 8u;
	};
	uint32_t mis_next = 0;
	if (blockIdx.x < a.n_jobs) {
		mis_next = stage_group(blockIdx.x, stage2[0]);
	}
	for (uint32_t j = blockIdx.x; j < a.n_jobs; j += gridDim.x, parity ^= 1u) {
		const CodecGroup *gp = a.groups + j;
		const uint32_t mode = __ldg(&gp->mode), w = __ldg(&gp->width), cnt = __ldg(&gp->count);
		const uint64_t frame = __ldg(&gp->frame);
		T *out = MODE == 0 ? (T *)a.out + (__ldg(&gp->out_row) - a.out_row0) : nullptr;
		uint32_t *stage = stage2[parity];
		const uint32_t mis_bits = mis_next;
		asm volatile("cp.async.wait_all;" ::: "memory");
		__syncthreads();  // group j is in stage[parity]; every thread is done with group j - 1 (stage[parity ^ 1])
		if (j + gridDim.x < a.n_jobs) {
			mis_next = stage_group(j + gridDim.x, stage2[parity ^ 1u]);
		}
		if (mode == kBpDeltaFor) {
			// blocked: thread t owns values [8t, 8t + 8); running sum of (field + for) in T arithmetic
			uint64_t d[8];
			uint64_t run = 0;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const uint32_t i = tid * 8u + k;
				const uint64_t f = i < cnt ? field_at(stage, mis_bits + i * w, w) + frame : 0ull;
				run += f;
				d[k] = run;
			}
			uint64_t incl = run;
#pragma unroll
			for (int s = 1; s < 32; s <<= 1) {
				const uint64_t o = __shfl_up_sync(0xffffffffu, incl, s);
				if (lane >= (uint32_t)s) {
					incl += o;
				}
			}
			if (lane == 31) {
				warp_tot[warp] = incl;
			}
			__syncthreads();
			uint64_t base = __ldg(&gp->delta_offset) + incl - run;
			for (uint32_t wi = 0; wi < warp; wi++) {
				base += warp_tot[wi];
			}
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const uint32_t i = tid * 8u + k;
				if (i < cnt) {
					const uint64_t v = trunc_T<S>(base + d[k]);
					if (MODE == 0) {
						out[i] = (T)v;
					} else {
						acc.add<SIGNED>(extend_low<S, SIGNED>(v));
						n_rows++;
					}
				}
			}
		} else if (mode == kBpFor) {
#pragma unroll 4
			for (uint32_t i = tid; i < cnt; i += kCodecThreads) {
				const uint64_t v = trunc_T<S>(field_at(stage, mis_bits + i * w, w) + frame);
				if (MODE == 0) {
					out[i] = (T)v;
				} else {
					acc.add<SIGNED>(extend_low<S, SIGNED>(v));
					n_rows++;
				}
			}
		} else {
			const uint64_t constant = __ldg(&gp->constant);
			for (uint32_t i = tid; i < cnt; i += kCodecThreads) {
				// CONSTANT, or (group_offset + i) * constant + frame in idx_t arithmetic, truncated to T (bitpacking.cpp:757-760)
				const uint64_t v = trunc_T<S>(mode == kBpConstant ? constant
				                                                  : (uint64_t)i * extend_low<S, SIGNED>(constant) +
				                                                        extend_low<S, SIGNED>(frame));
				if (MODE == 0) {
					out[i] = (T)v;
				} else {
					acc.add<SIGNED>(extend_low<S, SIGNED>(v));
					n_rows++;
				}
			}
		}
	}
	if (MODE == 1) {
		block_reduce_agg(acc.lo64(), acc.hi64(), (uint64_t)n_rows, red, a.agg);
	}
}

// one warp per (segment, row): BitpackingFetchRow
template <int S, bool SIGNED>
__global__ void __launch_bounds__(kCodecThreads) bitpacking_fetch_kernel(CodecArgs a, const uint32_t *seg_idx,
                                                                         const uint32_t *row_in_seg, uint64_t n) {
	using T = typename UIntOf<S>::type;
	const uint32_t lane = threadIdx.x & 31u;
	const uint64_t q = ((uint64_t)blockIdx.x * kCodecThreads + threadIdx.x) >> 5;
	if (q >= n) {
		return;
	}
	const CodecSeg seg = a.segs[seg_idx[q]];
	const uint8_t *img = a.images + seg.image_offset;
	const uint32_t row = row_in_seg[q], g = row / kGroupValues, r = row % kGroupValues;
	GroupHeader h;
	if (lane == 0) {
		h = read_group_header<S>(img, seg.count, g);
	}
	h.frame = __shfl_sync(0xffffffffu, h.frame, 0);
	h.constant = __shfl_sync(0xffffffffu, h.constant, 0);
	h.delta_offset = __shfl_sync(0xffffffffu, h.delta_offset, 0);
	h.mode = __shfl_sync(0xffffffffu, h.mode, 0);
	h.width = __shfl_sync(0xffffffffu, h.width, 0);
	h.data_off = __shfl_sync(0xffffffffu, h.data_off, 0);
	const uint8_t *src = img + h.data_off;
	const uint32_t w = h.width;
	auto field = [&](uint32_t i) -> uint64_t {
		const uint64_t bit = (uint64_t)i * w;
		const uint8_t *p = src + (bit >> 3);
		const uint32_t sh = (uint32_t)(bit & 7u);
		const uint32_t need = (sh + w + 7u) >> 3;  // <= 9 bytes
		uint64_t lo = load_bytes(p, need < 8 ? (int)need : 8);
		uint64_t v = lo >> sh;
		if (need > 8) {
			v |= (uint64_t)__ldg(p + 8) << (64u - sh);
		}
		return v & lo_set64(w);
	};
	uint64_t v = 0;
	if (h.mode == kBpConstant) {
		v = h.constant;
	} else if (h.mode == kBpConstantDelta) {
		v = (uint64_t)r * extend_low<S, SIGNED>(h.constant) + extend_low<S, SIGNED>(h.frame);
	} else if (h.mode == kBpFor) {
		v = lane == 0 ? field(r) + h.frame : 0ull;
	} else {
		uint64_t part = 0;
		for (uint32_t i = lane; i <= r; i += 32u) {
			part += field(i) + h.frame;
		}
#pragma unroll
		for (int s = 16; s > 0; s >>= 1) {
			part += __shfl_down_sync(0xffffffffu, part, s);
		}
		v = part + h.delta_offset;
	}
	if (lane == 0) {
		((T *)a.out)[q] = (T)trunc_T<S>(v);
	}
}

// ---- RLE ----
// exclusive prefix sums of the run lengths of every segment (one CTA per segment): run_start[first_run + k]
__global__ void __launch_bounds__(kCodecThreads) rle_run_starts_kernel(CodecArgs a, uint32_t seg_b, uint32_t n_segs) {
	__shared__ uint32_t warp_tot[kCodecThreads / 32];
	__shared__ uint32_t carry;
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	for (uint32_t s = seg_b + blockIdx.x; s < seg_b + n_segs; s += gridDim.x) {
		const CodecSeg seg = a.segs[s];
		const uint8_t *img = a.images + seg.image_offset;
		const uint64_t counts_off = *(const uint64_t *)img;
		const uint16_t *counts = (const uint16_t *)(img + counts_off);
		const uint32_t n_runs = (seg.bytes - (uint32_t)counts_off) / 2u;
		if (tid == 0) {
			carry = 0;
		}
		__syncthreads();
		for (uint32_t base = 0; base < n_runs; base += kCodecThreads) {
			const uint32_t k = base + tid;
			const uint32_t c = k < n_runs ? (uint32_t)__ldg(counts + k) : 0u;
			uint32_t incl = c;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
				if (lane >= (uint32_t)d) {
					incl += o;
				}
			}
			if (lane == 31) {
				warp_tot[warp] = incl;
			}
			__syncthreads();
			uint32_t off = carry;
			for (uint32_t wi = 0; wi < warp; wi++) {
				off += warp_tot[wi];
			}
			if (k < n_runs) {
				a.run_start[seg.first_run + k] = off + incl - c;
			}
			__syncthreads();
			if (tid == kCodecThreads - 1) {
				carry = off + incl;
			}
			__syncthreads();
		}
	}
}

__device__ __forceinline__ uint32_t rle_find_run(const uint32_t *starts, uint32_t n_runs, uint32_t row) {
	uint32_t lo = 0, hi = n_runs;  // last run whose start <= row
	while (hi - lo > 1) {
		const uint32_t mid = (lo + hi) >> 1;
		if (__ldg(starts + mid) <= row) {
			lo = mid;
		} else {
			hi = mid;
		}
	}
	return lo;
}

// decode 2048-row tiles: the heads of the runs that start inside the tile are scattered to their row, an inclusive
// max-scan over the 2048 rows gives every row its run (RLEScanPartial, rle.cpp:298-318, without the per-row walk)
struct RleTile {
	uint32_t cnt, run_lo, run_hi, first;
	uint64_t payload, out_row;
};
__device__ __forceinline__ RleTile load_rle_tile(const CodecGroup *gp) {
	RleTile t;
	t.cnt = __ldg(&gp->count);
	t.run_lo = __ldg(&gp->run_lo);
	t.run_hi = __ldg(&gp->run_hi);
	t.first = __ldg(&gp->first);
	t.payload = __ldg(&gp->payload);
	t.out_row = __ldg(&gp->out_row);
	return t;
}

template <int S>
__global__ void __launch_bounds__(kCodecThreads) rle_decode_kernel(CodecArgs a) {
	using T = typename UIntOf<S>::type;
	__shared__ __align__(16) uint32_t head[kGroupValues];  // run (relative to the tile's first run) that starts at this row, else 0
	__shared__ uint32_t warp_max[kCodecThreads / 32];
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	// the record of the NEXT tile is loaded while the current one is expanded (one global latency less per tile), and
	// its run starts are pulled towards L2 (ncu: long-scoreboard 21.6 of 28 stall cycles per issue before)
	RleTile next;
	next.cnt = 0;
	if (blockIdx.x < a.n_jobs) {
		next = load_rle_tile(a.groups + blockIdx.x);
	}
	for (uint32_t j = blockIdx.x; j < a.n_jobs; j += gridDim.x) {
		const RleTile t = next;
		const Job job = a.jobs[j];
		const uint32_t seg_first_run = __ldg(&a.segs[job.seg].first_run);
		if (j + gridDim.x < a.n_jobs) {
			next = load_rle_tile(a.groups + j + gridDim.x);
		}
		if (t.cnt == 0) {
			continue;
		}
		const uint32_t cnt = t.cnt, run_lo = t.run_lo, run_hi = t.run_hi, first = t.first;
		const T *values = (const T *)(a.images + t.payload);
		T *out = (T *)a.out + (t.out_row - a.out_row0);
		__syncthreads();  // the previous tile's head[] has been consumed
		// 16-byte shared accesses: a thread's 8 consecutive words as two vectors (word-by-word they are 8-way conflicts)
		*(uint4 *)&head[tid * 8u] = make_uint4(0u, 0u, 0u, 0u);
		*(uint4 *)&head[tid * 8u + 4u] = make_uint4(0u, 0u, 0u, 0u);
		__syncthreads();
		for (uint32_t r = run_lo + 1u + tid; r <= run_hi; r += kCodecThreads) {
			const uint32_t pos = __ldg(a.run_start + r) - first;  // > 0: run_lo is the last run that starts at or before `first`
			if (pos < cnt) {
				// a run of 65535 rows is followed by a ZERO-length run when the next value differs (RLEState::Update,
				// rle.cpp:66-72 then :53-60): such runs share a row with their successor, the last one holds it
				atomicMax(&head[pos], r - run_lo);
			}
		}
		if (next.cnt) {
			// the next tile's run starts: 4 bytes per run, one prefetch per 128-byte line
			for (uint32_t r = next.run_lo + tid * 32u; r <= next.run_hi; r += kCodecThreads * 32u) {
				asm volatile("prefetch.global.L2 [%0];" ::"l"(a.run_start + r));
			}
		}
		__syncthreads();
		// thread t owns rows [8t, 8t + 8)
		const uint4 h0 = *(const uint4 *)&head[tid * 8u], h1 = *(const uint4 *)&head[tid * 8u + 4u];
		uint32_t m[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
		for (int k = 1; k < 8; k++) {
			m[k] = max(m[k], m[k - 1]);
		}
		uint32_t incl = m[7];
#pragma unroll
		for (int s = 1; s < 32; s <<= 1) {
			const uint32_t o = __shfl_up_sync(0xffffffffu, incl, s);
			if (lane >= (uint32_t)s) {
				incl = max(incl, o);
			}
		}
		if (lane == 31) {
			warp_max[warp] = incl;
		}
		uint32_t before = __shfl_up_sync(0xffffffffu, incl, 1);
		if (lane == 0) {
			before = 0u;
		}
		__syncthreads();
		for (uint32_t wi = 0; wi < warp; wi++) {
			before = max(before, warp_max[wi]);
		}
		// the run index of every row goes back to shared memory so that the stores are row-strided (coalesced)
		*(uint4 *)&head[tid * 8u] = make_uint4(max(before, m[0]), max(before, m[1]), max(before, m[2]), max(before, m[3]));
		*(uint4 *)&head[tid * 8u + 4u] = make_uint4(max(before, m[4]), max(before, m[5]), max(before, m[6]), max(before, m[7]));
		__syncthreads();
		const T *tile_values = values + (run_lo - seg_first_run);
#pragma unroll 4
		for (uint32_t i = tid; i < cnt; i += kCodecThreads) {
			out[i] = __ldg(tile_values + head[i]);
		}
	}
}

// SUM = sum_runs value * length (job = 2048 runs of a segment)
template <int S, bool SIGNED, int MODE>
__global__ void __launch_bounds__(kCodecThreads) rle_kernel(CodecArgs a) {
	using T = typename UIntOf<S>::type;
	__shared__ uint64_t red[96];
	const uint32_t tid = threadIdx.x;
	uint64_t lo = 0, hi = 0, rows = 0;
	for (uint32_t j = blockIdx.x; j < a.n_jobs; j += gridDim.x) {
		const Job job = a.jobs[j];
		const CodecSeg seg = a.segs[job.seg];
		const uint8_t *img = a.images + seg.image_offset;
		const uint64_t counts_off = *(const uint64_t *)img;
		const uint32_t n_runs = (seg.bytes - (uint32_t)counts_off) / 2u;
		const T *values = (const T *)(img + 8);
		{
			const uint16_t *counts = (const uint16_t *)(img + counts_off);
			const uint32_t first = job.k * kGroupValues;
			for (uint32_t k = first + tid; k < min(n_runs, first + kGroupValues); k += kCodecThreads) {
				const uint64_t v = extend_low<S, SIGNED>((uint64_t)__ldg(values + k));
				const uint64_t c = (uint64_t)__ldg(counts + k);
				// 128-bit v * c (c < 2^16)
				uint64_t plo = v * c;
				uint64_t phi = __umul64hi(v, c);
				if (SIGNED && (int64_t)v < 0) {
					phi -= c;
				}
				add128(lo, hi, plo, phi);
				rows += c;
			}
		}
	}
	block_reduce_agg(lo, hi, rows, red, a.agg);
}

template <int S, bool SIGNED>
__global__ void __launch_bounds__(kCodecThreads) rle_fetch_kernel(CodecArgs a, const uint32_t *seg_idx,
                                                                  const uint32_t *row_in_seg, uint64_t n) {
	using T = typename UIntOf<S>::type;
	const uint64_t q = (uint64_t)blockIdx.x * kCodecThreads + threadIdx.x;
	if (q >= n) {
		return;
	}
	const CodecSeg seg = a.segs[seg_idx[q]];
	const uint8_t *img = a.images + seg.image_offset;
	const uint64_t counts_off = *(const uint64_t *)img;
	const uint32_t n_runs = (seg.bytes - (uint32_t)counts_off) / 2u;
	((T *)a.out)[q] = __ldg((const T *)(img + 8) + rle_find_run(a.run_start + seg.first_run, n_runs, row_in_seg[q]));
}

template <int S, bool SIGNED>
cudaError_t launch_typed(int codec, int mode, const CodecArgs &a, uint32_t grid, cudaStream_t stream) {
	if (codec == 0) {
		if (mode == 0) {
			bitpacking_groups_kernel<S, SIGNED, 0><<<grid, kCodecThreads, 0, stream>>>(a);
		} else {
			bitpacking_groups_kernel<S, SIGNED, 1><<<grid, kCodecThreads, 0, stream>>>(a);
		}
	} else {
		if (mode == 0) {
			rle_decode_kernel<S><<<grid, kCodecThreads, 0, stream>>>(a);
		} else {
			rle_kernel<S, SIGNED, 1><<<grid, kCodecThreads, 0, stream>>>(a);
		}
	}
	return cudaGetLastError();
}

template <int S, bool SIGNED>
cudaError_t groups_typed(int codec, const CodecArgs &a, CodecGroup *groups, cudaStream_t stream) {
	codec_groups_kernel<S><<<(a.n_jobs + kCodecThreads - 1) / kCodecThreads, kCodecThreads, 0, stream>>>(a, groups, codec);
	return cudaGetLastError();
}

template <int S, bool SIGNED>
cudaError_t fetch_typed(int codec, const CodecArgs &a, const uint32_t *seg_idx, const uint32_t *row_in_seg, uint64_t n,
                        cudaStream_t stream) {
	if (codec == 0) {
		const uint64_t blocks = (n * 32 + kCodecThreads - 1) / kCodecThreads;
		bitpacking_fetch_kernel<S, SIGNED><<<(uint32_t)blocks, kCodecThreads, 0, stream>>>(a, seg_idx, row_in_seg, n);
	} else {
		const uint64_t blocks = (n + kCodecThreads - 1) / kCodecThreads;
		rle_fetch_kernel<S, SIGNED><<<(uint32_t)blocks, kCodecThreads, 0, stream>>>(a, seg_idx, row_in_seg, n);
	}
	return cudaGetLastError();
}

#define CODEC_DISPATCH(fn, ...)                                                                             \
	switch (t.size * 2 + (t.is_signed ? 1 : 0)) {                                                           \
	case 2: return fn<1, false>(__VA_ARGS__);                                                               \
	case 3: return fn<1, true>(__VA_ARGS__);                                                                \
	case 4: return fn<2, false>(__VA_ARGS__);                                                               \
	case 5: return fn<2, true>(__VA_ARGS__);                                                                \
	case 8: return fn<4, false>(__VA_ARGS__);                                                               \
	case 9: return fn<4, true>(__VA_ARGS__);                                                                \
	case 16: return fn<8, false>(__VA_ARGS__);                                                              \
	case 17: return fn<8, true>(__VA_ARGS__);                                                               \
	default: return cudaErrorInvalidValue;                                                                  \
	}

}  // namespace

cudaError_t launch_codec_scan(TypeInfo t, int codec, int mode, const CodecArgs &a, int sm_count, cudaStream_t stream) {
	if (a.n_jobs == 0) {
		return cudaSuccess;
	}
	// 2048 values per CTA iteration: the tiles are latency-bound, so as many CTAs as fit (8 x 256 threads per SM; the
	// 8-byte bitpacking stage is 2 x 16 KiB, six of those fit)
	const uint32_t per_sm = (codec == 0 && t.size == 8) ? 6u : 8u;
	const uint32_t resident = (uint32_t)sm_count * per_sm;
	const uint32_t grid = a.n_jobs < resident ? a.n_jobs : resident;
	CODEC_DISPATCH(launch_typed, codec, mode, a, grid, stream)
}

cudaError_t launch_codec_groups(TypeInfo t, int codec, const CodecArgs &a, CodecGroup *groups, cudaStream_t stream) {
	if (a.n_jobs == 0) {
		return cudaSuccess;
	}
	CODEC_DISPATCH(groups_typed, codec, a, groups, stream)
}

cudaError_t launch_codec_fetch(TypeInfo t, int codec, const CodecArgs &a, const uint32_t *seg_idx, const uint32_t *row_in_seg,
                               uint64_t n, cudaStream_t stream) {
	if (n == 0) {
		return cudaSuccess;
	}
	CODEC_DISPATCH(fetch_typed, codec, a, seg_idx, row_in_seg, n, stream)
}

cudaError_t launch_rle_run_starts(const CodecArgs &a, uint32_t seg_b, uint32_t n_segs, int sm_count, cudaStream_t stream) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	const uint32_t resident = (uint32_t)sm_count * 4u;
	rle_run_starts_kernel<<<n_segs < resident ? n_segs : resident, kCodecThreads, 0, stream>>>(a, seg_b, n_segs);
	return cudaGetLastError();
}

}  // namespace adacom
