// kernels_scan.cu -- width-specialised fused scan+SUM over packed words (the "decode+scan" headline path).
//
// The reference decodes a vector with one scalar bits::read_int per value (succinct.cpp:123-144) and DuckDB's
// SUM then walks the materialised vector.  On a B200 the per-value formulation is ISSUE-bound below ~20 bits
// (at w bits per value the HBM roofline asks each SM for 184/w values per cycle, the generic extract + 96-bit
// accumulate costs ~12 instructions per value).  This kernel never materialises values:
//
//   * sum(decoded) = sum(fields) + count * extend(min_factor) whenever the decoded interval
//     [min, min + 2^w - 1] stays inside one sign/truncation zone of T (always true for data the reference can
//     represent; the host checks it per segment and routes the rest to the generic decode_kernel);
//   * the field stream has period P = w / gcd(w, 64) words holding V = 64 / gcd(w, 64) values; a lane owns
//     whole periods, so every shift and mask is a compile-time constant (template on w, 63 instantiations):
//     2 instructions per extraction for w <= 32, IADD3 folds two fields per add; w = 1, 2, 4, 8, 16 use
//     popc / SWAR + dp4a / dp2a;
//   * lanes take consecutive periods: the lane stride in shared memory is P words with P odd, which makes the
//     64-bit shared loads conflict-free;
//   * packed chunks (<= 16 KiB, 16-byte aligned by construction) are staged by the TMA engine
//     (cp.async.bulk + mbarrier) into a 4-deep ring by one elected thread, as in decode_kernel.
//
// Algorithmic HBM bytes: P_chunk read per chunk; the output is one atomic triple per CTA.
#include "adacom_kernels.cuh"

namespace adacom {

namespace {

constexpr int kConsumerWarps = 8;
constexpr int kConsumers = kConsumerWarps * 32;
constexpr int kThreads = kConsumers + 32;  // + one producer warp feeding the TMA ring
constexpr int kStages = 4;
constexpr uint32_t kStageBytes = 16384 + 32;

struct ChunkInfo {
	uint64_t ext_min;  // fast: extend_low<S,SIGNED>(min_factor) as a 64-bit pattern; slow: the value to add (or 0)
	uint32_t cnt;
	uint32_t w;
	uint32_t min_is_negative;  // fast: ext_min is to be read as a negative int64 (signed T)
	uint32_t bit_base;         // slow: bit position of the first field inside the staged bytes (0..127)
	uint32_t fast;
	uint32_t done;             // sentinel: the producer has no more chunks for this CTA
};

__host__ __device__ constexpr int gcd64(int w) {
	int g = 1;
	while (g < 64 && (w % (g * 2)) == 0) {
		g *= 2;
	}
	return g;
}

__device__ __forceinline__ uint32_t dp4a_ones(uint32_t x, uint32_t acc) {
	uint32_t r;
	asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(0x01010101u), "r"(acc));
	return r;
}

__device__ __forceinline__ uint32_t dp2a_ones(uint32_t x, uint32_t acc) {
	uint32_t r;
	asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(0x00000101u), "r"(acc));
	return r;
}

// Periods in a full chunk of width w.
__host__ __device__ constexpr int periods_per_chunk(int w) {
	return (w >= 43 ? 2048 : 128 * (1024 / w)) / (64 / gcd64(w));
}

// A period is split into Q sub-periods (power of two) so that a full chunk offers at least one piece of work
// per consumer thread; all lanes of a warp take the SAME sub-period index, so the code stays convergent and
// every shift is still a compile-time constant.
__host__ __device__ constexpr int subperiods(int w) {
	int q = 1;
	while (periods_per_chunk(w) * q < kConsumers && q < 64 / gcd64(w)) {
		q *= 2;
	}
	return q;
}

// Sum of the fields [q*V/Q, (q+1)*V/Q) of the period that starts at word p.
template <int W, int Q, int q>
__device__ __forceinline__ void sum_subperiod(const uint64_t *p, Acc96 &acc) {
	constexpr int G = gcd64(W);
	constexpr int V = 64 / G;
	constexpr int VQ = V / Q;
	constexpr int J0 = q * VQ;
	constexpr int K0 = (J0 * W) >> 6;                    // first 64-bit word touched
	constexpr int K1 = ((J0 + VQ) * W - 1) >> 6;         // last 64-bit word touched
	constexpr int NW = K1 - K0 + 1;
	constexpr int BASE = K0 * 64;
	uint32_t r[2 * NW + 2];
#pragma unroll
	for (int k = 0; k < NW; k++) {
		uint64_t x = p[K0 + k];
		r[2 * k] = (uint32_t)x;
		r[2 * k + 1] = (uint32_t)(x >> 32);
	}
	r[2 * NW] = 0;
	r[2 * NW + 1] = 0;
	if (W == 1) {
		acc.add<false>((uint64_t)(__popc(r[0]) + __popc(r[1])));
	} else if (W == 2) {
		uint32_t s = 0;
#pragma unroll
		for (int k = 0; k < 2; k++) {
			uint32_t t = (r[k] & 0x33333333u) + ((r[k] >> 2) & 0x33333333u);
			t = (t & 0x0F0F0F0Fu) + ((t >> 4) & 0x0F0F0F0Fu);
			s = dp4a_ones(t, s);
		}
		acc.add<false>((uint64_t)s);
	} else if (W == 4) {
		uint32_t s = 0;
#pragma unroll
		for (int k = 0; k < 2; k++) {
			uint32_t t = (r[k] & 0x0F0F0F0Fu) + ((r[k] >> 4) & 0x0F0F0F0Fu);
			s = dp4a_ones(t, s);
		}
		acc.add<false>((uint64_t)s);
	} else if (W == 8) {
		acc.add<false>((uint64_t)dp4a_ones(r[1], dp4a_ones(r[0], 0u)));
	} else if (W == 16) {
		acc.add<false>((uint64_t)dp2a_ones(r[1], dp2a_ones(r[0], 0u)));
	} else if (W <= 32) {
		constexpr uint32_t mask = W == 32 ? 0xFFFFFFFFu : ((1u << (W & 31)) - 1u);
		uint64_t s64 = 0;
		uint32_t s32 = 0;  // VQ * (2^W - 1) < 2^32 when W <= 26
#pragma unroll
		for (int j = 0; j < VQ; j++) {
			const int pos = (J0 + j) * W - BASE, i = pos >> 5, off = pos & 31;
			uint32_t f;
			if (off + W <= 32) {
				f = (off + W == 32) ? (r[i] >> off) : ((r[i] >> off) & mask);
			} else {
				f = __funnelshift_r(r[i], r[i + 1], off) & mask;
			}
			if (W <= 26) {
				s32 += f;
			} else {
				s64 += f;
			}
		}
		acc.add<false>(W <= 26 ? (uint64_t)s32 : s64);
	} else {
		constexpr uint32_t mask_hi = (1u << ((W - 32) & 31)) - 1u;
#pragma unroll
		for (int j = 0; j < VQ; j++) {
			const int pos = (J0 + j) * W - BASE, i = pos >> 5, off = pos & 31;
			uint32_t lo, hi;
			if (off == 0) {
				lo = r[i];
				hi = r[i + 1] & mask_hi;
			} else {
				lo = __funnelshift_r(r[i], r[i + 1], off);
				hi = __funnelshift_r(r[i + 1], r[i + 2], off) & mask_hi;
			}
			acc.add<false>(((uint64_t)hi << 32) | lo);
		}
	}
}

template <int W, int Q, int q>
struct SubSwitch {
	static __device__ __forceinline__ void run(uint32_t sub, const uint64_t *p, Acc96 &acc) {
		if (sub == (uint32_t)q) {
			sum_subperiod<W, Q, q>(p, acc);
		} else {
			SubSwitch<W, Q, q + 1>::run(sub, p, acc);
		}
	}
};
template <int W, int Q>
struct SubSwitch<W, Q, Q> {
	static __device__ __forceinline__ void run(uint32_t, const uint64_t *, Acc96 &) {
	}
};

template <int W>
__device__ __forceinline__ void sum_chunk_fields(const uint64_t *words, uint32_t cnt, uint32_t tid, Acc96 &acc) {
	constexpr int G = gcd64(W);
	constexpr int P = W / G;
	constexpr int V = 64 / G;
	constexpr int Q = subperiods(W);
	const uint32_t n_units = cnt / V;
	const uint32_t n_rows = ((n_units + 31u) >> 5) * Q;  // a row = 32 periods x one sub-period index
	const uint32_t lane = tid & 31u;
	for (uint32_t row = tid >> 5; row < n_rows; row += kConsumerWarps) {
		const uint32_t sub = row % Q;               // warp-uniform
		const uint32_t u = (row / Q) * 32u + lane;
		if (u < n_units) {
			SubSwitch<W, Q, 0>::run(sub, words + (size_t)u * P, acc);
		}
	}
	// values behind the last whole period (only the last chunk of a segment has any)
	const uint32_t done = n_units * V;
	const uint32_t i = done + tid;
	if (i < cnt) {
		const uint32_t bit = i * W;
		uint64_t f = W <= 32 ? (uint64_t)(read_bits32((const uint32_t *)words, bit) & (uint32_t)lo_set64(W))
		                     : (read_bits64(words, bit) & lo_set64(W));
		acc.add<false>(f);
	}
}

// A full chunk of width w holds periods_per_chunk(w) periods.  The period-per-lane path needs at
// least one period per consumer thread to keep all warps busy; the other widths (odd w >= 9, ...) have a
// large per-value instruction budget (0.7 * w) and use the fully parallel lane-per-value extraction.
__host__ __device__ constexpr bool period_path(int w) {
	return w >= 1 && w <= 63;  // measured: the unrolled period path beats lane-per-value at every width
}

// lane-per-value: adjacent lanes read adjacent fields (conflict-free), runtime shifts
__device__ __forceinline__ void sum_chunk_generic(const uint64_t *words, uint32_t w, uint32_t cnt, uint32_t tid,
                                                  Acc96 &acc) {
	if (w <= 32u) {
		const uint32_t *words32 = (const uint32_t *)words;
		const uint32_t mask = (uint32_t)lo_set64(w);
		uint64_t s = 0;
#pragma unroll 4
		for (uint32_t i = tid; i < cnt; i += kConsumers) {
			s += read_bits32(words32, i * w) & mask;
		}
		acc.add<false>(s);
	} else {
		const uint64_t mask = lo_set64(w);
#pragma unroll 4
		for (uint32_t i = tid; i < cnt; i += kConsumers) {
			acc.add<false>(read_bits64(words, i * w) & mask);
		}
	}
}

// not re-packed: the payload is the slot array itself -- aligned loads of T, sign/zero-extended like DuckDB reads
// them; `add` (literal mode only) is added before the truncation to T (succinct.cpp:138-143)
template <typename T, typename WideT>
__device__ __forceinline__ void sum_raw_typed(const uint8_t *p, uint32_t cnt, uint64_t add, uint32_t tid, Acc96 &acc) {
	const T *v = (const T *)p;
	constexpr bool kSigned = ((WideT)-1) < 0;
	if (sizeof(T) == 8) {
#pragma unroll 4
		for (uint32_t i = tid; i < cnt; i += kConsumers) {
			acc.add<kSigned>((uint64_t)(WideT)(T)(v[i] + (T)add));
		}
	} else {
		WideT s = 0;  // <= 2^14 values of <= 32 bits per thread and chunk: no overflow
#pragma unroll 4
		for (uint32_t i = tid; i < cnt; i += kConsumers) {
			s += (WideT)(T)(v[i] + (T)add);
		}
		acc.add<kSigned>((uint64_t)s);
	}
}

__device__ __forceinline__ void sum_raw_slots(const uint8_t *p, uint32_t cnt, uint64_t add, int type_size, int is_signed,
                                              uint32_t tid, Acc96 &acc, Acc96 &sacc) {
	if (is_signed) {
		switch (type_size) {
		case 1: sum_raw_typed<int8_t, int64_t>(p, cnt, add, tid, sacc); break;
		case 2: sum_raw_typed<int16_t, int64_t>(p, cnt, add, tid, sacc); break;
		case 4: sum_raw_typed<int32_t, int64_t>(p, cnt, add, tid, sacc); break;
		default: sum_raw_typed<int64_t, int64_t>(p, cnt, add, tid, sacc); break;
		}
	} else {
		switch (type_size) {
		case 1: sum_raw_typed<uint8_t, uint64_t>(p, cnt, add, tid, acc); break;
		case 2: sum_raw_typed<uint16_t, uint64_t>(p, cnt, add, tid, acc); break;
		case 4: sum_raw_typed<uint32_t, uint64_t>(p, cnt, add, tid, acc); break;
		default: sum_raw_typed<uint64_t, uint64_t>(p, cnt, add, tid, acc); break;
		}
	}
}

template <int W0>
struct WidthSwitch {
	static __device__ __forceinline__ void run(uint32_t w, const uint64_t *words, uint32_t cnt, uint32_t tid, Acc96 &acc) {
		if (period_path(W0) && w == W0) {
			sum_chunk_fields<W0>(words, cnt, tid, acc);
		} else {
			WidthSwitch<W0 + 1>::run(w, words, cnt, tid, acc);
		}
	}
};
template <>
struct WidthSwitch<64> {
	static __device__ __forceinline__ void run(uint32_t w, const uint64_t *words, uint32_t cnt, uint32_t tid, Acc96 &acc) {
		sum_chunk_generic(words, w, cnt, tid, acc);
	}
};

struct ResolvedChunk {
	ChunkInfo info;
	uint64_t src16;
	uint32_t bytes;  // 0: this (segment, chunk index) does not exist
};

__device__ __forceinline__ uint64_t extend_runtime(uint64_t x, int type_size, int is_signed) {
	const uint32_t shift = 64u - 8u * (uint32_t)type_size;
	return is_signed ? (uint64_t)(((int64_t)(x << shift)) >> shift) : ((x << shift) >> shift);
}

// sum(decoded) == sum(fields) + count * extend(min) iff every decoded value min + field stays inside the sign /
// truncation zone of T that min lies in.  Fields never exceed max - min when the descriptor carries the analyzed
// maximum (they are (v - min) & mask for v in [min, max]); otherwise the bound is the full field range 2^w - 1.
__device__ __forceinline__ bool linear_zone_dev(uint64_t min_factor, uint64_t max_factor, bool analyzed, uint32_t width,
                                                bool repacked, int type_size, int is_signed) {
	if (!repacked || width >= 64) {
		return false;
	}
	const uint64_t lo = min_factor == ~0ull ? 0 : min_factor;
	uint64_t range = (1ull << width) - 1;
	if (analyzed && min_factor != ~0ull && max_factor >= lo && max_factor - lo <= range) {
		range = max_factor - lo;
	}
	const uint64_t hi = lo + range;
	if (hi < lo) {
		return false;
	}
	const uint64_t e_lo = extend_runtime(lo, type_size, is_signed), e_hi = extend_runtime(hi, type_size, is_signed);
	if (e_hi - e_lo != range) {
		return false;
	}
	return is_signed ? (int64_t)e_lo <= (int64_t)e_hi : e_lo <= e_hi;
}

// Work decomposition without any table: job j = (segment seg_b + j / kScanParts, chunk index j % kScanParts).
// A segment has at most kScanParts chunks (16 KiB stages vs <= 256 KiB segments); jobs whose chunk does not
// exist cost one descriptor read by one producer lane.  The launch uses an odd grid so that every CTA's strided
// job sequence cycles through all chunk indices (even load).
__device__ __forceinline__ uint32_t job_segment(const ScanSumArgs &a, uint32_t job, uint32_t parts) {
	return a.seg_b + job / parts;
}

__device__ __forceinline__ ResolvedChunk resolve_chunk(const ScanSumArgs &a, uint32_t job, uint32_t parts, const DescQ &dq) {
	ResolvedChunk r;
	const uint32_t k = job % parts;
	const uint64_t row_start = dq.row_start();
	const uint64_t packed_ptr = dq.packed_ptr();
	const uint64_t mn = dq.min_factor();
	const uint32_t count = dq.count();
	const uint32_t w = dq.width();
	const bool repacked = dq.state() & kSegRepacked;
	const uint32_t gran = scan_chunk_values(w);
	const uint32_t first = k * gran;
	r.bytes = 0;
	r.src16 = 0;
	if (first >= count) {
		return r;
	}
	const uint32_t cnt = min(gran, count - first);
	const uint64_t src = repacked ? packed_ptr + (((uint64_t)first * w) >> 3)
	                              : (uint64_t)a.raw + (row_start + first) * (uint64_t)a.type_size;
	r.src16 = src & ~15ull;
	const uint32_t byte_base = (uint32_t)(src - r.src16);
	r.bytes = (byte_base + ((cnt * w + 7u) >> 3) + 15u) & ~15u;
	const bool add = (mn != ~0ull) && (repacked || a.literal);  // succinct.cpp:138-140 (+ Q1 in literal mode)
	// fast: 1 = field sums (linear zone), 2 = raw slots of T (not re-packed), 0 = generic per-value extraction
	const uint32_t fast = !repacked ? 2u
	                                : (linear_zone_dev(mn, dq.max_factor(), dq.state() & kSegAnalyzed, w, repacked,
	                                                   a.type_size, a.is_signed)
	                                       ? 1u
	                                       : 0u);
	uint64_t ext = 0;
	uint32_t neg = 0;
	if (fast == 1u) {
		if (add) {
			ext = extend_runtime(mn, a.type_size, a.is_signed);
			neg = a.is_signed && (int64_t)ext < 0;
		}
	} else {
		ext = add ? mn : 0ull;
	}
	r.info.ext_min = ext;
	r.info.cnt = cnt;
	r.info.w = w;
	r.info.min_is_negative = neg;
	r.info.bit_base = byte_base * 8u;
	r.info.fast = fast;
	r.info.done = 0;
	return r;
}

__global__ void __launch_bounds__(kThreads, 3) scan_sum_fields_kernel(const ScanSumArgs a) {
	extern __shared__ __align__(128) uint8_t stage_mem[];
	__shared__ __align__(8) uint64_t full_bar[kStages];
	__shared__ __align__(8) uint64_t empty_bar[kStages];
	__shared__ ChunkInfo infos[kStages];
	__shared__ uint64_t red[96];
	const uint32_t tid = threadIdx.x;
	const uint32_t lane = tid & 31u;
	const uint32_t warp = tid >> 5;
	const uint32_t n_units = a.seg_e - a.seg_b;
	// chunks per segment the column currently needs (maintained on the device by layout / uncompact; <= 16)
	const uint32_t parts = min(max(__ldg(a.parts), 1u), kScanParts);
	const uint32_t n_jobs = n_units * parts;

	if (tid == 0) {
		for (int s = 0; s < kStages; s++) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&empty_bar[s], kConsumerWarps);
		}
		mbar_fence_init();
	}
	__syncthreads();

	Acc96 acc;              // sum of the fields (unsigned) / of unsigned slow-path values
	Acc96 sacc;             // slow path, signed T: sign-extended values
	uint64_t corr_lo = 0;   // thread 0: sum over fast chunks of cnt * ext_min (128-bit two's complement)
	uint64_t corr_hi = 0;
	uint64_t n_values = 0;

	if (warp == kConsumerWarps) {
		// ---------------- producer warp: every lane resolves one job, the existing chunks are issued in order ----
		uint32_t seq = 0;  // stage sequence number, uniform across the warp
		const uint32_t batch = 32u * gridDim.x;
		DescQ dq_next;
		{
			const uint32_t job = blockIdx.x + lane * gridDim.x;
			if (job < n_jobs) {
				dq_next = load_desc(a.desc + job_segment(a, job, parts));
			}
		}
		for (uint32_t j0 = blockIdx.x; j0 < n_jobs; j0 += batch) {
			const uint32_t job = j0 + lane * gridDim.x;
			const DescQ dq = dq_next;
			if (job + batch < n_jobs) {  // descriptors of the next batch: in flight while this batch is issued
				dq_next = load_desc(a.desc + job_segment(a, job + batch, parts));
			}
			ResolvedChunk r;
			r.bytes = 0;
			if (job < n_jobs) {
				r = resolve_chunk(a, job, parts, dq);
			}
			// publish in job order, one lane at a time: the slot's previous user (seq - kStages) has then
			// already been issued, so the parity wait below is never more than one phase ahead
			uint32_t todo = __ballot_sync(0xffffffffu, r.bytes != 0);
			while (todo) {
				const uint32_t l = (uint32_t)__ffs((int)todo) - 1u;
				todo &= todo - 1u;
				if (lane == l) {
					const uint32_t s = seq % kStages;
					mbar_wait(&empty_bar[s], ((seq / kStages) & 1u) ^ 1u);
					infos[s] = r.info;
					mbar_arrive_expect_tx(&full_bar[s], r.bytes);
					tma_load_1d(stage_mem + s * kStageBytes, (const void *)r.src16, r.bytes, &full_bar[s]);
				}
				seq++;
				__syncwarp();
			}
		}
		if (lane == 0) {
			const uint32_t s = seq % kStages;
			mbar_wait(&empty_bar[s], ((seq / kStages) & 1u) ^ 1u);
			infos[s].done = 1;
			mbar_arrive(&full_bar[s]);
		}
	} else {
		for (uint32_t it = 0;; it++) {
			const uint32_t s = it % kStages;
			mbar_wait(&full_bar[s], (it / kStages) & 1u);
			const ChunkInfo in = infos[s];
			if (in.done) {
				break;
			}
			const uint64_t *words = (const uint64_t *)(stage_mem + s * kStageBytes);
			if (in.fast == 2u) {
				// raw slots of T (segment not re-packed, Q1: `min` is only added in literal mode)
				sum_raw_slots((const uint8_t *)words + (in.bit_base >> 3), in.cnt, in.ext_min, a.type_size, a.is_signed, tid,
				              acc, sacc);
			} else if (in.fast) {
				WidthSwitch<1>::run(in.w, words, in.cnt, tid, acc);
				if (tid == 0) {
					const uint64_t c = in.cnt;
					uint64_t lo = in.ext_min * c;
					uint64_t hi = __umul64hi(in.ext_min, c);
					if (in.min_is_negative) {
						hi -= c;  // signed x unsigned high part
					}
					add128(corr_lo, corr_hi, lo, hi);
				}
			} else {
				// generic path: raw (not re-packed) segments, or a decoded interval that crosses a sign /
				// truncation boundary of T -- one value per lane per step, extended like DuckDB sees it
				const uint32_t w = in.w;
				const uint64_t mask = lo_set64(w);
				for (uint32_t i = tid; i < in.cnt; i += kConsumers) {
					const uint32_t bit = in.bit_base + i * w;
					uint64_t f = w <= 32u ? (uint64_t)(read_bits32((const uint32_t *)words, bit) & (uint32_t)mask)
					                      : (read_bits64(words, bit) & mask);
					uint64_t v = extend_runtime(f + in.ext_min, a.type_size, a.is_signed);
					if (a.is_signed) {
						sacc.add<true>(v);
					} else {
						acc.add<false>(v);
					}
				}
			}
			if (tid == 0) {
				n_values += in.cnt;
			}
			__syncwarp();
			if (lane == 0) {
				mbar_arrive(&empty_bar[s]);
			}
		}
	}
	uint64_t lo = acc.lo64(), hi = (uint64_t)acc.a2;
	add128(lo, hi, sacc.lo64(), sacc.hi64());
	add128(lo, hi, corr_lo, corr_hi);
	block_reduce_agg(lo, hi, n_values, red, a.agg);
}

}  // namespace

cudaError_t launch_scan_sum_fields(const ScanSumArgs &args, int sm_count, cudaStream_t stream) {
	static bool configured[64] = {};
	const size_t smem = (size_t)kStages * kStageBytes;
	int dev = 0;
	cudaGetDevice(&dev);
	if (!configured[dev & 63]) {
		cudaError_t e = cudaFuncSetAttribute(scan_sum_fields_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		configured[dev & 63] = true;
	}
	if (args.seg_b >= args.seg_e) {
		return cudaSuccess;
	}
	// persistent CTAs; an odd grid makes every CTA's strided job sequence cycle through all 16 chunk indices
	const uint32_t grid = ((uint32_t)sm_count * 3u - 1u) | 1u;  // odd and <= the resident capacity
	scan_sum_fields_kernel<<<grid, kThreads, smem, stream>>>(args);
	return cudaGetLastError();
}

}  // namespace adacom
