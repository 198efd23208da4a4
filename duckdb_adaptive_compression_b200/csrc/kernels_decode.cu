// kernels_decode.cu -- decode / scan side of the succinct codec for sm_100a.
//
// Replaces the reference's SuccinctScanPartial<T> (src/storage/compression/succinct.cpp:123-144: per call a
// deep copy of the packed segment + one scalar bits::read_int per value), the separate filter pass
// (TemplatedFilterSelection, src/storage/table/column_segment.cpp:575-658) and the aggregate that DuckDB
// runs on the materialised vector, with ONE persistent kernel:
//
//   * work unit = (segment, k): a 2048-value tile (flat decode, == one DuckDB vector counted from the segment
//     start) or a <=16 KiB chunk of packed words (fused scans);
//   * the unit's packed words are staged global -> shared by the TMA engine (cp.async.bulk, SASS UBLKCP)
//     into a kStages-deep ring, completion signalled on an mbarrier per stage; one elected thread issues the
//     copies kStages units ahead, so HBM latency is covered by bytes in flight rather than by occupancy;
//   * every lane extracts one field per step with a funnel shift (fields at i*w are adjacent across lanes:
//     conflict-free shared loads), adds the frame of reference and either stores T values coalesced
//     (flat), or folds them into a 96-bit per-thread accumulator (SUM / filter), reduced once per CTA.
//
// Algorithmic HBM bytes per unit: flat = P_tile read + cnt*S written; SUM / filter = P_chunk read.
#include "adacom_kernels.cuh"

namespace adacom {

namespace {

constexpr int kConsumerWarps = 8;
constexpr int kConsumers = kConsumerWarps * 32;  // threads that unpack
constexpr int kThreads = kConsumers + 32;        // + one producer warp that only feeds the TMA ring
constexpr int kEmptyArrivals = 1;  // the consumers meet on a named barrier, one of them frees the stage
constexpr int kStages = 4;
constexpr uint32_t kStagePayload = 16384;
constexpr uint32_t kStageBytes = kStagePayload + 32;  // + 16 B misalignment + 16 B pad (read_bits over-read)

struct StageInfo {
	uint64_t min_add;      // frame of reference to add (0 when none)
	uint64_t row;          // column row index of the unit's first value
	uint32_t bit_base;     // bit position of the first field inside the staged bytes (0..127)
	uint32_t cnt;          // values in this unit
	uint32_t w;
	uint32_t seg;
	uint32_t first_value;  // index of the unit's first value inside its segment
	uint32_t job;
};

struct Resolved {
	StageInfo info;
	uint64_t src16;
	uint32_t bytes;
};

// job -> segment descriptor -> what to stage (two dependent global loads; the producer warp resolves
// kStages units in parallel, one per lane, so this latency is paid once per ring revolution)
template <int S>
__device__ __forceinline__ Resolved resolve_unit(const DecodeArgs &a, uint32_t job_idx) {
	Resolved r;
	Job j = a.jobs[job_idx];
	j.k &= ~kJobFastBit;
	const SegDesc *dp = a.desc + j.seg;
	uint64_t row_start = dp->row_start;
	uint64_t packed_ptr = dp->packed_ptr;
	uint64_t min_factor = dp->min_factor;
	uint32_t count = dp->count;
	uint32_t w = dp->width;
	uint32_t state = dp->state;
	bool repacked = state & kSegRepacked;
	uint32_t gran = a.tile_granularity ? a.tile_granularity : scan_chunk_values(w);
	uint32_t first = j.k * gran;
	uint32_t cnt = min(gran, count - first);
	if (a.uncompact) {
		// Uncompact() is a no-op for segments that are not compacted (column_segment.cpp:327-330); a compacted
		// segment that kept its slot width already sits in the raw arena (unless literal mode adds min, Q1)
		bool touch = (state & kSegCompacted) && (repacked || (a.literal && min_factor != ~0ull));
		if (!touch) {
			cnt = 0;
		}
	}
	uint64_t src = repacked ? packed_ptr + (((uint64_t)first * w) >> 3)
	                        : (uint64_t)a.raw + (row_start + first) * (uint64_t)S;
	r.src16 = src & ~15ull;
	uint32_t byte_base = (uint32_t)(src - r.src16);
	r.bytes = cnt ? ((byte_base + ((cnt * w + 7u) >> 3) + 15u) & ~15u) : 0u;
	bool add = (min_factor != ~0ull) && (repacked || a.literal);
	r.info.min_add = add ? min_factor : 0ull;
	r.info.row = row_start + first;
	r.info.bit_base = byte_base * 8u;
	r.info.cnt = cnt;
	r.info.w = w;
	r.info.seg = j.seg;
	r.info.first_value = first;
	r.info.job = job_idx;
	return r;
}

template <int S, bool SIGNED, int MODE>
__global__ void __launch_bounds__(kThreads) decode_kernel(DecodeArgs a) {
	extern __shared__ __align__(128) uint8_t stage_mem[];
	__shared__ __align__(8) uint64_t full_bar[kStages];
	__shared__ __align__(8) uint64_t empty_bar[kStages];
	__shared__ StageInfo infos[kStages];
	__shared__ uint64_t red[96];

	using T = typename UIntOf<S>::type;
	const uint32_t tid = threadIdx.x;
	const uint32_t lane = tid & 31u;
	const uint32_t warp = tid >> 5;
	if (a.chunk_first != nullptr) {  // job range read from the device-built chunk table
		const uint32_t j0 = a.chunk_first[a.seg_b];
		a.n_jobs = a.chunk_first[a.seg_e] - j0;
		a.jobs += j0;
	}

	if (tid == 0) {
		for (int s = 0; s < kStages; s++) {
			mbar_init(&full_bar[s], 1);                // the producer's arrive.expect_tx (+ the TMA bytes)
			mbar_init(&empty_bar[s], kEmptyArrivals);
		}
		mbar_fence_init();
	}
	__syncthreads();

	Acc96 acc;
	uint32_t n_pass = 0;

	if (warp == kConsumerWarps) {
		// ---------------- producer warp: lane s feeds stage s ----------------
		for (uint32_t it0 = 0;; it0 += kStages) {
			if (blockIdx.x + it0 * gridDim.x >= a.n_jobs) {
				break;
			}
			const uint32_t my_it = it0 + lane;
			const uint32_t job = blockIdx.x + my_it * gridDim.x;
			const bool have = lane < kStages && job < a.n_jobs;
			Resolved r;
			if (have) {
				r = resolve_unit<S>(a, job);
			}
#pragma unroll
			for (int s = 0; s < kStages; s++) {
				if ((int)lane == s && have) {
					mbar_wait(&empty_bar[s], ((my_it / kStages) & 1u) ^ 1u);  // slot drained by all consumer warps
					infos[s] = r.info;
					mbar_arrive_expect_tx(&full_bar[s], r.bytes);
					if (r.bytes) {
						tma_load_1d(stage_mem + s * kStageBytes, (const void *)r.src16, r.bytes, &full_bar[s]);
					}
				}
				__syncwarp();
			}
		}
	} else {
		// ---------------- consumer warps ----------------
		for (uint32_t it = 0;; it++) {
			const uint32_t job = blockIdx.x + it * gridDim.x;
			if (job >= a.n_jobs) {
				break;
			}
			const uint32_t s = it % kStages;
			mbar_wait(&full_bar[s], (it / kStages) & 1u);
			const StageInfo in = infos[s];
			const uint64_t *words = (const uint64_t *)(stage_mem + s * kStageBytes);
			const uint32_t *words32 = (const uint32_t *)words;
			const uint32_t w = in.w;
			const bool narrow = w <= 32u;
			const uint64_t mask64 = lo_set64(w);
			const uint32_t mask32 = (uint32_t)mask64;

			uint64_t constant = a.constant;
			uint32_t query = 0;
			if (MODE == kModePoint) {
				query = a.job_query[job];
				constant = (uint64_t)a.q_const[query];
			}

			for (uint32_t base = 0; base < in.cnt; base += kConsumers) {
				const uint32_t i = base + tid;
				const bool valid = i < in.cnt;
				uint64_t x = 0;
				if (valid) {
					const uint32_t bit = in.bit_base + i * w;
					uint64_t f = narrow ? (uint64_t)(read_bits32(words32, bit) & mask32) : (read_bits64(words, bit) & mask64);
					x = f + in.min_add;
				}
				if (MODE == kModeFlat) {
					if (valid) {
						((T *)a.out)[in.row - a.out_row0 + i] = (T)x;
					}
				} else if (MODE == kModeSum) {
					if (valid) {
						acc.add<SIGNED>(extend_low<S, SIGNED>(x));
						n_pass++;
					}
				} else if (MODE == kModeFilter) {
					uint64_t v = extend_low<S, SIGNED>(x);
					bool pass = valid && cmp_passes<SIGNED>(v, constant, a.op_mask);
					if (pass) {
						acc.add<SIGNED>(v);
						n_pass++;
					}
					if (a.bitmap32 != nullptr) {
						uint32_t ballot = __ballot_sync(0xffffffffu, pass);
						if (lane == 0 && ballot) {
							uint64_t bit0 = in.row - a.out_row0 + base + (tid & ~31u);
							uint32_t sh = (uint32_t)(bit0 & 31u);
							atomicOr(a.bitmap32 + (bit0 >> 5), ballot << sh);
							if (sh) {
								atomicOr(a.bitmap32 + (bit0 >> 5) + 1, ballot >> (32u - sh));
							}
						}
					}
				} else {  // kModePoint: equality look-up, matches are rare
					uint64_t v = extend_low<S, SIGNED>(x);
					if (valid && v == constant) {
						atomicAdd(a.q_count + query, 1ull);
						atomicMin(a.q_first + query, (unsigned long long)(in.first_value + i));
					}
				}
			}
			// Keep the consumer warps of a CTA in lockstep (named barrier 1, the producer is not involved): their
			// stores then hit the same 16 KiB output tile close in time, which measured 5-15 % faster on HBM3e
			// than letting the warps drift apart.
			asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
			if (tid == 0) {
				mbar_arrive(&empty_bar[s]);  // stage s (payload and info) is free again
			}
		}
	}

	if (MODE == kModeSum || MODE == kModeFilter) {
		block_reduce_agg(acc.lo64(), acc.hi64(), (uint64_t)n_pass, red, a.agg);
	}
}

template <int S, bool SIGNED, int MODE>
cudaError_t launch_one(const DecodeArgs &args, int sm_count, cudaStream_t stream) {
	static bool configured[64] = {};  // per device: the attribute is per (function, device)
	const size_t smem = (size_t)kStages * kStageBytes;
	auto kernel = decode_kernel<S, SIGNED, MODE>;
	int dev = 0;
	cudaGetDevice(&dev);
	if (!configured[dev & 63]) {
		cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		configured[dev & 63] = true;
	}
	if (args.n_jobs == 0 && args.chunk_first == nullptr) {
		return cudaSuccess;
	}
	uint32_t grid = (uint32_t)sm_count * 3u;  // 3 CTAs x 64.1 KiB of stages per SM
	if (args.chunk_first == nullptr && grid > args.n_jobs) {
		grid = args.n_jobs;
	}
	kernel<<<grid, kThreads, smem, stream>>>(args);
	return cudaGetLastError();
}

template <int S, bool SIGNED>
cudaError_t launch_mode(DecodeMode mode, const DecodeArgs &args, int sm_count, cudaStream_t stream) {
	switch (mode) {
	case kModeFlat:
		return launch_one<S, false, kModeFlat>(args, sm_count, stream);
	case kModeSum:
		return launch_one<S, SIGNED, kModeSum>(args, sm_count, stream);
	case kModeFilter:
		return launch_one<S, SIGNED, kModeFilter>(args, sm_count, stream);
	default:
		return launch_one<S, SIGNED, kModePoint>(args, sm_count, stream);
	}
}

template <int S>
cudaError_t launch_signed(bool is_signed, DecodeMode mode, const DecodeArgs &args, int sm_count, cudaStream_t stream) {
	return is_signed ? launch_mode<S, true>(mode, args, sm_count, stream)
	                 : launch_mode<S, false>(mode, args, sm_count, stream);
}

}  // namespace

cudaError_t launch_decode(TypeInfo t, DecodeMode mode, const DecodeArgs &args, int sm_count, cudaStream_t stream) {
	switch (t.size) {
	case 1:
		return launch_signed<1>(t.is_signed, mode, args, sm_count, stream);
	case 2:
		return launch_signed<2>(t.is_signed, mode, args, sm_count, stream);
	case 4:
		return launch_signed<4>(t.is_signed, mode, args, sm_count, stream);
	case 8:
		return launch_signed<8>(t.is_signed, mode, args, sm_count, stream);
	default:
		return cudaErrorInvalidValue;
	}
}

}  // namespace adacom
