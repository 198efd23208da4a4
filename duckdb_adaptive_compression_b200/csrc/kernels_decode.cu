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

constexpr int kThreads = 256;
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

template <int S>
__device__ __forceinline__ void issue_unit(const DecodeArgs &a, uint32_t job_idx, uint8_t *buf, StageInfo *info,
                                           uint64_t *bar) {
	Job j = a.jobs[job_idx];
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
	uint64_t src16 = src & ~15ull;
	uint32_t byte_base = (uint32_t)(src - src16);
	uint32_t bytes = (byte_base + ((cnt * w + 7u) >> 3) + 15u) & ~15u;
	bool add = (min_factor != ~0ull) && (repacked || a.literal);
	info->min_add = add ? min_factor : 0ull;
	info->row = row_start + first;
	info->bit_base = byte_base * 8u;
	info->cnt = cnt;
	info->w = w;
	info->seg = j.seg;
	info->first_value = first;
	info->job = job_idx;
	mbar_arrive_expect_tx(bar, cnt ? bytes : 0u);
	if (cnt) {
		tma_load_1d(buf, (const void *)src16, bytes, bar);
	}
}

template <int S, bool SIGNED, int MODE>
__global__ void __launch_bounds__(kThreads) decode_kernel(const DecodeArgs a) {
	extern __shared__ __align__(128) uint8_t stage_mem[];
	__shared__ __align__(8) uint64_t bars[kStages];
	__shared__ StageInfo infos[kStages];
	__shared__ uint64_t red[96];

	using T = typename UIntOf<S>::type;
	const uint32_t tid = threadIdx.x;
	const uint32_t lane = tid & 31u;

	if (tid == 0) {
		for (int s = 0; s < kStages; s++) {
			mbar_init(&bars[s], 1);
		}
		mbar_fence_init();
	}
	__syncthreads();
	if (tid == 0) {
		for (int s = 0; s < kStages; s++) {
			uint32_t job = blockIdx.x + (uint32_t)s * gridDim.x;
			if (job < a.n_jobs) {
				issue_unit<S>(a, job, stage_mem + s * kStageBytes, &infos[s], &bars[s]);
			}
		}
	}

	Acc96 acc;
	uint32_t n_pass = 0;

	for (uint32_t it = 0;; it++) {
		uint32_t job = blockIdx.x + it * gridDim.x;
		if (job >= a.n_jobs) {
			break;
		}
		const uint32_t s = it % kStages;
		mbar_wait(&bars[s], (it / kStages) & 1u);
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

		for (uint32_t base = 0; base < in.cnt; base += kThreads) {
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

		__syncthreads();  // every thread is done with stage s (payload and info)
		if (tid == 0) {
			uint32_t next = job + (uint32_t)kStages * gridDim.x;
			if (next < a.n_jobs) {
				issue_unit<S>(a, next, stage_mem + s * kStageBytes, &infos[s], &bars[s]);
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
	if (args.n_jobs == 0) {
		return cudaSuccess;
	}
	uint32_t grid = (uint32_t)sm_count * 3u;  // 3 CTAs x 64.1 KiB of stages per SM
	if (grid > args.n_jobs) {
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
