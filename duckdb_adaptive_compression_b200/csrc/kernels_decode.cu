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
#include "swar8.hpp"

namespace adacom {

namespace {

constexpr int kConsumerWarps = 8;
constexpr int kConsumers = kConsumerWarps * 32;  // threads that unpack
constexpr int kThreads = kConsumers + 32;        // + one producer warp that only feeds the TMA ring
constexpr int kEmptyArrivals = 1;  // the consumers meet on a named barrier, one of them frees the stage
constexpr int kMaxStages = 8;
// Stage payload: a fused-scan chunk is <= 16 KiB for every type; a flat tile is 2048 values = at most 2048*S bytes
// (slot width), so the narrower types stage less and fit more CTAs per SM.
// + 16 B misalignment + 16 B pad (read_bits over-read)
//
// Consumer groups: the 8 consumer warps split into kGroups groups that take alternate stages, so that several tiles
// are unpacked per CTA at a time.  That hides the per-tile latency chain (barrier wait -> stage info -> extract ->
// store -> named barrier) which bounds the narrow types, whose 2048-value tiles carry few bytes (measured on B200,
// flat decode: UTINYINT 0.365 -> 0.258 ms per 2^29 values with 4 groups, UINT32 +2-3 % with 2, INT64 neutral).
// kStages must be a multiple of kGroups: a stage is then always drained by the same group, and a group's wait on
// full_bar[s] can never run more than one phase ahead of the barrier (parity aliasing).
template <int S, int MODE> struct StageCfg {
	static constexpr uint32_t kBytes = (MODE == kModeFlat ? kVectorSize * (uint32_t)S : 16384u) + 32u;
	static constexpr int kGroups = MODE != kModeFlat ? 1 : (S == 8 ? 1 : (S == 4 ? 2 : 4));
	// INT64 flat decode: decoded tiles leave through the TMA engine (two 16 KiB output buffers per CTA, 4 input
	// stages).  Measured on B200 against 8-byte stores from the consumer lanes with 6 input stages: -3 % time at 1-3
	// bits, +2-4 % at 8-13 bits, -2.5..-6 % from 17 bits up; -2.4 % over the sweep.
	static constexpr bool kTmaStore = MODE == kModeFlat && S == 8;
	static constexpr uint32_t kOutBytes = kTmaStore ? kVectorSize * (uint32_t)S + 16u : 0u;
	static constexpr int kStages = kTmaStore ? 4 : (kGroups == 4 ? 8 : 6);
	static constexpr uint32_t kSmem = (uint32_t)kStages * kBytes + 2u * kOutBytes;
#ifndef ADACOM_DECODE_MINB
#define ADACOM_DECODE_MINB 4
#endif
	// resident CTAs per SM the kernel is compiled for (register cap) and launched with
	static constexpr int kMinCtas = MODE != kModeFlat ? 2 : (S == 8 ? 2 : (S == 4 ? 4 : ADACOM_DECODE_MINB));
	static_assert(kStages % kGroups == 0 && kStages <= kMaxStages, "a stage belongs to one consumer group");
};

struct StageInfo {
	uint64_t min_add;      // frame of reference to add (0 when none)
	uint64_t row;          // column row index of the unit's first value
	uint32_t bit_base;     // bit position of the first field inside the staged bytes (0..127)
	uint32_t cnt;          // values in this unit
	uint32_t w;
	uint32_t seg;
	uint32_t first_value;  // index of the unit's first value inside its segment
	uint32_t unit;         // fused scans: index of the scanned segment inside the call (== query for kModePoint)
	uint32_t done;         // sentinel: the producer has no more work for this CTA
	uint32_t pad;
};

struct Resolved {
	StageInfo info;
	uint64_t src16;
	uint32_t bytes;
};

// (segment, k) -> segment descriptor -> what to stage.  `gran` = values per work unit: 2048-value tiles for the
// flat decode, scan_chunk_values(w) (gran == 0) for the fused scans.  bytes == 0: nothing to stage.
template <int S>
__device__ __forceinline__ Resolved resolve_unit(const DecodeArgs &a, uint32_t seg, uint32_t k, uint32_t gran,
                                                 const DescQ &dq) {
	Resolved r;
	const uint64_t row_start = dq.row_start();
	const uint64_t packed_ptr = dq.packed_ptr();
	const uint64_t min_factor = dq.min_factor();
	const uint32_t count = dq.count();
	const uint32_t w = dq.width();
	const uint32_t state = dq.state();
	const bool repacked = state & kSegRepacked;
	if (gran == 0) {
		gran = scan_chunk_values(w);
	}
	const uint32_t first = k * gran;
	uint32_t cnt = first < count ? min(gran, count - first) : 0u;
	if (a.uncompact) {
		// Uncompact() is a no-op for segments that are not compacted (column_segment.cpp:327-330); a compacted
		// segment that kept its slot width already sits in the raw arena (unless literal mode adds min, Q1)
		bool touch = (state & kSegCompacted) && (repacked || (a.literal && min_factor != ~0ull));
		if (!touch) {
			cnt = 0;
		}
	}
	const uint64_t src = repacked ? packed_ptr + (((uint64_t)first * w) >> 3)
	                              : (uint64_t)a.raw + (row_start + first) * (uint64_t)S;
	r.src16 = src & ~15ull;
	const uint32_t byte_base = (uint32_t)(src - r.src16);
	r.bytes = cnt ? ((byte_base + ((cnt * w + 7u) >> 3) + 15u) & ~15u) : 0u;
	const bool add = (min_factor != ~0ull) && (repacked || a.literal);
	r.info.min_add = add ? min_factor : 0ull;
	r.info.row = row_start + first;
	r.info.bit_base = byte_base * 8u;
	r.info.cnt = cnt;
	r.info.w = w;
	r.info.seg = seg;
	r.info.first_value = first;
	r.info.unit = 0;
	r.info.done = 0;
	r.info.pad = 0;
	return r;
}

template <int S, bool SIGNED, int MODE>
__global__ void __launch_bounds__(kThreads, StageCfg<S, MODE>::kMinCtas) decode_kernel(DecodeArgs a) {
	extern __shared__ __align__(128) uint8_t stage_mem[];
	constexpr int kStages = StageCfg<S, MODE>::kStages;
	constexpr int kGroups = StageCfg<S, MODE>::kGroups;
	constexpr int kGroupThreads = kConsumers / kGroups;
	__shared__ __align__(8) uint64_t full_bar[kStages];
	__shared__ __align__(8) uint64_t empty_bar[kStages];
	__shared__ StageInfo infos[kStages];
	__shared__ uint64_t red[96];

	using T = typename UIntOf<S>::type;
	constexpr uint32_t kStageBytes = StageCfg<S, MODE>::kBytes;
	const uint32_t tid = threadIdx.x;
	const uint32_t lane = tid & 31u;
	const uint32_t warp = tid >> 5;
	// Flat decode: jobs come from the column's static (segment, tile) table.  Fused scans: no table -- job j =
	// (unit j / kScanParts, chunk index j % kScanParts), unit -> segment seg_b + unit or unit_seg[unit]; chunks
	// that do not exist cost one descriptor read by one producer lane, and the launch uses an odd grid so that a
	// CTA's strided job sequence cycles through every chunk index (even load).
	const uint32_t parts = MODE == kModeFlat ? 1u : min(max(__ldg(a.parts), 1u), kScanParts);
	const uint32_t n_jobs = MODE == kModeFlat ? a.n_jobs : a.n_units * parts;

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
		// ---------------- producer warp: every lane resolves one job; existing units are issued in job order ----
		uint32_t seq = 0;  // stage sequence number, uniform across the warp
		const uint32_t batch = 32u * gridDim.x;
		// job -> (segment, k) -> descriptor, fetched ahead: the pair two batches ahead (flat: a table read), the
		// descriptor one batch ahead, so neither latency sits between two issue loops
		auto job_pair = [&](uint32_t job) -> Job {
			if (MODE == kModeFlat) {
				return a.jobs[job];
			}
			const uint32_t unit = job / parts;
			Job j;
			j.seg = a.unit_seg != nullptr ? __ldg(a.unit_seg + unit) : a.seg_b + unit;
			j.k = job - unit * parts;
			return j;
		};
		const uint32_t job0 = blockIdx.x + lane * gridDim.x;
		Job jb = {0u, 0u}, jb_next = {0u, 0u};
		DescQ dq_next;
		if (job0 < n_jobs) {
			jb = job_pair(job0);
			dq_next = load_desc(a.desc + jb.seg);
		}
		if (job0 + batch < n_jobs) {
			jb_next = job_pair(job0 + batch);
		}
		for (uint32_t j0 = blockIdx.x; j0 < n_jobs; j0 += batch) {
			const uint32_t job = j0 + lane * gridDim.x;
			const DescQ dq = dq_next;
			const Job cur = jb;
			jb = jb_next;
			if (job + batch < n_jobs) {
				dq_next = load_desc(a.desc + jb.seg);
				if (job + 2u * batch < n_jobs) {
					jb_next = job_pair(job + 2u * batch);
				}
			}
			Resolved r;
			r.bytes = 0;
			if (job < n_jobs) {
				r = resolve_unit<S>(a, cur.seg, cur.k, MODE == kModeFlat ? kVectorSize : 0u, dq);
				r.info.unit = MODE == kModeFlat ? 0u : job / parts;
			}
			// one lane at a time, in job order: the slot's previous user (seq - kStages) has then already been
			// issued, so the parity wait is never more than one phase ahead
			uint32_t todo = __ballot_sync(0xffffffffu, r.bytes != 0);
			while (todo) {
				const uint32_t l = (uint32_t)__ffs((int)todo) - 1u;
				todo &= todo - 1u;
				if (lane == l) {
					const uint32_t s = seq % kStages;
					mbar_wait(&empty_bar[s], ((seq / kStages) & 1u) ^ 1u);  // slot drained by the consumers
					infos[s] = r.info;
					mbar_arrive_expect_tx(&full_bar[s], r.bytes);
					tma_load_1d(stage_mem + s * kStageBytes, (const void *)r.src16, r.bytes, &full_bar[s]);
				}
				seq++;
				__syncwarp();
			}
		}
		if (lane == 0) {
			for (int g = 0; g < kGroups; g++, seq++) {  // one sentinel per consumer group
				const uint32_t s = seq % kStages;
				mbar_wait(&empty_bar[s], ((seq / kStages) & 1u) ^ 1u);
				infos[s].done = 1;
				mbar_arrive(&full_bar[s]);
			}
		}
	} else {
		// ---------------- consumer warps ----------------
		// kGroups independent groups take alternate stages (see StageCfg)
		const uint32_t group = warp / (kConsumerWarps / kGroups);
		const uint32_t gtid = tid - group * kGroupThreads;
		for (uint32_t it = group;; it += kGroups) {
			const uint32_t s = it % kStages;
			mbar_wait(&full_bar[s], (it / kStages) & 1u);
			const StageInfo in = infos[s];
			if (in.done) {
				break;
			}
			const uint64_t *words = (const uint64_t *)(stage_mem + s * kStageBytes);
			const uint32_t *words32 = (const uint32_t *)words;
			const uint32_t w = in.w;
			const bool narrow = w <= 32u;
			const uint64_t mask64 = lo_set64(w);
			const uint32_t mask32 = (uint32_t)mask64;

			uint64_t constant = a.constant;
			uint32_t query = 0;
			if (MODE == kModePoint) {
				query = in.unit;
				constant = (uint64_t)a.q_const[query];
			}

			bool point_done = false;
			if (MODE == kModePoint) {
				// Equality look-up in the FIELD domain: value == c  <=>  field == (c - min) mod 2^(8S), and a field
				// is < 2^w -- a chunk whose target does not fit w bits cannot match and is not touched at all.
				// Byte-aligned widths (8 / 16 / 32 bits: the reference's benchmark columns pack to 16) are compared
				// 16 bytes per lane and step with the SIMD-in-a-word compares; a hit (rare) is resolved per field.
				const uint64_t tf = (constant - in.min_add) & lo_set64(8 * S);
				if (w < 64u && (tf >> w) != 0) {
					point_done = true;
				} else if ((w == 8u || w == 16u || w == 32u) && (in.bit_base & 127u) == 0) {
					const uint4 *vec = (const uint4 *)((const uint8_t *)words + (in.bit_base >> 3));
					const uint32_t per = 128u / w;  // fields per 16 bytes
					const uint32_t n_vec = (in.cnt + per - 1u) / per;
					const uint32_t t32 = (uint32_t)tf;
					const uint32_t kk = w == 32u ? t32 : (w == 16u ? t32 * 0x00010001u : t32 * 0x01010101u);
					for (uint32_t j = gtid; j < n_vec; j += kGroupThreads) {
						const uint4 x = vec[j];
						uint32_t m0, m1, m2, m3;
						if (w == 32u) {
							m0 = x.x == kk;
							m1 = x.y == kk;
							m2 = x.z == kk;
							m3 = x.w == kk;
						} else if (w == 16u) {
							m0 = __vcmpeq2(x.x, kk);
							m1 = __vcmpeq2(x.y, kk);
							m2 = __vcmpeq2(x.z, kk);
							m3 = __vcmpeq2(x.w, kk);
						} else {
							m0 = __vcmpeq4(x.x, kk);
							m1 = __vcmpeq4(x.y, kk);
							m2 = __vcmpeq4(x.z, kk);
							m3 = __vcmpeq4(x.w, kk);
						}
						if (m0 | m1 | m2 | m3) {
							const uint32_t xs[4] = {x.x, x.y, x.z, x.w};
							for (uint32_t f = 0; f < per; f++) {
								const uint32_t i = j * per + f;
								const uint32_t fb = f * w;  // bit of the field inside the 16 bytes
								const uint32_t fv = w == 32u ? xs[fb >> 5] : ((xs[fb >> 5] >> (fb & 31u)) & mask32);
								if (i < in.cnt && fv == t32) {
									atomicAdd(a.q_count + query, 1ull);
									atomicMin(a.q_first + query, (unsigned long long)(in.first_value + i));
								}
							}
						}
					}
					point_done = true;
				}
			}
			if (MODE == kModePoint && point_done) {
				// fall through to the barrier / stage release below
			} else
			if (StageCfg<S, MODE>::kTmaStore) {
				// decoded values go to a shared output tile placed so that shared and global addresses agree mod 16;
				// the 16-byte aligned middle leaves as ONE bulk store, the (at most one each) head / tail elements
				// as ordinary stores
				constexpr uint32_t EPB = 16u / S;  // elements per 16 bytes
				uint8_t *obuf = stage_mem + kStages * kStageBytes + (it & 1u) * StageCfg<S, MODE>::kOutBytes;
				T *gdst = (T *)a.out + (in.row - a.out_row0);
				const uint32_t sh = (uint32_t)(((uint64_t)gdst & 15ull) / S);
				const uint32_t head = sh ? min(EPB - sh, in.cnt) : 0u;
				const uint32_t bulk = ((in.cnt - head) / EPB) * EPB;  // elements in the bulk store
				const uint32_t bulk_end = head + bulk;
				T *ob = (T *)obuf + sh;
				if (in.cnt == kVectorSize) {
					constexpr int PER = kVectorSize / kGroupThreads;
					uint64_t x[PER];
					if (narrow) {
#pragma unroll
						for (int j = 0; j < PER; j++) {
							const uint32_t bit = in.bit_base + (j * kGroupThreads + gtid) * w;
							x[j] = (uint64_t)(read_bits32(words32, bit) & mask32) + in.min_add;
						}
					} else {
#pragma unroll
						for (int j = 0; j < PER; j++) {
							const uint32_t bit = in.bit_base + (j * kGroupThreads + gtid) * w;
							x[j] = (read_bits64(words, bit) & mask64) + in.min_add;
						}
					}
#pragma unroll
					for (int j = 0; j < PER; j++) {
						ob[j * kGroupThreads + gtid] = (T)x[j];
					}
					if (gtid < head) {
						gdst[gtid] = (T)x[0];
					}
					if ((PER - 1) * kGroupThreads + gtid >= bulk_end) {  // tail elements live in the last slice
						gdst[(PER - 1) * kGroupThreads + gtid] = (T)x[PER - 1];
					}
				} else {
					for (uint32_t i = gtid; i < in.cnt; i += kGroupThreads) {
						const uint32_t bit = in.bit_base + i * w;
						const uint64_t f = narrow ? (uint64_t)(read_bits32(words32, bit) & mask32) : (read_bits64(words, bit) & mask64);
						const T v = (T)(f + in.min_add);
						ob[i] = v;
						if (i < head || i >= bulk_end) {
							gdst[i] = v;
						}
					}
				}
				fence_proxy_async_smem();
				if (gtid == 0) {
					tma_store_wait_read();  // the previous tile's store no longer reads the other buffer
				}
				asm volatile("bar.sync 1, %0;" ::"n"(kGroupThreads) : "memory");
				if (gtid == 0) {
					mbar_arrive(&empty_bar[s]);
					if (bulk) {
						tma_store_1d(gdst + head, ob + head, bulk * S);
					}
				}
				continue;
			}
			if (MODE == kModeFlat && S < 8 && in.cnt == kVectorSize &&
			    ((((uint64_t)a.out + (in.row - a.out_row0) * S) & 7ull) == 0)) {
				// narrow types, full tile: a lane unpacks 8 / S adjacent values and leaves ONE 8-byte store
				// (the kernel is bound by values per second here, not by bytes)
				constexpr int VS = 8 / S;                          // values per store
				constexpr int GROUPS = kVectorSize / kGroupThreads / VS;  // stores per thread
				uint64_t *dst = (uint64_t *)((T *)a.out + (in.row - a.out_row0)) + gtid;
				if (S == 1) {
					// one-byte slots: the 8 values of a store are exactly w payload bytes (w <= 7 re-packed, w = 8 for a
					// slot-width segment).  They are spread word-parallel -- 4w bits -> two 2w-bit pairs at 16-bit positions
					// -> four fields at byte positions -- and the frame of reference is added to four slots at once: the
					// low 7 bits per byte cannot carry into the next byte, and the top bit is an XOR.  ~30 instructions per
					// 8 values against ~80 for eight separate field reads.
					const bool spread = w < 8u;
					const uint32_t fm = ((1u << w) - 1u) * 0x01010101u, fm_even = fm & 0x00ff00ffu, fm_odd = fm & 0xff00ff00u;
					const uint32_t m4 = spread ? (1u << (4u * w)) - 1u : 0u, m2 = (1u << (2u * w)) - 1u;
					const uint32_t add7 = ((uint32_t)in.min_add & 0x7fu) * 0x01010101u;
					const uint32_t add80 = ((uint32_t)in.min_add & 0x80u) * 0x01010101u;
					const uint32_t byte0 = in.bit_base >> 3;
#pragma unroll
					for (int j = 0; j < GROUPS; j++) {
						const uint32_t at = byte0 + (uint32_t)(j * kGroupThreads + gtid) * w;
						const uint32_t *p32 = words32 + (at >> 2);
						const uint32_t sh = (at & 3u) * 8u;
						const uint32_t r0 = p32[0], r1 = p32[1], r2 = p32[2];
						uint32_t x0 = __funnelshift_r(r0, r1, sh), x1 = __funnelshift_r(r1, r2, sh);
						if (spread) {
							const uint32_t q0 = x0 & m4, q1 = __funnelshift_r(x0, x1, 4u * w) & m4;
							x0 = swar8_spread(q0, w, m2, fm_even, fm_odd);
							x1 = swar8_spread(q1, w, m2, fm_even, fm_odd);
						}
						const uint32_t y0 = swar8_add(x0, add7, add80), y1 = swar8_add(x1, add7, add80);
						dst[j * kGroupThreads] = (uint64_t)y0 | ((uint64_t)y1 << 32);
					}
				} else
#pragma unroll
				for (int j = 0; j < GROUPS; j++) {
					const uint32_t i0 = (j * kGroupThreads + gtid) * VS;
					uint64_t packed = 0;
#pragma unroll
					for (int v = 0; v < VS; v++) {
						const uint32_t bit = in.bit_base + (i0 + v) * w;  // S < 8: w <= 32
						const uint64_t x = (uint64_t)(read_bits32(words32, bit) & mask32) + in.min_add;
						packed |= (x & lo_set64(8 * S)) << (8 * S * v);
					}
					dst[j * kGroupThreads] = packed;
				}
			} else if (MODE == kModeFlat && in.cnt == kVectorSize && (w <= 2u || w >= 28u)) {
				// full tile: the 8 fields of a thread are extracted back to back (independent shared loads in flight
				// together), then stored; adjacent lanes still own adjacent values (coalesced 8-byte stores).
				// Measured A/B on B200: -6 % time at w = 1, -1..3 % above 28 bits, but +5..10 % between (the store
				// bursts of the 3..27-bit tiles collide), hence the width window.
				constexpr int PER = kVectorSize / kGroupThreads;
				T *dst = (T *)a.out + (in.row - a.out_row0) + gtid;
				uint64_t x[PER];
				if (narrow) {
#pragma unroll
					for (int j = 0; j < PER; j++) {
						const uint32_t bit = in.bit_base + (j * kGroupThreads + gtid) * w;
						x[j] = (uint64_t)(read_bits32(words32, bit) & mask32) + in.min_add;
					}
				} else {
#pragma unroll
					for (int j = 0; j < PER; j++) {
						const uint32_t bit = in.bit_base + (j * kGroupThreads + gtid) * w;
						x[j] = (read_bits64(words, bit) & mask64) + in.min_add;
					}
				}
#pragma unroll
				for (int j = 0; j < PER; j++) {
					dst[j * kGroupThreads] = (T)x[j];
				}
			} else
			for (uint32_t base = 0; base < in.cnt && !(MODE == kModePoint && point_done); base += kGroupThreads) {
				const uint32_t i = base + gtid;
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
							uint64_t bit0 = in.row - a.out_row0 + base + (gtid & ~31u);
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
			// (immediate barrier ids: a register id makes ptxas reserve all 16 named barriers)
			if (kGroups == 1 || group == 0) {
				asm volatile("bar.sync 1, %0;" ::"n"(kGroupThreads) : "memory");
			} else if (group == 1) {
				asm volatile("bar.sync 2, %0;" ::"n"(kGroupThreads) : "memory");
			} else if (group == 2) {
				asm volatile("bar.sync 3, %0;" ::"n"(kGroupThreads) : "memory");
			} else {
				asm volatile("bar.sync 4, %0;" ::"n"(kGroupThreads) : "memory");
			}
			if (gtid == 0) {
				mbar_arrive(&empty_bar[s]);  // stage s (payload and info) is free again
			}
		}
		if (StageCfg<S, MODE>::kTmaStore && tid == 0) {
			tma_store_wait_all();
		}
	}

	if (MODE == kModeSum || MODE == kModeFilter) {
		block_reduce_agg(acc.lo64(), acc.hi64(), (uint64_t)n_pass, red, a.agg);
	}
}

template <int S, bool SIGNED, int MODE>
cudaError_t launch_one(const DecodeArgs &args, int sm_count, cudaStream_t stream) {
	static int configured[64] = {};  // per device: the attribute is per (function, device); holds resident CTAs / SM
	const size_t smem = StageCfg<S, MODE>::kSmem;
	auto kernel = decode_kernel<S, SIGNED, MODE>;
	int dev = 0;
	cudaGetDevice(&dev);
	if (!configured[dev & 63]) {
		cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		int resident = 0;
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, kernel, kThreads, smem);
		if (e != cudaSuccess) {
			return e;
		}
		configured[dev & 63] = resident > 0 ? resident : 1;
	}
	const uint32_t n_jobs = MODE == kModeFlat ? args.n_jobs : args.n_units * kScanParts;
	if (n_jobs == 0) {
		return cudaSuccess;
	}
	// INT64 flat: (4 input stages + 2 output tiles) x 16 KiB x 2 CTAs per SM; fused scans: 6 stages x 16 KiB x 2.
	// The narrower types are bound by values per second, not bytes: their flat stages are smaller, so 4 CTAs per
	// SM hide the per-value latency.  All CTAs resident; fused scans: odd grid (see the kernel)
	// Never more CTAs than are resident at once: a persistent grid with a second wave doubles the tail.
	const uint32_t want_per_sm = (uint32_t)StageCfg<S, MODE>::kMinCtas;
	const uint32_t per_sm = want_per_sm < (uint32_t)configured[dev & 63] ? want_per_sm : (uint32_t)configured[dev & 63];
	uint32_t grid = (uint32_t)sm_count * per_sm;
	if (MODE != kModeFlat) {
		grid = (grid - 1u) | 1u;
	}
	if (grid > n_jobs) {
		grid = n_jobs;
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
		return cudaErrorInvalidValue;  // fused SUM lives in kernels_scan.cu (width-specialised)
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
