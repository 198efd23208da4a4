// kernels_pack.cu -- the re-pack loop of the succinct codec, width-specialised for sm_100a.
//
// Replaces the scalar in-place loop of ColumnSegment::BitCompressFromSuccinct / BitCompressFromUncompressed
// (src/storage/table/column_segment.cpp:365-378, 427-446): for i < count
//     field = (slot[i] - min_factor) & lo_set[w];  bits::write_int_and_move(field, w)
// (third_party/sdsl/include/sdsl/bits.hpp:478-499), LSB first, tail bits of the last word zero
// (memory_management.hpp:360-365).
//
// Work unit = one 2048-value tile counted from the segment start (256*w output bytes, 16-byte aligned).
// 64 consecutive values always fill exactly w whole 64-bit words, so a tile is 32 independent "rows" of
// 64 values -> w words.  A producer warp stages the raw tiles of a CTA in shared memory (TMA ring, see
// pack_staged_kernel); the 8 consumer warps take one tile at a time through three phases:
//
//   A  every thread reads 8 raw slots of the staged tile, subtracts the frame of reference, masks to w bits
//      and parks the field in shared memory at an XOR-swizzled position (pos = i ^ (row & 31|15)):
//      conflict-free both for this linear write and for phase B's strided read;
//   B  lane = row, warp q = the q-th eighth of the row's w output words.  The template fixes (w, q), so the
//      list of fields touching those words and every shift amount are compile-time constants: a field costs
//      one shared load and one or two shift/or (gather formulation: no atomics, no inter-thread carries,
//      boundary fields are simply read by both neighbours).  Words go to a staging row of odd stride
//      (w | 1 words): conflict-free 64-bit shared stores;
//   C  the tile's packed words leave as coalesced 64-bit global stores (+ the zeroed gap up to the next
//      128-byte boundary behind the last tile of a segment).
//
// Algorithmic HBM bytes per tile: cnt*S read + 8*ceil(cnt*w/64) written; nothing is re-read.
#include "adacom_kernels.cuh"
#include "swar8.hpp"

namespace adacom {

namespace {

constexpr int kPackThreads = 256;
constexpr int kRowValues = 64;                        // values per row: always w whole output words
constexpr int kRows = kVectorSize / kRowValues;       // 32 rows per tile == lanes of a warp
constexpr int kSubs = kPackThreads / 32;              // 8 warps share the words of a row

struct PackTile {
	uint64_t sub;        // value subtracted from every slot (0 when min_factor is unset)
	uint64_t *out;       // first output word of this tile
	const uint8_t *src;  // first raw value of this tile
	uint32_t cnt;        // values in the tile (0: segment not repacked by this encode call)
	uint32_t w;
	uint32_t gap;        // last tile of a segment: zero words up to the next 128-byte boundary
};

template <int S>
__device__ __forceinline__ PackTile resolve_pack_tile(const DescQ &dq, uint32_t k, const uint8_t *raw, uint32_t extent,
                                                      uint64_t *base) {
	PackTile t;
	const uint64_t row_start = dq.row_start();
	// segments laid out by this encode call are not bound yet (packed_ptr == 0): their payload goes to
	// base + word_offset; bind_extent_kernel publishes the pointer behind this kernel, off the critical path
	const uint64_t packed_ptr = (uint64_t)(base + dq.word_offset());
	const uint64_t mn = dq.min_factor();
	const uint32_t count = dq.count();
	const uint32_t w = dq.width();
	const uint32_t state = dq.state();
	const uint32_t first = k * kVectorSize;
	const bool repacked = (state & kSegRepacked) && dq.extent() == extent && dq.packed_ptr() == 0 && first < count;
	t.cnt = repacked ? min(kVectorSize, count - first) : 0u;
	t.src = raw + (row_start + first) * (uint64_t)S;
	t.sub = mn == ~0ull ? 0ull : mn;
	t.out = (uint64_t *)packed_ptr + (uint64_t)k * (kVectorSize / 64) * w;
	t.w = w;
	uint32_t gap = 0;
	if (t.cnt && first + t.cnt == count) {
		const uint64_t total = ((uint64_t)count * w + 63) >> 6;
		gap = (uint32_t)(((total + kSegAlignWords - 1) & ~(uint64_t)(kSegAlignWords - 1)) - total);
	}
	t.gap = gap;
	return t;
}

// ---- phase A helpers ------------------------------------------------------------------------------------

template <typename F> struct Swz { static constexpr uint32_t kMask = sizeof(F) == 4 ? 31u : 15u; };

// ---- phase B: fields of one row -> the words of sub-range q -------------------------------------------

template <int W, int Q, typename F>
__device__ __forceinline__ void pack_sub(const F *fld_row, uint32_t x, uint64_t *out_row) {
	constexpr int K0 = Q * W / kSubs, K1 = (Q + 1) * W / kSubs;  // output words [K0, K1) of the row
	if constexpr (K1 > K0) {
		constexpr int V0 = 64 * K0 / W, V1 = (64 * K1 - 1) / W;  // fields touching them (inclusive)
		constexpr int NU = 2 * (K1 - K0);
		constexpr uint32_t XM = Swz<F>::kMask;
		uint32_t unit[NU];  // 32-bit output units, unit n covers bits [64*K0 + 32n, +32) of the row
#pragma unroll
		for (int n = 0; n < NU; n++) {
			unit[n] = 0u;
		}
#pragma unroll
		for (int v = V0; v <= V1; v++) {
			const F f = fld_row[(v & ~XM) | ((v & XM) ^ x)];
			const int p = v * W - 64 * K0;      // bit position relative to the first owned word (may be < 0)
			const int n0 = (p + 64) / 32 - 2;   // floor(p / 32)
			const int s = (p + 64) & 31;
			const uint32_t lo = (uint32_t)f;
			const uint32_t hi = sizeof(F) == 8 ? (uint32_t)((uint64_t)f >> 32) : 0u;
			// the field as a 96-bit string shifted left by s, cut into three units
			if (n0 >= 0 && n0 < NU) {
				unit[n0] |= lo << s;
			}
			if (s + W > 32 && n0 + 1 >= 0 && n0 + 1 < NU) {
				unit[n0 + 1] |= s == 0 ? hi : __funnelshift_l(lo, hi, s);
			}
			if (s + W > 64 && n0 + 2 >= 0 && n0 + 2 < NU) {
				unit[n0 + 2] |= hi >> (32 - s);
			}
		}
#pragma unroll
		for (int k = 0; k < K1 - K0; k++) {
			out_row[K0 + k] = ((uint64_t)unit[2 * k + 1] << 32) | unit[2 * k];
		}
	}
}

template <int W, typename F>
__device__ __forceinline__ void phase_b(const F *fld, uint64_t *ostage, uint32_t cnt, uint32_t warp, uint32_t lane) {
	constexpr uint32_t WP = W | 1;  // odd row stride: conflict-free 64-bit stores across the lanes
	if (lane * kRowValues >= cnt) {
		return;
	}
	const F *fld_row = fld + lane * kRowValues;
	const uint32_t x = lane & Swz<F>::kMask;
	uint64_t *out_row = ostage + lane * WP;
	switch (warp) {
	case 0: pack_sub<W, 0, F>(fld_row, x, out_row); break;
	case 1: pack_sub<W, 1, F>(fld_row, x, out_row); break;
	case 2: pack_sub<W, 2, F>(fld_row, x, out_row); break;
	case 3: pack_sub<W, 3, F>(fld_row, x, out_row); break;
	case 4: pack_sub<W, 4, F>(fld_row, x, out_row); break;
	case 5: pack_sub<W, 5, F>(fld_row, x, out_row); break;
	case 6: pack_sub<W, 6, F>(fld_row, x, out_row); break;
	default: pack_sub<W, 7, F>(fld_row, x, out_row); break;
	}
}

template <int W>
__device__ __forceinline__ void phase_b_w(const void *fld, uint64_t *ostage, uint32_t cnt, uint32_t warp, uint32_t lane) {
	if constexpr (W <= 32) {
		phase_b<W, uint32_t>((const uint32_t *)fld, ostage, cnt, warp, lane);
	} else {
		phase_b<W, uint64_t>((const uint64_t *)fld, ostage, cnt, warp, lane);
	}
}

// widths a type of S bytes can be re-packed to: 1 .. 8*S - 1 (re-pack iff 8*S > w, column_segment.cpp:363)
template <int S, int W>
__device__ __forceinline__ void phase_b_case(const void *fld, uint64_t *ostage, uint32_t cnt, uint32_t warp, uint32_t lane) {
	if constexpr (W >= 1 && W < 8 * S) {
		phase_b_w<W>(fld, ostage, cnt, warp, lane);
	}
}

#define PACK_CASE(W) case W: phase_b_case<S, W>(fld, ostage, t.cnt, warp, lane); break;
#define PACK_CASES_8(B) PACK_CASE(B) PACK_CASE(B + 1) PACK_CASE(B + 2) PACK_CASE(B + 3) PACK_CASE(B + 4) PACK_CASE(B + 5) PACK_CASE(B + 6) PACK_CASE(B + 7)

// ---- the kernel: raw tiles staged by the TMA engine, 8 consumer warps pack -------------------------------
//
// With register loads issued at the top of every tile (the first version of this kernel) a CTA had its 2-16 KiB of
// input in flight for about a third of the time and every thread re-derived the tile from the descriptor: ~650
// Gvalues/s for every type, 0.80 of the HBM peak for INT64.  Here a producer warp walks the tile table, resolves the
// descriptors (32 tiles per batch, the next batch's table entries and descriptors already in flight) and keeps a
// ring of raw tiles in flight with cp.async.bulk; the consumer warps never wait on global memory and never touch a
// descriptor: phase A reads the slots from shared memory (measured: INT64 encode 5.2 -> 5.9 TB/s at 17 bits,
// UINT8 0.79 -> 0.55 ms per 2^29 values).
constexpr int kStagedThreads = kPackThreads + 32;

struct PackStage {
	PackTile t;
	uint32_t mis;      // byte offset of the tile's first slot inside the staged (16-byte aligned) bytes
	uint32_t done;     // sentinel
	uint32_t n_words;  // output words of the tile
	uint32_t magic;    // ceil(2^20 / w): m / w for m < 2048 as a multiply-shift (phase C's row of a word)
};

// staged phase A: slots (already in shared memory) -> swizzled fields
template <typename T, typename F>
__device__ __forceinline__ void park_fields(const T *vals, F *fld, const PackTile &t, uint32_t tid) {
	const F mask = (F)lo_set64(t.w);
	const F sub = (F)t.sub;
	constexpr uint32_t XM = Swz<F>::kMask;
	const uint32_t sw = tid ^ (tid >> 6);
	constexpr int PER = kVectorSize / kPackThreads;
	if (t.cnt == kVectorSize) {
#pragma unroll
		for (int j = 0; j < PER; j++) {
			fld[j * kPackThreads + (sw ^ ((4 * j) & XM))] = ((F)vals[j * kPackThreads] - sub) & mask;
		}
	} else {
#pragma unroll
		for (int j = 0; j < PER; j++) {
			const F v = j * kPackThreads + tid < t.cnt ? (F)vals[j * kPackThreads] : sub;  // padding -> zero fields
			fld[j * kPackThreads + (sw ^ ((4 * j) & XM))] = (v - sub) & mask;
		}
	}
}

template <int S> struct StagedSmem {
	static constexpr int kRing = S == 8 ? 2 : 3;      // raw tiles in flight per CTA
	static constexpr int kMinCtas = S == 8 ? 3 : 4;   // resident CTAs per SM the kernel is compiled for
	static constexpr uint32_t kFieldBytes = kVectorSize * (S == 8 ? 8u : 4u);
	static constexpr uint32_t kOstageBytes = kRows * ((8 * S - 1) | 1u) * 8u;
	static constexpr uint32_t kRawBytes = kVectorSize * S + 32u;  // + misalignment + rounding to 16
	static constexpr uint32_t kBytes = kFieldBytes + kOstageBytes + kRing * kRawBytes;
};

template <int S>
__global__ void __launch_bounds__(kStagedThreads, StagedSmem<S>::kMinCtas) pack_staged_kernel(const SegDesc *desc, const Job *tiles,
                                                                        uint32_t n_tiles, const uint8_t *raw,
                                                                        uint32_t extent, uint64_t *base) {
	constexpr int kRing = StagedSmem<S>::kRing;
	extern __shared__ __align__(128) uint8_t pack_smem[];
	__shared__ __align__(8) uint64_t full_bar[kRing];
	__shared__ __align__(8) uint64_t empty_bar[kRing];
	__shared__ PackStage infos[kRing];
	void *fld = pack_smem;
	uint64_t *ostage = (uint64_t *)(pack_smem + StagedSmem<S>::kFieldBytes);
	uint8_t *ring = pack_smem + StagedSmem<S>::kFieldBytes + StagedSmem<S>::kOstageBytes;
	const uint32_t tid = threadIdx.x;
	const uint32_t lane = tid & 31u;
	const uint32_t warp = tid >> 5;
	using T = typename UIntOf<S>::type;

	if (tid == 0) {
		for (int s = 0; s < kRing; s++) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&empty_bar[s], 1);
		}
		mbar_fence_init();
	}
	__syncthreads();

	if (warp == kSubs) {
		// ---------------- producer warp: 32 tiles resolved per batch, published in table order ----------------
		uint32_t seq = 0;
		const uint32_t batch = 32u * gridDim.x;
		Job job_next;
		DescQ dq_next;
		{
			const uint32_t ti = blockIdx.x + lane * gridDim.x;
			if (ti < n_tiles) {
				job_next = tiles[ti];
				dq_next = load_desc(desc + job_next.seg);
			}
		}
		for (uint32_t t0 = blockIdx.x; t0 < n_tiles; t0 += batch) {
			const uint32_t ti = t0 + lane * gridDim.x;
			const Job job = job_next;
			const DescQ dq = dq_next;
			if (ti + batch < n_tiles) {  // table entry and descriptor of the next batch: in flight meanwhile
				job_next = tiles[ti + batch];
				dq_next = load_desc(desc + job_next.seg);
			}
			PackStage st;
			st.t.cnt = 0;
			uint64_t src16 = 0;
			uint32_t bytes = 0;
			if (ti < n_tiles) {
				st.t = resolve_pack_tile<S>(dq, job.k, raw, extent, base);
				src16 = (uint64_t)st.t.src & ~15ull;
				st.mis = (uint32_t)((uint64_t)st.t.src - src16);
				bytes = (st.mis + st.t.cnt * S + 15u) & ~15u;  // the raw arena is padded: the tail over-read stays inside
				st.done = 0;
				// once per tile here instead of once per consumer thread (the division was 7 % of the kernel's instructions)
				st.n_words = (st.t.cnt * st.t.w + 63u) >> 6;
				st.magic = st.t.w ? ((1u << 20) + st.t.w - 1u) / st.t.w : 0u;
			}
			uint32_t todo = __ballot_sync(0xffffffffu, st.t.cnt != 0);
			while (todo) {
				const uint32_t l = (uint32_t)__ffs((int)todo) - 1u;
				todo &= todo - 1u;
				if (lane == l) {
					const uint32_t s = seq % kRing;
					mbar_wait(&empty_bar[s], ((seq / kRing) & 1u) ^ 1u);
					infos[s] = st;
					mbar_arrive_expect_tx(&full_bar[s], bytes);
					tma_load_1d(ring + s * StagedSmem<S>::kRawBytes, (const void *)src16, bytes, &full_bar[s]);
				}
				seq++;
				__syncwarp();
			}
		}
		if (lane == 0) {
			const uint32_t s = seq % kRing;
			mbar_wait(&empty_bar[s], ((seq / kRing) & 1u) ^ 1u);
			infos[s].done = 1;
			mbar_arrive(&full_bar[s]);
		}
		return;
	}

	// ---------------- consumer warps ----------------
	for (uint32_t it = 0;; it++) {
		const uint32_t s = it % kRing;
		mbar_wait(&full_bar[s], (it / kRing) & 1u);
		if (infos[s].done) {
			break;
		}
		const PackTile t = infos[s].t;
		const uint32_t n_words = infos[s].n_words, magic = infos[s].magic;
		if constexpr (S == 1) {
			// One-byte slots, w <= 7: 8 consecutive values are exactly w output bytes, so a thread packs its own 8
			// slots with word-parallel arithmetic and the phases A / B (a shared-memory round trip per field and a
			// barrier) fall away.  Four slots per register: (x | 0x80..) - (min & 0x7f..) cannot borrow across bytes
			// and its low 7 bits per byte are the field's (w <= 7 masks the rest); two shift-and-merge steps squeeze the
			// four w-bit fields together.  The w bytes go to a double-buffered byte image of the tile (the field area).
			const uint32_t a = infos[s].mis + 8u * tid;
			const uint32_t *w32 = (const uint32_t *)(ring + s * StagedSmem<S>::kRawBytes) + (a >> 2);
			const uint32_t sh = (a & 3u) * 8u;
			const uint32_t r0 = w32[0], r1 = w32[1], r2 = w32[2];
			uint32_t x0 = __funnelshift_r(r0, r1, sh), x1 = __funnelshift_r(r1, r2, sh);
			const uint32_t w = t.w;
			const uint32_t sub7 = ((uint32_t)t.sub & 0x7fu) * 0x01010101u, fmask = ((1u << w) - 1u) * 0x01010101u;
			x0 = swar8_fields(x0, sub7, fmask);
			x1 = swar8_fields(x1, sub7, fmask);
			if (t.cnt != kVectorSize) {  // slots behind the segment's last value: zero fields
				const int v = (int)t.cnt - (int)(8u * tid);
				x0 &= v >= 4 ? 0xffffffffu : (v <= 0 ? 0u : (1u << (8 * v)) - 1u);
				x1 &= v >= 8 ? 0xffffffffu : (v <= 4 ? 0u : (1u << (8 * (v - 4))) - 1u);
			}
			const uint64_t c = (uint64_t)swar8_squeeze(x0, w) | ((uint64_t)swar8_squeeze(x1, w) << (4u * w));
			uint8_t *image = (uint8_t *)fld + (it & 1u) * (kVectorSize * 7u / 8u);
			uint8_t *o = image + tid * w;
#pragma unroll
			for (uint32_t k = 0; k < 7u; k++) {
				if (k < w) {
					o[k] = (uint8_t)(c >> (8u * k));
				}
			}
			asm volatile("bar.sync 1, %0;" ::"n"(kPackThreads) : "memory");
			if (tid == 0) {
				mbar_arrive(&empty_bar[s]);
			}
			if (tid < n_words) {
				t.out[tid] = ((const uint64_t *)image)[tid];
			}
			if (tid < t.gap) {
				t.out[n_words + tid] = 0ull;
			}
			continue;
		}
		{
			// phase A from shared memory: slot -> field, parked at its swizzled position.  Value i = j*256 + tid sits
			// in row j*4 + tid/64, so its position i ^ (row & 31) is j*256 + ((tid ^ tid/64) ^ 4j): one XOR with a
			// constant per value, everything else folds into immediate offsets.
			const T *vals = (const T *)(ring + s * StagedSmem<S>::kRawBytes + infos[s].mis) + tid;
			if (S < 8 || t.w <= 32u) {
				park_fields<T, uint32_t>(vals, (uint32_t *)fld, t, tid);
			} else {
				park_fields<T, uint64_t>(vals, (uint64_t *)fld, t, tid);
			}
		}
		asm volatile("bar.sync 1, %0;" ::"n"(kPackThreads) : "memory");
		if (tid == 0) {
			mbar_arrive(&empty_bar[s]);  // raw stage s (and its info) may be refilled
		}
		switch (t.w) {
			PACK_CASES_8(0)
			PACK_CASES_8(8)
			PACK_CASES_8(16)
			PACK_CASES_8(24)
			PACK_CASES_8(32)
			PACK_CASES_8(40)
			PACK_CASES_8(48)
			PACK_CASES_8(56)
		default:
			break;
		}
		asm volatile("bar.sync 1, %0;" ::"n"(kPackThreads) : "memory");
		const uint32_t pad = (t.w & 1u) ^ 1u;
		for (uint32_t m = tid; m < n_words; m += kPackThreads) {
			const uint32_t row = (m * magic) >> 20;
			t.out[m] = ostage[m + row * pad];
		}
		if (tid < t.gap) {
			t.out[n_words + tid] = 0ull;
		}
	}
}

template <int S>
cudaError_t launch_pack_staged(const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw,
                               uint32_t extent, uint64_t *base, int sm_count, cudaStream_t stream) {
	static int per_sm[64] = {};
	const size_t smem = StagedSmem<S>::kBytes;
	auto kernel = pack_staged_kernel<S>;
	int dev = 0;
	cudaGetDevice(&dev);
	if (per_sm[dev & 63] == 0) {
		cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		int n = 0;
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kStagedThreads, smem);
		if (e != cudaSuccess) {
			return e;
		}
		per_sm[dev & 63] = n > 0 ? n : 1;
	}
	if (n_tiles == 0) {
		return cudaSuccess;
	}
	uint32_t grid = (uint32_t)sm_count * (uint32_t)per_sm[dev & 63];
	if (grid > n_tiles) {
		grid = n_tiles;
	}
	kernel<<<grid, kStagedThreads, smem, stream>>>(desc, tiles, n_tiles, raw, extent, base);
	return cudaGetLastError();
}

}  // namespace

cudaError_t launch_encode(TypeInfo t, const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw,
                          uint32_t extent, uint64_t *base, int sm_count, cudaStream_t stream) {
	switch (t.size) {
	case 1:
		return launch_pack_staged<1>(desc, tiles, n_tiles, raw, extent, base, sm_count, stream);
	case 2:
		return launch_pack_staged<2>(desc, tiles, n_tiles, raw, extent, base, sm_count, stream);
	case 4:
		return launch_pack_staged<4>(desc, tiles, n_tiles, raw, extent, base, sm_count, stream);
	case 8:
		return launch_pack_staged<8>(desc, tiles, n_tiles, raw, extent, base, sm_count, stream);
	default:
		return cudaErrorInvalidValue;
	}
}

}  // namespace adacom
