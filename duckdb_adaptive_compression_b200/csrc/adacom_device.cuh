// adacom_device.cuh -- device-side building blocks shared by the succinct codec kernels (sm_100a).
//
// Bit layout contract (== sdsl::int_vector<0>, reference third_party/sdsl/include/sdsl/bits.hpp:456-529,
// int_vector.hpp:1359-1363): a little-endian array of 64-bit words; value i of width w occupies bits
// [i*w, (i+1)*w) of the concatenated bit string, LSB first, and may straddle two words.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace adacom {

constexpr uint32_t kVectorSize = 2048;   // STANDARD_VECTOR_SIZE (vector_size.hpp:17)
constexpr uint32_t kBlockSize = 262136;  // Storage::BLOCK_SIZE (constants.hpp:101-106)
constexpr uint32_t kSegAlignWords = 16;  // every packed segment starts on a 128-byte boundary

enum : uint8_t { kSegCompacted = 1, kSegRepacked = 2, kSegAnalyzed = 4, kSegUncompressedFn = 8 };
enum : int { kFlagPadded = 1, kFlagLiteral = 2, kFlagZeroExt = 4 };

// Device-resident segment descriptor (64 bytes, one per ColumnSegment).
struct __align__(16) SegDesc {
	uint64_t row_start;    // first value of the segment inside the column / raw arena
	uint64_t packed_ptr;   // device address of the payload words (0 when not repacked)
	uint64_t word_offset;  // payload offset inside its packed extent, in words (multiple of kSegAlignWords)
	uint64_t min_factor;   // UINT64_MAX = unset (column_segment.cpp:96)
	uint64_t max_factor;
	uint32_t count;
	uint32_t capacity;     // slots (segment_size / sizeof(T))
	uint32_t extent;       // id of the packed extent holding the payload
	uint32_t prev_wst;     // width | state << 8 as they were before the last layout_kernel rewrote them (rollback)
	uint8_t width;         // current payload width (8*S when not repacked)
	uint8_t state;         // kSeg* bits
	uint8_t pad[6];
};
static_assert(sizeof(SegDesc) == 64, "SegDesc must stay 64 bytes");
// the kernels read descriptors as four 16-byte quarters
static_assert(offsetof(SegDesc, packed_ptr) == 8 && offsetof(SegDesc, min_factor) == 24 &&
                  offsetof(SegDesc, max_factor) == 32 && offsetof(SegDesc, count) == 40 &&
                  offsetof(SegDesc, extent) == 48 && offsetof(SegDesc, width) == 56 && offsetof(SegDesc, state) == 57,
              "SegDesc field offsets are baked into the quarter loads");

// A descriptor as the kernels read it: four independent 16-byte loads (one memory round trip), decoded later
// so that a producer can fetch the descriptors of its NEXT batch of jobs while it issues the current one.
struct DescQ {
	uint4 q0;  // row_start, packed_ptr
	uint4 q1;  // word_offset, min_factor
	uint4 q2;  // max_factor, count, capacity
	uint4 q3;  // extent, prev_wst, width | state << 8 | pad, pad
#ifdef __CUDACC__
	__device__ __forceinline__ uint64_t row_start() const { return ((uint64_t)q0.y << 32) | q0.x; }
	__device__ __forceinline__ uint64_t packed_ptr() const { return ((uint64_t)q0.w << 32) | q0.z; }
	__device__ __forceinline__ uint64_t word_offset() const { return ((uint64_t)q1.y << 32) | q1.x; }
	__device__ __forceinline__ uint64_t min_factor() const { return ((uint64_t)q1.w << 32) | q1.z; }
	__device__ __forceinline__ uint64_t max_factor() const { return ((uint64_t)q2.y << 32) | q2.x; }
	__device__ __forceinline__ uint32_t count() const { return q2.z; }
	__device__ __forceinline__ uint32_t extent() const { return q3.x; }
	__device__ __forceinline__ uint32_t width() const { return q3.z & 0xffu; }
	__device__ __forceinline__ uint32_t state() const { return (q3.z >> 8) & 0xffu; }
#endif
};

#ifdef __CUDACC__
__device__ __forceinline__ DescQ load_desc(const SegDesc *d) {
	DescQ q;
	const uint4 *p = (const uint4 *)d;
	q.q0 = __ldg(p);
	q.q1 = __ldg(p + 1);
	q.q2 = __ldg(p + 2);
	q.q3 = __ldg(p + 3);
	return q;
}
#endif

// A unit of work for the tile kernels: (segment, index of the tile/chunk inside the segment).
struct Job {
	uint32_t seg;
	uint32_t k;
};

// Values per scan chunk for a payload of width w: the largest multiple of 128 values whose packed bytes
// (16*w*floor(1024/w)) fit one 16 KiB stage and stay 16-byte aligned (cp.async.bulk rule).  From 43 bits up that
// would be 46 .. 32 periods of 64 values -- a poor fit for the period-per-lane scan kernels, whose warps take 32
// periods at a time (measured: odd widths 10-15 % behind the even ones) -- so those widths use exactly 32 periods
// (2048 values, 10.75 .. 16 KiB).  Below 43 bits the 16 KiB chunk wins (measured both ways).
__host__ __device__ inline uint32_t scan_chunk_values(uint32_t w) {
	return w >= 43u ? 2048u : 128u * (1024u / w);
}
// A segment holds <= 262136 payload bytes, a chunk >= 15.4 KiB (or 2048 values of an 8-byte type): never more
// than 16 chunks per segment (checked for every (S, w) in tests/test_capi_boundary.py).
constexpr uint32_t kScanParts = 16;

__host__ __device__ inline uint64_t lo_set64(uint32_t len) {
	return len >= 64 ? ~0ull : ((1ull << len) - 1ull);
}

// sdsl::bits::hi(x) + 1 with hi(0) == 0 (bits.hpp:392-414)
__host__ __device__ inline uint32_t width_for_range(uint64_t range, bool padded) {
#ifdef __CUDA_ARCH__
	uint32_t w = range == 0 ? 1u : (uint32_t)(64 - __clzll((long long)range));
#else
	uint32_t w = range == 0 ? 1u : (uint32_t)(64 - __builtin_clzll(range));
#endif
	if (padded) {
		w = (w + 7u) & ~7u;
	}
	return w;
}

// ---- typed views ---------------------------------------------------------------------------------

template <int S> struct UIntOf;
template <> struct UIntOf<1> { using type = uint8_t; };
template <> struct UIntOf<2> { using type = uint16_t; };
template <> struct UIntOf<4> { using type = uint32_t; };
template <> struct UIntOf<8> { using type = uint64_t; };

// low S bytes of x, then sign- or zero-extended to 64 bits (what DuckDB sees as the T value)
template <int S, bool SIGNED>
__device__ __forceinline__ uint64_t extend_low(uint64_t x) {
	if (S == 8) {
		return x;
	}
	if (SIGNED) {
		if (S == 4) return (uint64_t)(int64_t)(int32_t)(uint32_t)x;
		if (S == 2) return (uint64_t)(int64_t)(int16_t)(uint16_t)x;
		return (uint64_t)(int64_t)(int8_t)(uint8_t)x;
	}
	if (S == 4) return (uint64_t)(uint32_t)x;
	if (S == 2) return (uint64_t)(uint16_t)x;
	return (uint64_t)(uint8_t)x;
}

// ---- bit-field extraction from a staged word array --------------------------------------------------

// 64-bit field at bit position `bit` of a 64-bit word array (may straddle); caller masks.
__device__ __forceinline__ uint64_t read_bits64(const uint64_t *words, uint32_t bit) {
	uint32_t idx = bit >> 6;
	uint32_t off = bit & 63u;
	uint64_t lo = words[idx];
	uint64_t hi = words[idx + 1];  // stage buffers carry one pad word, so this never faults
	uint32_t l0 = (uint32_t)lo, l1 = (uint32_t)(lo >> 32), h0 = (uint32_t)hi, h1 = (uint32_t)(hi >> 32);
	bool upper = off & 32u;
	uint32_t a = upper ? l1 : l0;
	uint32_t b = upper ? h0 : l1;
	uint32_t c = upper ? h1 : h0;
	uint32_t r0 = __funnelshift_r(a, b, off);  // shift amount is taken mod 32
	uint32_t r1 = __funnelshift_r(b, c, off);
	return ((uint64_t)r1 << 32) | r0;
}

// <=32-bit field at bit position `bit` of the same array viewed as 32-bit words; caller masks.
__device__ __forceinline__ uint32_t read_bits32(const uint32_t *words32, uint32_t bit) {
	uint32_t idx = bit >> 5;
	return __funnelshift_r(words32[idx], words32[idx + 1], bit);
}

// ---- mbarrier + 1-D TMA (cp.async.bulk) wrappers -----------------------------------------------------

__device__ __forceinline__ uint32_t smem_addr(const void *p) {
	return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}

__device__ __forceinline__ void mbar_fence_init() {
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes)
	             : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
	uint32_t addr = smem_addr(bar);
	uint32_t done;
	do {
		asm volatile(
		    "{\n"
		    ".reg .pred p;\n"
		    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
		    "selp.u32 %0, 1, 0, p;\n"
		    "}\n"
		    : "=r"(done)
		    : "r"(addr), "r"(parity)
		    : "memory");
	} while (!done);
}

// The same wait for a warp that may have to wait long (a producer far ahead of slow consumers): a spinning warp is
// always eligible and takes issue slots from the warps it is waiting for, so retries back off with nanosleep.
__device__ __forceinline__ void mbar_wait_backoff(uint64_t *bar, uint32_t parity, uint32_t ns) {
	uint32_t addr = smem_addr(bar);
	uint32_t done;
	for (;;) {
		asm volatile(
		    "{\n"
		    ".reg .pred p;\n"
		    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
		    "selp.u32 %0, 1, 0, p;\n"
		    "}\n"
		    : "=r"(done)
		    : "r"(addr), "r"(parity)
		    : "memory");
		if (done) {
			break;
		}
		__nanosleep(ns);
	}
}

// shared -> global bulk copy (TMA store): dst/src 16-byte aligned, bytes % 16 == 0; completion via bulk groups
__device__ __forceinline__ void tma_store_1d(void *gmem_dst, const void *smem_src, uint32_t bytes) {
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_addr(smem_src)),
	             "r"(bytes)
	             : "memory");
	asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// all bulk stores of this thread have finished READING shared memory (the source may be overwritten)
__device__ __forceinline__ void tma_store_wait_read() {
	asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_all() {
	asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
// make this thread's generic-proxy writes to shared memory visible to the async proxy (TMA) before a bulk store
__device__ __forceinline__ void fence_proxy_async_smem() {
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// global -> shared bulk copy (TMA engine, SASS UBLKCP); src/dst 16-byte aligned, bytes % 16 == 0
__device__ __forceinline__ void tma_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
	asm volatile(
	    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
	        smem_addr(smem_dst)),
	    "l"(gmem_src), "r"(bytes), "r"(smem_addr(bar))
	    : "memory");
}

// ---- 128-bit aggregate ---------------------------------------------------------------------------------

// Per-thread accumulator: 96 bits in three 32-bit limbs; the top limb absorbs carries and sign extension.
struct Acc96 {
	uint32_t a0 = 0, a1 = 0, a2 = 0;
	template <bool SIGNED>
	__device__ __forceinline__ void add(uint64_t v) {
		uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
		uint32_t ext = SIGNED ? (uint32_t)((int32_t)hi >> 31) : 0u;
		asm("add.cc.u32 %0, %0, %3;\n\t"
		    "addc.cc.u32 %1, %1, %4;\n\t"
		    "addc.u32 %2, %2, %5;"
		    : "+r"(a0), "+r"(a1), "+r"(a2)
		    : "r"(lo), "r"(hi), "r"(ext));
	}
	// + (hi : lo) as a 96-bit two's-complement number
	__device__ __forceinline__ void add96(uint64_t lo, uint32_t hi) {
		asm("add.cc.u32 %0, %0, %3;\n\t"
		    "addc.cc.u32 %1, %1, %4;\n\t"
		    "addc.u32 %2, %2, %5;"
		    : "+r"(a0), "+r"(a1), "+r"(a2)
		    : "r"((uint32_t)lo), "r"((uint32_t)(lo >> 32)), "r"(hi));
	}
	__device__ __forceinline__ uint64_t lo64() const { return ((uint64_t)a1 << 32) | a0; }
	__device__ __forceinline__ uint64_t hi64() const { return (uint64_t)(int64_t)(int32_t)a2; }
};

struct Agg {  // == adacom_agg
	uint64_t sum_lo, sum_hi, count, reserved;
};

__device__ __forceinline__ void add128(uint64_t &lo, uint64_t &hi, uint64_t olo, uint64_t ohi) {
	uint64_t s = lo + olo;
	hi += ohi + (s < lo ? 1ull : 0ull);
	lo = s;
}

// Block-wide reduction of (lo, hi, count) and ONE atomic triple per block into *out.
// scratch: 3 * 32 uint64_t of shared memory.
__device__ __forceinline__ void block_reduce_agg(uint64_t lo, uint64_t hi, uint64_t cnt, uint64_t *scratch, Agg *out) {
	const unsigned full = 0xffffffffu;
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) {
		uint64_t olo = __shfl_down_sync(full, lo, d);
		uint64_t ohi = __shfl_down_sync(full, hi, d);
		uint64_t oc = __shfl_down_sync(full, cnt, d);
		add128(lo, hi, olo, ohi);
		cnt += oc;
	}
	int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	int nwarps = (blockDim.x + 31) >> 5;
	if (lane == 0) {
		scratch[warp] = lo;
		scratch[32 + warp] = hi;
		scratch[64 + warp] = cnt;
	}
	__syncthreads();
	if (warp == 0) {
		lo = lane < nwarps ? scratch[lane] : 0;
		hi = lane < nwarps ? scratch[32 + lane] : 0;
		cnt = lane < nwarps ? scratch[64 + lane] : 0;
#pragma unroll
		for (int d = 16; d > 0; d >>= 1) {
			uint64_t olo = __shfl_down_sync(full, lo, d);
			uint64_t ohi = __shfl_down_sync(full, hi, d);
			uint64_t oc = __shfl_down_sync(full, cnt, d);
			add128(lo, hi, olo, ohi);
			cnt += oc;
		}
		if (lane == 0 && (cnt | lo | hi)) {
			unsigned long long old = atomicAdd((unsigned long long *)&out->sum_lo, (unsigned long long)lo);
			unsigned long long carry = (old + lo < old) ? 1ull : 0ull;
			atomicAdd((unsigned long long *)&out->sum_hi, (unsigned long long)(hi + carry));
			atomicAdd((unsigned long long *)&out->count, (unsigned long long)cnt);
		}
	}
	__syncthreads();
}

// compare two sign/zero-extended T values; truth table packed as bits {lt, eq, gt} of `op_mask`
__host__ __device__ inline uint32_t cmp_mask_for_op(int op) {
	// EQ, NE, LT, GT, LE, GE  (TemplatedFilterSelection, column_segment.cpp:593-655)
	const uint32_t table[6] = {0x2u, 0x5u, 0x1u, 0x4u, 0x3u, 0x6u};
	return table[op];
}

template <bool SIGNED>
__device__ __forceinline__ bool cmp_passes(uint64_t v, uint64_t c, uint32_t op_mask) {
	bool lt = SIGNED ? ((int64_t)v < (int64_t)c) : (v < c);
	uint32_t sel = lt ? 0u : (v == c ? 1u : 2u);
	return (op_mask >> sel) & 1u;
}

}  // namespace adacom
