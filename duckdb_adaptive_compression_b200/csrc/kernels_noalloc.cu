// kernels_noalloc.cu -- adacom_packed_sum_noalloc: fused decode + SUM/COUNT of ONE packed payload WITHOUT any allocation.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Everything the call needs beyond the
// caller's input (packed words) and output (one adacom_agg in device-accessible memory) lives in ONE file-scope device
// array, g_noalloc_scratch: a ticket counter (first 256 bytes) and one 32-byte partial aggregate per 4096-value tile.
//
//   * The array is resident on every GPU that loads the library until the process exits (kNaScratchWords * 8 = 524,544 bytes).
//   * In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the device address is looked up
//     once per device with cudaGetSymbolAddress (scratch_of_current_device) and handed to the kernel as an argument.
//   * Every call shares the array and finds the last call's contents in it: the ticket is cleared on the caller's
//     stream before the launch (and again by the folding block); the partial slots need no clearing -- the fold reads
//     only slots the same launch has written.  Calls that overlap on the GPU (two streams, two threads) race on it.
//
// Value semantics = the fused SUM of adacom_scan_sum over one re-packed segment: value_i = T(min_factor + field_i),
// field_i = bits [i * width, (i + 1) * width) of the LSB-first 64-bit word string (sdsl::int_vector layout), summed as
// sign- or zero-extended T values into a 128-bit two's-complement total.
#include "../../include/adacom_gpu.h"
#include "adacom_kernels.cuh"

#include <mutex>
#include <string>

namespace adacom {

namespace {

constexpr uint32_t kNaThreads = 256;
constexpr uint32_t kNaTileValues = 4096;                           // values per CTA
constexpr uint64_t kNaMaxValues = 1ull << 26;                      // == ADACOM_NOALLOC_MAX_VALUES
constexpr uint32_t kNaMaxTiles = (uint32_t)(kNaMaxValues / kNaTileValues);  // 16384
constexpr uint32_t kNaHeaderWords = 32;                            // 256 bytes: word 0 = ticket
constexpr uint32_t kNaScratchWords = kNaHeaderWords + 4u * kNaMaxTiles;
constexpr int kNaMaxDevices = 64;

// the scratch: ticket | partial aggregates {sum_lo, sum_hi, count, unused} per tile
static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNaScratchWords];

template <int S, bool SIGNED>
__global__ void __launch_bounds__(kNaThreads) packed_sum_noalloc_kernel(const uint64_t *__restrict__ words, uint64_t count,
                                                                         uint32_t w, uint64_t min_factor, uint32_t n_tiles,
                                                                         unsigned long long *scratch, Agg *out) {
	__shared__ uint64_t s_lo[kNaThreads / 32], s_hi[kNaThreads / 32], s_cnt[kNaThreads / 32];
	__shared__ uint32_t s_last;
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	const uint64_t mask = lo_set64(w);
	const uint64_t n_words = (count * w + 63ull) >> 6;
	const uint64_t v0 = (uint64_t)blockIdx.x * kNaTileValues;
	Acc96 acc;
	uint32_t n = 0;
	for (uint32_t k = tid; k < kNaTileValues; k += kNaThreads) {
		const uint64_t i = v0 + k;
		if (i >= count) {
			break;
		}
		const uint64_t bit = i * w;
		const uint64_t idx = bit >> 6;
		const uint32_t off = (uint32_t)(bit & 63ull);
		uint64_t f = __ldg(words + idx) >> off;
		if (off + w > 64u && idx + 1 < n_words) {
			f |= __ldg(words + idx + 1) << (64u - off);
		}
		acc.add<SIGNED>(extend_low<S, SIGNED>(min_factor + (f & mask)));
		n++;
	}
	// block reduction -> one partial
	uint64_t lo = acc.lo64(), hi = acc.hi64(), cnt = n;
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) {
		const uint64_t olo = __shfl_down_sync(0xffffffffu, lo, d);
		const uint64_t ohi = __shfl_down_sync(0xffffffffu, hi, d);
		cnt += __shfl_down_sync(0xffffffffu, cnt, d);
		add128(lo, hi, olo, ohi);
	}
	if (lane == 0) {
		s_lo[warp] = lo;
		s_hi[warp] = hi;
		s_cnt[warp] = cnt;
	}
	__syncthreads();
	if (tid == 0) {
		for (uint32_t x = 1; x < kNaThreads / 32; x++) {
			add128(lo, hi, s_lo[x], s_hi[x]);
			cnt += s_cnt[x];
		}
		unsigned long long *slot = scratch + kNaHeaderWords + 4ull * blockIdx.x;
		slot[0] = lo;
		slot[1] = hi;
		slot[2] = cnt;
		__threadfence();  // the partial is visible before the ticket moves
		const unsigned long long t = atomicAdd(scratch, 1ull);
		s_last = t == (unsigned long long)(n_tiles - 1u) ? 1u : 0u;
	}
	__syncthreads();
	if (!s_last) {
		return;
	}
	// the last CTA to finish folds every partial of THIS launch (slots [0, n_tiles)) into the caller's aggregate
	__threadfence();
	lo = hi = cnt = 0;
	for (uint32_t p = tid; p < n_tiles; p += kNaThreads) {
		const unsigned long long *slot = scratch + kNaHeaderWords + 4ull * p;
		add128(lo, hi, __ldcg(slot), __ldcg(slot + 1));
		cnt += __ldcg(slot + 2);
	}
#pragma unroll
	for (int d = 16; d > 0; d >>= 1) {
		const uint64_t olo = __shfl_down_sync(0xffffffffu, lo, d);
		const uint64_t ohi = __shfl_down_sync(0xffffffffu, hi, d);
		cnt += __shfl_down_sync(0xffffffffu, cnt, d);
		add128(lo, hi, olo, ohi);
	}
	__syncthreads();  // s_lo / s_hi / s_cnt are free again
	if (lane == 0) {
		s_lo[warp] = lo;
		s_hi[warp] = hi;
		s_cnt[warp] = cnt;
	}
	__syncthreads();
	if (tid == 0) {
		for (uint32_t x = 1; x < kNaThreads / 32; x++) {
			add128(lo, hi, s_lo[x], s_hi[x]);
			cnt += s_cnt[x];
		}
		out->sum_lo = lo;
		out->sum_hi = hi;
		out->count = cnt;
		out->reserved = 0;
		scratch[0] = 0ull;  // leave the ticket clean (the next call clears it again on its own stream)
	}
}

std::mutex g_na_mutex;
unsigned long long *g_na_scratch_ptr[kNaMaxDevices] = {};

// device address of g_noalloc_scratch on the CURRENT device: looked up once per device
cudaError_t scratch_of_current_device(unsigned long long **out, int *device_out) {
	int dev = -1;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) {
		return e;
	}
	if (dev < 0 || dev >= kNaMaxDevices) {
		return cudaErrorInvalidDevice;
	}
	std::lock_guard<std::mutex> lock(g_na_mutex);
	if (g_na_scratch_ptr[dev] == nullptr) {
		void *p = nullptr;
		e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
		if (e != cudaSuccess) {
			return e;
		}
		if (p == nullptr) {
			return cudaErrorInvalidSymbol;
		}
		g_na_scratch_ptr[dev] = (unsigned long long *)p;
	}
	*out = g_na_scratch_ptr[dev];
	*device_out = dev;
	return cudaSuccess;
}

adacom_status na_fail(adacom_status st, const std::string &msg) {
	set_process_error(msg);
	return st;
}

template <int S, bool SIGNED>
cudaError_t launch_na(const uint64_t *words, uint64_t count, uint32_t w, uint64_t min_factor, uint32_t n_tiles,
                      unsigned long long *scratch, Agg *out, cudaStream_t stream) {
	packed_sum_noalloc_kernel<S, SIGNED><<<n_tiles, kNaThreads, 0, stream>>>(words, count, w, min_factor, n_tiles, scratch, out);
	return cudaGetLastError();
}

}  // namespace

}  // namespace adacom

using namespace adacom;

static_assert(ADACOM_NOALLOC_MAX_VALUES == (1ull << 26), "header and kernel disagree on the no-alloc maximum");

extern "C" adacom_status adacom_packed_sum_noalloc(const uint64_t *device_words, uint64_t count, uint32_t width,
                                                   uint64_t min_factor, adacom_ptype type, void *stream,
                                                   adacom_agg *device_out) {
	int S = 0;
	bool is_signed = false;
	switch (type) {
	case ADACOM_UINT8: S = 1; break;
	case ADACOM_INT8: S = 1; is_signed = true; break;
	case ADACOM_UINT16: S = 2; break;
	case ADACOM_INT16: S = 2; is_signed = true; break;
	case ADACOM_UINT32: S = 4; break;
	case ADACOM_INT32: S = 4; is_signed = true; break;
	case ADACOM_UINT64: S = 8; break;
	case ADACOM_INT64: S = 8; is_signed = true; break;
	default:
		return na_fail(ADACOM_ERR_UNSUPPORTED, "adacom_packed_sum_noalloc: unsupported physical type " + std::to_string((int)type));
	}
	if (count > kNaMaxValues) {
		return na_fail(ADACOM_ERR_INVALID, "adacom_packed_sum_noalloc: count " + std::to_string(count) +
		                                       " exceeds ADACOM_NOALLOC_MAX_VALUES (" + std::to_string(kNaMaxValues) +
		                                       "): the static scratch holds one partial per 4096 values; split the call");
	}
	if (device_out == nullptr) {
		return na_fail(ADACOM_ERR_INVALID, "adacom_packed_sum_noalloc: device_out is NULL");
	}
	if (width == 0 || width > 8u * (uint32_t)S) {
		return na_fail(ADACOM_ERR_INVALID, "adacom_packed_sum_noalloc: width " + std::to_string(width) + " outside [1, " +
		                                       std::to_string(8 * S) + "]");
	}
	if (count != 0 && device_words == nullptr) {
		return na_fail(ADACOM_ERR_INVALID, "adacom_packed_sum_noalloc: device_words is NULL");
	}
	if (((uintptr_t)device_words & 7u) != 0 || ((uintptr_t)device_out & 7u) != 0) {
		return na_fail(ADACOM_ERR_INVALID, "adacom_packed_sum_noalloc: device_words / device_out must be 8-byte aligned");
	}
	unsigned long long *scratch = nullptr;
	int dev = 0;
	cudaError_t e = scratch_of_current_device(&scratch, &dev);
	if (e != cudaSuccess) {
		return na_fail(ADACOM_ERR_CUDA, std::string("adacom_packed_sum_noalloc: cudaGetSymbolAddress(g_noalloc_scratch): ") +
		                                    cudaGetErrorString(e));
	}
	cudaStream_t s = (cudaStream_t)stream;
	// the ticket must start at zero whatever the previous call left behind
	e = cudaMemsetAsync(scratch, 0, kNaHeaderWords * sizeof(unsigned long long), s);
	if (e != cudaSuccess) {
		return na_fail(ADACOM_ERR_CUDA, std::string("adacom_packed_sum_noalloc: cudaMemsetAsync: ") + cudaGetErrorString(e));
	}
	const uint32_t n_tiles = count == 0 ? 1u : (uint32_t)((count + kNaTileValues - 1) / kNaTileValues);
	Agg *out = (Agg *)device_out;
	switch (S * 2 + (is_signed ? 1 : 0)) {
	case 2: e = launch_na<1, false>(device_words, count, width, min_factor, n_tiles, scratch, out, s); break;
	case 3: e = launch_na<1, true>(device_words, count, width, min_factor, n_tiles, scratch, out, s); break;
	case 4: e = launch_na<2, false>(device_words, count, width, min_factor, n_tiles, scratch, out, s); break;
	case 5: e = launch_na<2, true>(device_words, count, width, min_factor, n_tiles, scratch, out, s); break;
	case 8: e = launch_na<4, false>(device_words, count, width, min_factor, n_tiles, scratch, out, s); break;
	case 9: e = launch_na<4, true>(device_words, count, width, min_factor, n_tiles, scratch, out, s); break;
	case 16: e = launch_na<8, false>(device_words, count, width, min_factor, n_tiles, scratch, out, s); break;
	default: e = launch_na<8, true>(device_words, count, width, min_factor, n_tiles, scratch, out, s); break;
	}
	if (e != cudaSuccess) {
		return na_fail(ADACOM_ERR_CUDA, std::string("adacom_packed_sum_noalloc: launch: ") + cudaGetErrorString(e));
	}
	return ADACOM_OK;
}
