// kernels_misc.cu -- single-vector scan (generic unaligned path), batched fetch_row gather and the
// synthetic-data generator.
//
//   scan_range  <- SuccinctScanPartial<T> for ONE call (src/storage/compression/succinct.cpp:123-144):
//                  arbitrary `start` (DuckDB vectors are aligned to the row group, segments are not:
//                  start = 2048k + c), arbitrary scan_count, result_offset.  Launch-latency bound; reads the
//                  two covering words straight from global memory.
//   fetch_rows  <- the intended SuccinctFetchRow<T> (succinct.cpp:244-260):
//                  out[q] = T(vec[row] + min) for a batch of (segment, row) pairs; one thread per fetch.
#include "adacom_kernels.cuh"

namespace adacom {

namespace {

template <int S>
__device__ __forceinline__ uint64_t fetch_value(const SegDesc *d, const uint8_t *raw, uint64_t row, int literal) {
	using T = typename UIntOf<S>::type;
	const uint32_t state = d->state;
	const uint64_t mn = d->min_factor;
	const bool repacked = state & kSegRepacked;
	uint64_t f;
	if (repacked) {
		const uint32_t w = d->width;
		const uint64_t *words = (const uint64_t *)d->packed_ptr;
		const uint64_t bit = row * w;
		const uint64_t idx = bit >> 6;
		const uint32_t off = (uint32_t)(bit & 63u);
		uint64_t lo = __ldg(words + idx);
		f = lo >> off;
		if (off + w > 64u) {
			uint64_t hi = __ldg(words + idx + 1);
			f |= hi << (64u - off);
		}
		f &= lo_set64(w);
	} else {
		f = (uint64_t)__ldg((const T *)raw + d->row_start + row);
	}
	const bool add = (mn != ~0ull) && (repacked || literal);
	return add ? f + mn : f;
}

template <int S>
__global__ void scan_range_kernel(const SegDesc *desc, uint32_t seg, const uint8_t *raw, uint64_t start,
                                  uint64_t scan_count, int literal, void *result, uint64_t result_offset) {
	using T = typename UIntOf<S>::type;
	const SegDesc *d = desc + seg;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < scan_count;
	     i += (uint64_t)gridDim.x * blockDim.x) {
		((T *)result)[result_offset + i] = (T)fetch_value<S>(d, raw, start + i, literal);
	}
}

template <int S>
__global__ void fetch_rows_kernel(const SegDesc *desc, const uint8_t *raw, uint64_t n, const uint32_t *seg_idx,
                                  const uint32_t *row_in_seg, int literal, void *out) {
	using T = typename UIntOf<S>::type;
	for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (uint64_t)gridDim.x * blockDim.x) {
		const SegDesc *d = desc + __ldg(seg_idx + q);
		((T *)out)[q] = (T)fetch_value<S>(d, raw, (uint64_t)__ldg(row_in_seg + q), literal);
	}
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
	x += 0x9E3779B97F4A7C15ull;
	x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
	x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
	return x ^ (x >> 31);
}

template <int S>
__global__ void fill_splitmix_kernel(void *raw, uint64_t n, uint64_t seed, uint64_t first_index, uint64_t mask) {
	using T = typename UIntOf<S>::type;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		((T *)raw)[i] = (T)(splitmix64(seed + first_index + i) & mask);
	}
}

}  // namespace

#define ADACOM_DISPATCH_SIZE(size, CALL) \
	switch (size) {                      \
	case 1: { constexpr int S = 1; CALL; } break; \
	case 2: { constexpr int S = 2; CALL; } break; \
	case 4: { constexpr int S = 4; CALL; } break; \
	case 8: { constexpr int S = 8; CALL; } break; \
	default: return cudaErrorInvalidValue; \
	}

cudaError_t launch_scan_range(TypeInfo t, const SegDesc *desc, uint32_t seg, const uint8_t *raw, uint64_t start,
                              uint64_t scan_count, int literal, void *result, uint64_t result_offset,
                              cudaStream_t stream) {
	if (scan_count == 0) {
		return cudaSuccess;
	}
	uint32_t grid = (uint32_t)((scan_count + 255) / 256);
	if (grid > 1184) {
		grid = 1184;
	}
	ADACOM_DISPATCH_SIZE(t.size, (scan_range_kernel<S><<<grid, 256, 0, stream>>>(desc, seg, raw, start, scan_count,
	                                                                            literal, result, result_offset)));
	return cudaGetLastError();
}

cudaError_t launch_fetch_rows(TypeInfo t, const SegDesc *desc, const uint8_t *raw, uint64_t n, const uint32_t *seg_idx,
                              const uint32_t *row_in_seg, int literal, void *out, cudaStream_t stream) {
	if (n == 0) {
		return cudaSuccess;
	}
	uint64_t blocks = (n + 255) / 256;
	uint32_t grid = (uint32_t)(blocks > 148u * 16u ? 148u * 16u : blocks);
	ADACOM_DISPATCH_SIZE(t.size,
	                     (fetch_rows_kernel<S><<<grid, 256, 0, stream>>>(desc, raw, n, seg_idx, row_in_seg, literal, out)));
	return cudaGetLastError();
}

cudaError_t launch_fill_splitmix(TypeInfo t, void *raw, uint64_t n, uint64_t seed, uint64_t first_index, uint64_t mask,
                                 cudaStream_t stream) {
	if (n == 0) {
		return cudaSuccess;
	}
	uint64_t blocks = (n + 255) / 256;
	uint32_t grid = (uint32_t)(blocks > 148u * 32u ? 148u * 32u : blocks);
	ADACOM_DISPATCH_SIZE(t.size, (fill_splitmix_kernel<S><<<grid, 256, 0, stream>>>(raw, n, seed, first_index, mask)));
	return cudaGetLastError();
}

}  // namespace adacom
