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

// `bad` (optional, host-mapped or device word): set to 1 + the index of a query whose (segment, row) lies outside the
// column; such a query reads nothing and answers 0.  The host-buffer form of adacom_fetch_rows validates here instead of
// walking the descriptors on the host (that loop cost more than the transfer and the kernel together).
template <int S>
__global__ void fetch_rows_kernel(const SegDesc *desc, const uint8_t *raw, uint64_t n, const uint32_t *seg_idx,
                                  const uint32_t *row_in_seg, int literal, void *out, uint32_t n_segs,
                                  unsigned long long *bad) {
	using T = typename UIntOf<S>::type;
	for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (uint64_t)gridDim.x * blockDim.x) {
		const uint32_t seg = __ldg(seg_idx + q), row = __ldg(row_in_seg + q);
		if (bad != nullptr && (seg >= n_segs || row >= desc[seg].count)) {
			atomicMin(bad, (unsigned long long)q);
			((T *)out)[q] = (T)0;
			continue;
		}
		((T *)out)[q] = (T)fetch_value<S>(desc + seg, raw, (uint64_t)row, literal);
	}
}

// ONE fetch with (segment, row) in the kernel parameters: the single-row slot of the vtable (SuccinctFetchRow) has no
// id array for the device to read back over PCIe -- the launch carries the pair, the answer is one posted write into
// the host-mapped page.
template <int S>
__global__ void fetch_one_kernel(const SegDesc *desc, const uint8_t *raw, uint32_t seg, uint32_t row, int literal, void *out) {
	using T = typename UIntOf<S>::type;
	*(T *)out = (T)fetch_value<S>(desc + seg, raw, (uint64_t)row, literal);
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

// ---- batched equality look-ups on byte-aligned widths, queries of a segment share its chunks -------------------
// The Zipf look-up benchmarks (benchmark/micro/succinct/zipf_distribution.cpp:13-61) send most of their keys to a few
// segments (q = 1: 60 % of 10 000 look-ups land in the first one).  One job = (segment, 16 KiB chunk of its payload,
// slice of <= kPointSlice of the segment's queries): the chunk is read ONCE into registers (4 x 16 bytes per thread)
// and compared against every target of the slice in the FIELD domain (value == c  <=>  field == (c - min) mod 2^(8S),
// and the target must fit w bits) with the SIMD-in-a-word compares; a hit (rare) is resolved per field.
constexpr int kPointThreads = 256;
constexpr uint32_t kPointChunkBytes = 16384;

__global__ void __launch_bounds__(kPointThreads) point_shared_kernel(const SegDesc *desc, const PointJob *jobs,
                                                                      uint32_t n_jobs, const int64_t *q_const,
                                                                      unsigned long long *q_count,
                                                                      unsigned long long *q_first, uint32_t type_bits) {
	__shared__ uint32_t target[kPointSlice];
	__shared__ uint32_t target_ok[kPointSlice];
	const uint32_t tid = threadIdx.x;
	for (uint32_t jb = blockIdx.x; jb < n_jobs; jb += gridDim.x) {
		const PointJob job = jobs[jb];
		const SegDesc *d = desc + job.seg;
		const uint32_t w = d->width;  // 8, 16 or 32 (checked by the host against its mirror)
		const uint32_t count = d->count;
		const uint64_t mn = d->min_factor;
		const uint64_t min_add = mn == ~0ull ? 0ull : mn;
		const uint32_t per = 128u / w;                        // fields per 16 bytes
		const uint32_t chunk_vals = kPointChunkBytes * 8u / w;
		const uint32_t first = job.chunk * chunk_vals;
		const uint32_t cnt = min(chunk_vals, count - first);
		const uint32_t n_vec = (cnt + per - 1u) / per;        // <= 1024
		const uint4 *vec = (const uint4 *)(d->packed_ptr + (uint64_t)job.chunk * kPointChunkBytes);
		uint4 x[4];
#pragma unroll
		for (int u = 0; u < 4; u++) {
			const uint32_t j = tid + u * kPointThreads;
			x[u] = j < n_vec ? __ldg(vec + j) : make_uint4(0u, 0u, 0u, 0u);
		}
		const uint32_t nq = job.q_hi - job.q_lo;
		if (tid < nq) {
			const uint64_t tf = ((uint64_t)__ldg(q_const + job.q_lo + tid) - min_add) & lo_set64(type_bits);
			target[tid] = (uint32_t)tf;
			target_ok[tid] = (tf >> w) == 0 ? 1u : 0u;  // a target that does not fit w bits cannot match
		}
		__syncthreads();
		const uint32_t mask32 = w == 32u ? 0xffffffffu : ((1u << w) - 1u);
		for (uint32_t q = 0; q < nq; q++) {
			if (!target_ok[q]) {
				continue;
			}
			const uint32_t t32 = target[q];
			const uint32_t kk = w == 32u ? t32 : (w == 16u ? t32 * 0x00010001u : t32 * 0x01010101u);
			// "some field of the word equals the target" as a zero-field test of x ^ kk: (t - ones) & ~t & highs is
			// non-zero iff t has a zero field (a filter: the exact fields are resolved below), three instructions a word
			const uint32_t ones = w == 16u ? 0x00010001u : 0x01010101u;
			const uint32_t highs = ones << (w - 1u);
#pragma unroll
			for (int u = 0; u < 4; u++) {
				uint32_t m;
				if (w == 32u) {
					m = (x[u].x == kk) | (x[u].y == kk) | (x[u].z == kk) | (x[u].w == kk);
				} else {
					const uint32_t t0 = x[u].x ^ kk, t1 = x[u].y ^ kk, t2 = x[u].z ^ kk, t3 = x[u].w ^ kk;
					m = (((t0 - ones) & ~t0) | ((t1 - ones) & ~t1) | ((t2 - ones) & ~t2) | ((t3 - ones) & ~t3)) & highs;
				}
				if (m) {
					const uint32_t xs[4] = {x[u].x, x[u].y, x[u].z, x[u].w};
					const uint32_t j = tid + u * kPointThreads;
					for (uint32_t f = 0; f < per; f++) {
						const uint32_t i = j * per + f;
						const uint32_t fb = f * w;
						const uint32_t fv = w == 32u ? xs[fb >> 5] : ((xs[fb >> 5] >> (fb & 31u)) & mask32);
						if (i < cnt && fv == t32) {
							atomicAdd(q_count + job.q_lo + q, 1ull);
							atomicMin(q_first + job.q_lo + q, (unsigned long long)(first + i));
						}
					}
				}
			}
		}
		__syncthreads();  // the targets are rewritten by the next job
	}
}

}  // namespace

cudaError_t launch_point_shared(const SegDesc *desc, const PointJob *jobs, uint32_t n_jobs, const int64_t *q_const,
                                unsigned long long *q_count, unsigned long long *q_first, int type_size, int sm_count,
                                cudaStream_t stream) {
	if (n_jobs == 0) {
		return cudaSuccess;
	}
	const uint32_t resident = (uint32_t)sm_count * 8u;
	point_shared_kernel<<<n_jobs < resident ? n_jobs : resident, kPointThreads, 0, stream>>>(
	    desc, jobs, n_jobs, q_const, q_count, q_first, 8u * (uint32_t)type_size);
	return cudaGetLastError();
}

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
                              const uint32_t *row_in_seg, int literal, void *out, cudaStream_t stream, uint32_t n_segs,
                              unsigned long long *bad) {
	if (n == 0) {
		return cudaSuccess;
	}
	uint64_t blocks = (n + 255) / 256;
	uint32_t grid = (uint32_t)(blocks > 148u * 16u ? 148u * 16u : blocks);
	ADACOM_DISPATCH_SIZE(t.size,
	                     (fetch_rows_kernel<S><<<grid, 256, 0, stream>>>(desc, raw, n, seg_idx, row_in_seg, literal, out, n_segs, bad)));
	return cudaGetLastError();
}

cudaError_t launch_fetch_one(TypeInfo t, const SegDesc *desc, const uint8_t *raw, uint32_t seg, uint32_t row, int literal,
                             void *out, cudaStream_t stream) {
	ADACOM_DISPATCH_SIZE(t.size, (fetch_one_kernel<S><<<1, 1, 0, stream>>>(desc, raw, seg, row, literal, out)));
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
