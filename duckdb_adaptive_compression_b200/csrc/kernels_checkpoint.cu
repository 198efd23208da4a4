// kernels_checkpoint.cu -- the persisted form of a succinct column.
//
// The reference never persists succinct segments (column_segment.cpp:529-533 converts nothing; the
// init_compression/compress/compress_finalize slots are stubs, succinct.cpp:86-119).  What DuckDB does for every other
// codec is ColumnCheckpointState::FlushSegment (src/storage/table/column_checkpoint_state.cpp:91-163): the segment
// payload is copied into a 256 KiB block -- several small segments share one block (PartialBlockManager,
// src/storage/partial_block_manager.cpp:14-77) -- and FileBuffer::ChecksumAndWrite (src/common/file_buffer.cpp:113-119)
// stores Checksum(payload) (src/common/checksum.cpp:6-25) in the first 8 bytes of the block.
//
//   checkpoint_blocks_kernel : payloads (packed words, or the slot-width values of a segment that was not re-packed)
//                              -> block images, gap/tail bytes zeroed, checksum written; ONE pass: every payload byte
//                              is read once and every block byte written once (algorithmic bytes = Σ payload read +
//                              n_blocks * 262144 written)
//   restore_blocks_kernel    : block images -> packed extent (128-byte aligned payloads, alignment gap zeroed) / raw
//                              arena, verifying every block checksum in the same pass (n_blocks * 262144 read +
//                              Σ payload written)
//
// One CTA per block at a time (grid-stride); a block holds <= a handful of pieces, walked in offset order.  The
// checksum is an XOR of per-word products, so the order of accumulation is free.
#include "adacom_kernels.cuh"

namespace adacom {

namespace {

constexpr uint64_t kChecksumMul = 0xbf58476d1ce4e5b9ull;  // Checksum(uint64_t), checksum.cpp:6-8
constexpr uint64_t kChecksumSeed = 5381;                  // checksum.cpp:11
constexpr int kCkThreads = 512;

__device__ __forceinline__ uint64_t block_xor(uint64_t x, uint64_t *smem) {
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		x ^= __shfl_xor_sync(0xffffffffu, x, o);
	}
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	__syncthreads();  // smem reuse across grid-stride iterations
	if (lane == 0) {
		smem[warp] = x;
	}
	__syncthreads();
	if (warp == 0) {
		x = lane < (kCkThreads >> 5) ? smem[lane] : 0;
#pragma unroll
		for (int o = 8; o > 0; o >>= 1) {
			x ^= __shfl_xor_sync(0xffffffffu, x, o);
		}
	}
	return x;  // valid in thread 0
}

__device__ __forceinline__ uint64_t tail_mask(uint32_t bytes, uint32_t nw) {
	const uint32_t vb = bytes - 8u * (nw - 1u);  // valid bytes of the last word: 1..8
	return vb >= 8u ? ~0ull : ((1ull << (8u * vb)) - 1ull);
}

__global__ void __launch_bounds__(kCkThreads, 3) checkpoint_blocks_kernel(const CkPiece *__restrict__ pieces,
                                                                       const uint32_t *__restrict__ piece_begin,
                                                                       uint8_t *__restrict__ blocks, uint32_t n_blocks) {
	__shared__ uint64_t red[kCkThreads / 32];
	const uint32_t tid = threadIdx.x;
	for (uint32_t b = blockIdx.x; b < n_blocks; b += gridDim.x) {
		uint64_t *dst = (uint64_t *)(blocks + (size_t)b * kBlockAllocSize) + 1;
		uint64_t x = 0;
		uint32_t cursor = 0;
		const uint32_t pb = __ldg(piece_begin + b), pe = __ldg(piece_begin + b + 1);
		for (uint32_t p = pb; p < pe; p++) {
			const CkPiece pc = pieces[p];
			const uint32_t w0 = pc.block_off >> 3;
			for (uint32_t i = cursor + tid; i < w0; i += kCkThreads) {
				dst[i] = 0;
			}
			const uint32_t nw = (pc.bytes + 7u) >> 3;
			const uint32_t mis = (uint32_t)(pc.addr & 7ull);
			const uint64_t *src = (const uint64_t *)(pc.addr - mis);
			const uint64_t last_mask = tail_mask(pc.bytes, nw);
			if (mis == 0) {
#pragma unroll 4
				for (uint32_t i = tid; i < nw; i += kCkThreads) {
					uint64_t v = __ldg(src + i);
					if (i == nw - 1) {
						v &= last_mask;
					}
					dst[w0 + i] = v;
					x ^= v * kChecksumMul;
				}
			} else {
				// slot-width payload of a narrow type whose segment starts off an 8-byte boundary: two aligned
				// words, funnelled (the arena carries 256 bytes of slack behind its last value)
				const uint32_t sh = 8u * mis;
#pragma unroll 2
				for (uint32_t i = tid; i < nw; i += kCkThreads) {
					const uint64_t lo = __ldg(src + i), hi = __ldg(src + i + 1);
					uint64_t v = (lo >> sh) | (hi << (64u - sh));
					if (i == nw - 1) {
						v &= last_mask;
					}
					dst[w0 + i] = v;
					x ^= v * kChecksumMul;
				}
			}
			cursor = w0 + nw;
		}
		for (uint32_t i = cursor + tid; i < kBlockSize / 8u; i += kCkThreads) {
			dst[i] = 0;
		}
		x = block_xor(x, red);
		if (tid == 0) {
			dst[-1] = kChecksumSeed ^ x;
		}
	}
}

template <int S>
__device__ __forceinline__ void store_elems(uint64_t addr, uint32_t word, uint64_t v, uint32_t bytes) {
	using T = typename UIntOf<S>::type;
	T *d = (T *)addr + (size_t)word * (8 / S);
#pragma unroll
	for (int e = 0; e < 8 / S; e++) {
		if (word * 8u + (uint32_t)e * S < bytes) {
			d[e] = (T)(v >> (8 * e * S));
		}
	}
}

__global__ void __launch_bounds__(kCkThreads, 3) restore_blocks_kernel(const CkPiece *__restrict__ pieces,
                                                                    const uint32_t *__restrict__ piece_begin,
                                                                    const uint8_t *__restrict__ blocks,
                                                                    uint32_t n_blocks, unsigned long long *bad_block) {
	__shared__ uint64_t red[kCkThreads / 32];
	const uint32_t tid = threadIdx.x;
	for (uint32_t b = blockIdx.x; b < n_blocks; b += gridDim.x) {
		const uint64_t *src = (const uint64_t *)(blocks + (size_t)b * kBlockAllocSize) + 1;
		uint64_t x = 0;
		uint32_t cursor = 0;
		const uint32_t pb = __ldg(piece_begin + b), pe = __ldg(piece_begin + b + 1);
		for (uint32_t p = pb; p < pe; p++) {
			const CkPiece pc = pieces[p];
			const uint32_t w0 = pc.block_off >> 3;
			for (uint32_t i = cursor + tid; i < w0; i += kCkThreads) {
				x ^= __ldg(src + i) * kChecksumMul;
			}
			const uint32_t nw = (pc.bytes + 7u) >> 3;
			const uint32_t full = pc.bytes >> 3;  // words every byte of which belongs to the payload
			const bool aligned = (pc.addr & 7ull) == 0;
			if (aligned) {
				uint64_t *dst = (uint64_t *)pc.addr;
#pragma unroll 4
				for (uint32_t i = tid; i < nw; i += kCkThreads) {
					const uint64_t v = __ldg(src + w0 + i);
					x ^= v * kChecksumMul;
					if (i < full) {
						dst[i] = v;
					} else {
						switch (pc.elem_size) {
						case 1: store_elems<1>(pc.addr, i, v, pc.bytes); break;
						case 2: store_elems<2>(pc.addr, i, v, pc.bytes); break;
						default: store_elems<4>(pc.addr, i, v, pc.bytes); break;
						}
					}
				}
				// alignment gap behind a packed payload (the scans' 16-byte rounded bulk copies may touch it)
				for (uint32_t i = nw + tid; i < pc.pad_words; i += kCkThreads) {
					dst[i] = 0;
				}
			} else {
				for (uint32_t i = tid; i < nw; i += kCkThreads) {
					const uint64_t v = __ldg(src + w0 + i);
					x ^= v * kChecksumMul;
					switch (pc.elem_size) {
					case 1: store_elems<1>(pc.addr, i, v, pc.bytes); break;
					case 2: store_elems<2>(pc.addr, i, v, pc.bytes); break;
					default: store_elems<4>(pc.addr, i, v, pc.bytes); break;
					}
				}
			}
			cursor = w0 + nw;
		}
#pragma unroll 4
		for (uint32_t i = cursor + tid; i < kBlockSize / 8u; i += kCkThreads) {
			x ^= __ldg(src + i) * kChecksumMul;
		}
		x = block_xor(x, red);
		if (tid == 0 && (kChecksumSeed ^ x) != src[-1]) {
			atomicMin(bad_block, (unsigned long long)b);
		}
	}
}

}  // namespace

static uint32_t checkpoint_grid(uint32_t n_blocks, int sm_count) {
	const uint32_t resident = (uint32_t)sm_count * 3u;
	return n_blocks < resident ? n_blocks : resident;
}

cudaError_t launch_checkpoint_blocks(const CkPiece *pieces, const uint32_t *piece_begin, uint8_t *blocks,
                                     uint32_t n_blocks, int sm_count, cudaStream_t stream) {
	if (n_blocks == 0) {
		return cudaSuccess;
	}
	checkpoint_blocks_kernel<<<checkpoint_grid(n_blocks, sm_count), kCkThreads, 0, stream>>>(pieces, piece_begin, blocks,
	                                                                                        n_blocks);
	return cudaGetLastError();
}

cudaError_t launch_restore_blocks(const CkPiece *pieces, const uint32_t *piece_begin, const uint8_t *blocks,
                                  uint32_t n_blocks, unsigned long long *bad_block, int sm_count, cudaStream_t stream) {
	if (n_blocks == 0) {
		return cudaSuccess;
	}
	restore_blocks_kernel<<<checkpoint_grid(n_blocks, sm_count), kCkThreads, 0, stream>>>(pieces, piece_begin, blocks,
	                                                                                     n_blocks, bad_block);
	return cudaGetLastError();
}

}  // namespace adacom
