// kernels_filter.cu -- fused decode + (value <op> constant) + SUM/COUNT (+ selection bitmap) for columns whose live
// payloads are all <= 32 bits wide, evaluated in the FIELD domain.
//
// Reference: ColumnData::Select = scan, then ColumnSegment::FilterSelection -> TemplatedFilterSelection
// (src/storage/table/column_data.cpp:199-204, src/storage/table/column_segment.cpp:575-658), then the aggregate.
// decode_kernel<..., filter> does that per value (decode, extend, compare as T, 96-bit add) and is bound by its
// per-value instruction budget (~400 Gvalues/s).  Here, for a chunk whose values are extend(min) + f without wrap-around
// (the whole interval [min, min + 2^w - 1] lies inside one sign / truncation zone of T):
//     value <op> c   <=>   f <op> (c - extend(min))          (a constant outside the interval decides the whole chunk)
//     SUM(passing)    =    SUM(f) + n * extend(min)
// so a value costs a funnel shift, a mask, two compares, a select and a 64-bit add.  Chunks that are not linear (raw
// signed slots, wrap-around frames) take the per-value path inside the same kernel.
//
// Job = (segment, chunk of 4096 values): 512 * w payload bytes (16-byte aligned: 64 values are w whole words), staged
// in shared memory with 16-byte loads; 128 threads, 32 values each, lane = consecutive rows so that a ballot is 32
// consecutive selection bits.  Algorithmic bytes: the packed payload (+ 1 bit per row for the bitmap).
#include "adacom_kernels.cuh"
#include "adacom_pred32.cuh"

namespace adacom {

namespace {

constexpr int kFiltThreads = 128;  // small CTAs, 8 per SM: a job is a chain of dependent loads (descriptor -> payload)
constexpr uint32_t kFiltChunk = 4096;                      // values per job
constexpr uint32_t kFiltStageBytes = kFiltChunk * 32 / 8;  // widest staged chunk (w = 32)

template <int S, bool SIGNED>
__global__ void __launch_bounds__(kFiltThreads, 8) filter_fields_kernel(const SegDesc *desc, uint32_t seg_b, uint32_t n_segs,
                                                                        uint32_t cps_shift, const uint8_t *raw,
                                                                        int literal, uint32_t op_mask, uint64_t constant,
                                                                        uint32_t *bitmap32, uint64_t out_row0, Agg *agg) {
	using T = typename UIntOf<S>::type;
	__shared__ __align__(16) uint32_t stage[kFiltStageBytes / 4 + 4];
	__shared__ uint64_t red[96];  // block_reduce_agg scratch
	// selection bits of the chunk at their position inside the output words (the chunk's first row is not 32-aligned):
	// ballots are OR-ed in here, interior words then leave as plain stores, only the two boundary words (shared with the
	// neighbouring chunks) as global atomics
	__shared__ uint32_t bm[kFiltChunk / 32 + 2];
	const uint32_t tid = threadIdx.x, lane = tid & 31u;
	Acc96 acc;
	uint32_t n_pass = 0;
	const uint32_t n_jobs = n_segs << cps_shift;  // chunks per segment rounded up to a power of two, empty ones skipped
	for (uint32_t job = blockIdx.x; job < n_jobs; job += gridDim.x) {
		const SegDesc *d = desc + seg_b + (job >> cps_shift);
		const uint32_t first = (job & ((1u << cps_shift) - 1u)) * kFiltChunk;
		const uint32_t count = d->count;
		if (first >= count) {
			continue;  // uniform per CTA
		}
		const uint32_t cnt = min(kFiltChunk, count - first);
		const uint32_t state = d->state;
		const bool repacked = state & kSegRepacked;
		const uint32_t w = repacked ? d->width : 8u * S;
		const uint64_t mn = d->min_factor;
		const uint64_t min_add = (mn != ~0ull && (repacked || literal)) ? mn : 0ull;
		const uint64_t row0 = d->row_start + first;  // column row of the chunk's first value
		const uint64_t mask64 = lo_set64(w);
		const uint64_t e_min = extend_low<S, SIGNED>(min_add);
		const uint64_t top = min_add + mask64;
		const uint64_t e_top = extend_low<S, SIGNED>(top);
		const bool linear = repacked && w <= 32u && top >= min_add && e_top - e_min == mask64 &&
		                    (SIGNED ? (int64_t)e_min <= (int64_t)e_top : e_min <= e_top);
		const uint64_t rel0 = row0 - out_row0;  // chunk's first row relative to the bitmap
		const uint32_t bsh = (uint32_t)(rel0 & 31u);
		if (bitmap32 != nullptr) {
			__syncthreads();  // the previous chunk's words have been written out
			for (uint32_t t = tid; t < kFiltChunk / 32 + 2; t += kFiltThreads) {
				bm[t] = 0u;
			}
			__syncthreads();
		}
		if (repacked && w <= 32u) {
			// stage the chunk: 512 * w bytes rounded up to 16 (the payload is followed by its alignment gap / arena pad)
			const uint4 *src = (const uint4 *)(d->packed_ptr + (uint64_t)(first / 64u) * w * 8u);
			const uint32_t n16 = (cnt * w + 127u) >> 7;
			__syncthreads();  // the previous chunk has been consumed
			for (uint32_t j = tid; j < n16; j += kFiltThreads) {
				((uint4 *)stage)[j] = __ldg(src + j);
			}
			if (tid == 0) {
				stage[n16 * 4] = 0u;  // the funnel's second word behind the last field
			}
			__syncthreads();
		}
		if (linear) {
			const bool below = SIGNED ? (int64_t)constant < (int64_t)e_min : constant < e_min;
			const bool above = SIGNED ? (int64_t)constant > (int64_t)e_top : constant > e_top;
			const bool p_lt = op_mask & 1u, p_gt = (op_mask >> 2) & 1u;
			// a constant below the interval: every value is greater; above: every value is smaller
			const int whole = below ? (p_gt ? 1 : 0) : (above ? (p_lt ? 1 : 0) : 2);
			if (whole == 0) {
				continue;  // nothing passes, the bitmap was zeroed by the caller
			}
			const uint32_t fc = (uint32_t)(constant - e_min);
			const uint32_t mask32 = (uint32_t)mask64;
			// the operator as one inclusive field range [flo, flo + span] (inv: its complement), one compare per value
			uint32_t flo = 0, span = mask32;
			bool inv = false, none = false;
			if (whole == 2) {
				switch (op_mask & 7u) {
				case 1u: none = fc == 0u; span = fc - 1u; break;                      // <
				case 3u: span = fc; break;                                            // <=
				case 4u: none = fc == mask32; flo = fc + 1u; span = mask32 - flo; break;  // >
				case 6u: flo = fc; span = mask32 - fc; break;                         // >=
				case 2u: flo = fc; span = 0u; break;                                  // =
				case 5u: flo = fc; span = 0u; inv = true; break;                      // <>
				case 0u: none = true; break;
				default: break;                                                       // everything passes
				}
			}
			if (none) {
				continue;
			}
			uint64_t fsum = 0;
			uint32_t n = 0;
			if (cnt == kFiltChunk && bitmap32 == nullptr) {
				// full chunk, no bitmap: the distance between a thread's values is kFiltThreads * w bits, whole 32-bit
				// words, so its funnel shift never changes and the loads sit at fixed strides from one pointer
				const uint32_t bit0 = tid * w;
				const uint32_t *p32 = stage + (bit0 >> 5);
				const uint32_t sh = bit0 & 31u;
				const uint32_t stride = (uint32_t)kFiltThreads / 32u * w;
				if (w <= 26u) {
					uint32_t s32 = 0;  // 32 fields below 2^26 cannot overflow 32 bits
#pragma unroll
					for (int jj = 0; jj < (int)(kFiltChunk / kFiltThreads); jj++) {
						const uint32_t f = __funnelshift_r(p32[jj * stride], p32[jj * stride + 1], sh) & mask32;
						const bool pass = ((f - flo) <= span) != inv;
						s32 += pass ? f : 0u;
						n += pass ? 1u : 0u;
					}
					fsum = s32;
				} else {
#pragma unroll
					for (int jj = 0; jj < (int)(kFiltChunk / kFiltThreads); jj++) {
						const uint32_t f = __funnelshift_r(p32[jj * stride], p32[jj * stride + 1], sh) & mask32;
						const bool pass = ((f - flo) <= span) != inv;
						fsum += pass ? (uint64_t)f : 0ull;
						n += pass ? 1u : 0u;
					}
				}
			} else {
				for (uint32_t base = 0; base < cnt; base += kFiltThreads) {
					const uint32_t i = base + tid;
					const bool valid = i < cnt;
					const uint32_t bit = (valid ? i : 0u) * w;
					const uint32_t f = __funnelshift_r(stage[bit >> 5], stage[(bit >> 5) + 1], bit) & mask32;
					const bool pass = valid && (((f - flo) <= span) != inv);
					fsum += pass ? (uint64_t)f : 0ull;
					n += pass ? 1u : 0u;
					if (bitmap32 != nullptr) {
						const uint32_t ballot = __ballot_sync(0xffffffffu, pass);
						if (lane == 0 && ballot) {
							const uint32_t k = (base + (tid & ~31u)) >> 5;
							atomicOr(&bm[k], ballot << bsh);
							if (bsh) {
								atomicOr(&bm[k + 1], ballot >> (32u - bsh));
							}
						}
					}
				}
			}
			// + fsum + n * extend(min) as a 96-bit two's-complement number
			const uint64_t plo = e_min * (uint64_t)n;
			uint64_t phi = __umul64hi(e_min, (uint64_t)n);
			if (SIGNED && (int64_t)e_min < 0) {
				phi -= (uint64_t)n;  // signed x unsigned high part
			}
			uint64_t lo = plo, hi = phi;
			add128(lo, hi, fsum, 0ull);
			acc.add96(lo, (uint32_t)hi);
			n_pass += n;
		} else {
			// per-value path: raw slots (read from the raw arena) or a frame of reference that wraps inside T
			const T *slots = (const T *)raw + row0;
			for (uint32_t base = 0; base < cnt; base += kFiltThreads) {
				const uint32_t i = base + tid;
				const bool valid = i < cnt;
				bool pass = false;
				uint64_t v = 0;
				if (valid) {
					uint64_t f;
					if (repacked && w <= 32u) {
						const uint32_t bit = i * w;
						f = (uint64_t)(__funnelshift_r(stage[bit >> 5], stage[(bit >> 5) + 1], bit) & (uint32_t)mask64);
					} else if (repacked) {
						// wide payloads are not staged (the host routes such columns to decode_kernel): two words from HBM
						const uint64_t *words = (const uint64_t *)d->packed_ptr;
						const uint64_t bit = (uint64_t)(first + i) * w;
						const uint32_t off = (uint32_t)(bit & 63u);
						f = __ldg(words + (bit >> 6)) >> off;
						if (off + w > 64u) {
							f |= __ldg(words + (bit >> 6) + 1) << (64u - off);
						}
						f &= mask64;
					} else {
						f = (uint64_t)__ldg(slots + i);
					}
					v = extend_low<S, SIGNED>(f + min_add);
					pass = cmp_passes<SIGNED>(v, constant, op_mask);
				}
				if (pass) {
					acc.add<SIGNED>(v);
					n_pass++;
				}
				if (bitmap32 != nullptr) {
					const uint32_t ballot = __ballot_sync(0xffffffffu, pass);
					if (lane == 0 && ballot) {
						const uint32_t k = (base + (tid & ~31u)) >> 5;
						atomicOr(&bm[k], ballot << bsh);
						if (bsh) {
							atomicOr(&bm[k + 1], ballot >> (32u - bsh));
						}
					}
				}
			}
		}
		if (bitmap32 != nullptr) {
			__syncthreads();  // every ballot of the chunk is in bm
			const uint32_t words = (bsh + cnt + 31u) >> 5;
			uint32_t *dst = bitmap32 + (rel0 >> 5);
			for (uint32_t t = tid; t < words; t += kFiltThreads) {
				const uint32_t v = bm[t];
				if (v) {
					if (t == 0 || t == words - 1u) {
						atomicOr(dst + t, v);
					} else {
						dst[t] = v;
					}
				}
			}
		}
	}
	block_reduce_agg(acc.lo64(), acc.hi64(), (uint64_t)n_pass, red, agg);
}

// =====================================================================================================================
// filter_words_kernel -- the same operator with a LANE per 32-row group (pred32: compile-time field positions, five
// integer instructions per value instead of thirteen) and a warp per 2048-value tile: no CTA barrier, the tile is staged
// with cp.async into the warp's own window, the selection words leave as whole 32-bit stores (only the two words a tile
// shares with its neighbours are atomics).  Segments that are not field-comparable (frames that wrap inside T, payloads
// wider than 32 bits, slot-width segments) take a lane = row path in the same kernel, straight from global memory.
// =====================================================================================================================
template <bool SIGNED>
__device__ __forceinline__ bool t_less(uint64_t a, uint64_t b) {
	return SIGNED ? (int64_t)a < (int64_t)b : a < b;
}

constexpr int kFWarps = 8;
constexpr uint32_t kFRows = 2048;
constexpr uint32_t kFRun = 4;

__device__ __forceinline__ void f_cp_async16(void *smem_dst, const void *gmem_src) {
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(smem_dst)), "l"(gmem_src) : "memory");
}

template <int S, bool SIGNED, bool ALLOW>
__global__ void __launch_bounds__(kFWarps * 32, 3) filter_words_kernel(const SegDesc *desc, const Job *tiles, uint32_t n_tiles,
                                                                       const uint8_t *raw, int literal, const FilterRange fr,
                                                                       const uint32_t *valid32, const uint32_t *in32,
                                                                       uint32_t *bitmap32, uint64_t out_row0,
                                                                       Agg *agg, uint32_t window_bytes) {
	using T = typename UIntOf<S>::type;
	extern __shared__ __align__(128) uint8_t f_mem[];
	__shared__ uint64_t red[96];
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	uint32_t *window = (uint32_t *)(f_mem + (size_t)warp * window_bytes);
	Acc96 acc;
	uint32_t n_pass = 0;
	// per-segment state (warp-uniform), recomputed when the tile belongs to another segment
	uint32_t c_seg = ~0u, c_w = 0, c_count = 0, c_flo = 0, c_span = 0;
	uint64_t c_flo64 = 0, c_span64 = 0;  // the same range for payloads wider than 32 bits
	uint64_t c_row_start = 0, c_packed = 0, c_min_add = 0, c_emin = 0;
	int c_kind = 0;  // 0 nothing passes, 1 field range, 2 complement of one field value (<>), 3 per-value path
	bool c_repacked = false;

	const uint32_t total_warps = gridDim.x * kFWarps, wg = blockIdx.x * kFWarps + warp;
	for (uint32_t run = wg; run * kFRun < n_tiles; run += total_warps) {
		for (uint32_t t = run * kFRun; t < min(n_tiles, (run + 1u) * kFRun); t++) {
			const Job job = tiles[t];
			if (job.seg != c_seg) {
				const DescQ dq = load_desc(desc + job.seg);
				c_seg = job.seg;
				c_count = dq.count();
				c_row_start = dq.row_start();
				c_repacked = dq.state() & kSegRepacked;
				c_w = c_repacked ? dq.width() : 8u * S;
				c_packed = dq.packed_ptr();
				const uint64_t mn = dq.min_factor();
				c_min_add = (mn != ~0ull && (c_repacked || literal)) ? mn : 0ull;
				const uint64_t mask64 = lo_set64(c_w);
				c_emin = extend_low<S, SIGNED>(c_min_add);
				const uint64_t top = c_min_add + mask64;
				const uint64_t e_top = extend_low<S, SIGNED>(top);
				// (payloads wider than 32 bits: the same field-domain test with 64-bit fields, lane = row from global memory)
				const bool linear = c_repacked && top >= c_min_add && e_top - c_emin == mask64 &&
				                    (SIGNED ? (int64_t)c_emin <= (int64_t)e_top : c_emin <= e_top);
				c_kind = 3;
				if (linear) {
					// the value range [lo, hi] (or its complement) clipped to the segment's interval, in the field domain
					const FieldRange r = clip_range_to_segment<SIGNED>(fr, c_emin, e_top, mask64);
					c_flo64 = r.flo;
					c_span64 = r.span;
					c_kind = r.kind;
					c_flo = (uint32_t)c_flo64;
					c_span = (uint32_t)c_span64;
				}
			}
			const uint32_t first = job.k * kFRows;
			if (first >= c_count || c_kind == 0) {
				continue;  // warp-uniform; the bitmap was zeroed by the caller
			}
			const uint32_t cnt = min(kFRows, c_count - first);
			const uint64_t row0 = c_row_start + first;
			uint32_t sel[2] = {0u, 0u};
			// optional masks: validity (bit = column row) and the incoming selection (bit = row - out_row0).  allow_word:
			// the 32 bits of rows [32 g, 32 g + nrows) of the tile (every lane may ask for a different group)
			constexpr bool has_allow = ALLOW;  // (a kernel of its own: the plain filter keeps its registers)
			auto mask_bits = [&](const uint32_t *m, uint64_t r, uint32_t nrows) -> uint32_t {
				const uint32_t sh = (uint32_t)(r & 31u);
				const uint32_t *p = m + (r >> 5);
				return __funnelshift_r(__ldg(p), sh + nrows > 32u ? __ldg(p + 1) : 0u, sh);
			};
			auto allow_word = [&](uint32_t g, uint32_t nrows) -> uint32_t {
				uint32_t aw = 0xffffffffu;
				if (valid32 != nullptr) {
					aw &= mask_bits(valid32, row0 + 32u * g, nrows);
				}
				if (in32 != nullptr) {
					aw &= mask_bits(in32, row0 - out_row0 + 32u * g, nrows);
				}
				return aw;
			};
			if (S == 8 && c_kind != 3 && c_w > 32u) {
				// wide payloads (33 .. 64 bits) in the field domain: the tile is staged in two halves of 1024 rows (128 w
				// bytes each: 16-byte aligned, and the warp's window stays at 8 KiB so three CTAs fit an SM), lane = row,
				// (f - flo) <= span on 64 bits, the passing fields summed on 96 bits; SUM = sum(f) + n * extend(min).
				// (reading the fields straight from global memory, lane = row, was measured at 0.66 ms per 2^27 values
				// against 0.39 ms for decode_kernel<filter>: without the staging the warps wait on every group)
				const uint32_t w = c_w;
				const uint64_t mask64 = lo_set64(w);
				uint64_t fs_lo = 0;
				uint32_t fs_hi = 0, n = 0;
#pragma unroll 1
				for (uint32_t half = 0; half * 1024u < cnt; half++) {
					const uint32_t hcnt = min(1024u, cnt - half * 1024u);
					const uint8_t *src = (const uint8_t *)(c_packed + ((uint64_t)(first + half * 1024u) / 64u) * w * 8u);
					const uint32_t nbytes = ((hcnt * w + 127u) >> 7) * 16u;
					__syncwarp();  // the previous half has been consumed
					for (uint32_t off = lane * 16u; off < nbytes; off += 512u) {
						f_cp_async16((uint8_t *)window + off, src + off);
					}
					asm volatile("cp.async.wait_all;" ::: "memory");
					if (lane < 4) {
						window[(nbytes >> 2) + lane] = 0u;  // the funnel's words behind the last field
					}
					__syncwarp();
					const uint64_t *words = (const uint64_t *)window;
					const uint32_t n_g = (hcnt + 31u) >> 5;
					uint32_t mine = 0;
#pragma unroll 4
					for (uint32_t gg = 0; gg < n_g; gg++) {
						const uint32_t i = gg * 32u + lane;
						const bool inside = i < hcnt;
						const uint64_t f = read_bits64(words, (inside ? i : 0u) * w) & mask64;
						bool pass = inside && (((f - c_flo64) <= c_span64) != (c_kind == 2));
						if (has_allow) {
							pass = pass && ((allow_word(half * 32u + gg, min(32u, hcnt - gg * 32u)) >> lane) & 1u);
						}
						const uint32_t word = __ballot_sync(0xffffffffu, pass);
						const uint64_t add = pass ? f : 0ull;
						fs_lo += add;
						fs_hi += fs_lo < add ? 1u : 0u;
						mine = lane == gg ? word : mine;
					}
					n += __popc(mine);
					sel[0] = half == 0 ? mine : sel[0];
					sel[1] = half == 1 ? mine : sel[1];
				}
				// every lane holds the popcount of ITS groups; the per-lane partial sums and counts are reduced at the end
				const uint64_t plo = c_emin * (uint64_t)n;
				uint64_t phi = __umul64hi(c_emin, (uint64_t)n);
				if (SIGNED && (int64_t)c_emin < 0) {
					phi -= (uint64_t)n;
				}
				uint64_t lo = plo, hi = phi;
				add128(lo, hi, fs_lo, (uint64_t)fs_hi);
				acc.add96(lo, (uint32_t)hi);
				n_pass += n;
			} else if (c_kind != 3) {
				const uint32_t w = c_w;
				const uint8_t *src = (const uint8_t *)(c_packed + (uint64_t)(first / 64u) * w * 8u);
				const uint32_t nbytes = ((cnt * w + 127u) >> 7) * 16u;
				__syncwarp();  // the previous tile has been consumed
				for (uint32_t off = lane * 16u; off < nbytes; off += 512u) {
					f_cp_async16((uint8_t *)window + off, src + off);
				}
				asm volatile("cp.async.wait_all;" ::: "memory");
				if (lane == 0) {
					window[nbytes >> 2] = 0u;  // the funnel's word behind the last field
				}
				__syncwarp();
				uint32_t full_groups = cnt >> 5;
				uint64_t fsum = 0;
				uint32_t n = 0;
				if (w == 32u) {
					// whole-word fields: lane = row is as cheap as pred32 and free of the 32-way bank conflict a lane per
					// group would have at this width
					for (uint32_t gg = 0; gg < full_groups; gg++) {
						const uint32_t f = window[gg * 32u + lane];
						bool pass = ((f - c_flo) <= c_span) != (c_kind == 2);
						if (has_allow) {
							pass = pass && ((allow_word(gg, 32u) >> lane) & 1u);
						}
						const uint32_t word = __ballot_sync(0xffffffffu, pass);
						fsum += pass ? (uint64_t)f : 0ull;
						n += pass ? 1u : 0u;
						if (lane == (gg & 31u)) {
							sel[gg >> 5] = word;
						}
					}
					full_groups = 0;  // (the loop below has nothing left; the ragged tail uses cnt)
				}
#pragma unroll 1
				for (int j = 0; j < 2; j++) {
					const uint32_t g = lane + 32u * j;
					if (g < full_groups) {
						uint32_t word;
						if (c_kind == 1) {
							word = pred32_sum_any(w, window + g * w, 0u, c_flo, c_span, fsum);
						} else {
							// outside the field range: everything but the fields inside it
							uint64_t in_sum = 0, all = 0;
							const uint32_t inr = pred32_sum_any(w, window + g * w, 0u, c_flo, c_span, in_sum);
							pred32_sum_any(w, window + g * w, 0u, 0u, 0xffffffffu, all);
							word = ~inr;
							fsum += all - in_sum;
						}
						if (has_allow) {
							// rows the masks take out again: their fields leave the sum one by one (NULLs are few)
							const uint32_t aw = allow_word(g, 32u);
							uint32_t rem = word & ~aw;
							word &= aw;
							const uint32_t mask32 = w >= 32u ? 0xffffffffu : ((1u << w) - 1u);
							while (rem) {
								const uint32_t k = (uint32_t)__ffs((int)rem) - 1u;
								rem &= rem - 1u;
								fsum -= (uint64_t)(read_bits32(window, (g * 32u + k) * w) & mask32);
							}
						}
						n += __popc(word);
						sel[0] = j == 0 ? word : sel[0];
						sel[1] = j == 1 ? word : sel[1];
					}
				}
				if (cnt & 31u) {  // ragged tail group: lane = row
					full_groups = cnt >> 5;
					const uint32_t i = full_groups * 32u + lane;
					const bool inside = i < cnt;
					const uint32_t mask32 = w >= 32u ? 0xffffffffu : ((1u << w) - 1u);
					const uint32_t f = read_bits32(window, (inside ? i : 0u) * w) & mask32;
					bool pass = inside && (((f - c_flo) <= c_span) != (c_kind == 2));
					if (has_allow) {
						pass = pass && ((allow_word(full_groups, cnt & 31u) >> lane) & 1u);
					}
					const uint32_t word = __ballot_sync(0xffffffffu, pass);
					fsum += pass ? (uint64_t)f : 0ull;
					n += pass ? 1u : 0u;
					if (lane == (full_groups & 31u)) {
						sel[full_groups >> 5] = word;
					}
				}
				// + fsum + n * extend(min) as a 96-bit two's-complement number
				const uint64_t plo = c_emin * (uint64_t)n;
				uint64_t phi = __umul64hi(c_emin, (uint64_t)n);
				if (SIGNED && (int64_t)c_emin < 0) {
					phi -= (uint64_t)n;
				}
				uint64_t lo = plo, hi = phi;
				add128(lo, hi, fsum, 0ull);
				acc.add96(lo, (uint32_t)hi);
				n_pass += n;
			} else {
				// per-value path: raw slots, frames that wrap inside T, payloads wider than 32 bits
				const uint32_t w = c_w;
				const uint64_t *words = (const uint64_t *)c_packed;
				const T *slots = (const T *)raw + row0;
				const uint64_t mask64 = lo_set64(w);
				for (uint32_t gg = 0; gg * 32u < cnt; gg++) {
					const uint32_t i = gg * 32u + lane;
					bool pass = false;
					uint64_t v = 0;
					if (i < cnt) {
						uint64_t f;
						if (c_repacked) {
							const uint64_t bit = (uint64_t)(first + i) * w;
							const uint32_t off = (uint32_t)(bit & 63u);
							f = __ldg(words + (bit >> 6)) >> off;
							if (off + w > 64u) {
								f |= __ldg(words + (bit >> 6) + 1) << (64u - off);
							}
							f &= mask64;
						} else {
							f = (uint64_t)__ldg(slots + i);
						}
						v = extend_low<S, SIGNED>(f + c_min_add);
						const bool inr = !fr.empty && !t_less<SIGNED>(v, fr.lo) && !t_less<SIGNED>(fr.hi, v);
						pass = inr != (fr.complement != 0u);
						if (has_allow) {
							pass = pass && ((allow_word(gg, min(32u, cnt - gg * 32u)) >> lane) & 1u);
						}
					}
					if (pass) {
						acc.add<SIGNED>(v);
						n_pass++;
					}
					const uint32_t word = __ballot_sync(0xffffffffu, pass);
					if (lane == (gg & 31u)) {
						sel[gg >> 5] = word;
					}
				}
			}
			if (bitmap32 != nullptr) {
				// selection words S[g] (g = lane + 32 j) -> output words O[t] = S[t] << sh | S[t - 1] >> (32 - sh); the first
				// and the last word of the tile are shared with its neighbours (atomics), the others are whole stores
				const uint64_t rel0 = row0 - out_row0;
				const uint32_t sh = (uint32_t)(rel0 & 31u);
				uint32_t *dst = bitmap32 + (rel0 >> 5);
				const uint32_t n_words = (sh + cnt + 31u) >> 5;
				uint32_t prev0 = __shfl_up_sync(0xffffffffu, sel[0], 1);
				uint32_t prev1 = __shfl_up_sync(0xffffffffu, sel[1], 1);
				const uint32_t last0 = __shfl_sync(0xffffffffu, sel[0], 31);
				const uint32_t last1 = __shfl_sync(0xffffffffu, sel[1], 31);
				if (lane == 0) {
					prev0 = 0u;
					prev1 = last0;
				}
#pragma unroll
				for (int j = 0; j < 3; j++) {
					const uint32_t tw = lane + 32u * j;  // output word
					if (j == 2 && lane != 0) {
						break;
					}
					const uint32_t cur = j == 0 ? sel[0] : (j == 1 ? sel[1] : 0u);
					const uint32_t prv = j == 0 ? prev0 : (j == 1 ? prev1 : last1);
					const uint32_t o = sh ? ((cur << sh) | (prv >> (32u - sh))) : cur;
					if (tw < n_words && o) {
						if (tw == 0 || tw == n_words - 1u) {
							atomicOr(dst + tw, o);
						} else {
							dst[tw] = o;
						}
					}
				}
			}
		}
	}
	block_reduce_agg(acc.lo64(), acc.hi64(), (uint64_t)n_pass, red, agg);
}

template <int S>
cudaError_t launch_w(bool is_signed, const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw, int literal,
                     const FilterRange &range, const uint32_t *valid32, const uint32_t *in32, uint32_t *bitmap32, uint64_t out_row0,
                     Agg *agg, uint32_t wmax,
                     int sm_count, cudaStream_t stream) {
	// a whole 2048-row tile of a payload <= 32 bits (256 w bytes), or half a tile of a wider one (128 w bytes)
	const uint32_t narrow = 256u * (wmax < 32u ? wmax : 32u), wide = wmax > 32u ? 128u * (wmax < 64u ? wmax : 64u) : 0u;
	const uint32_t window = ((narrow > wide ? narrow : wide) + 16u + 127u) & ~127u;
	const size_t smem = (size_t)kFWarps * window;
	void (*kernel)(const SegDesc *, const Job *, uint32_t, const uint8_t *, int, const FilterRange, const uint32_t *, const uint32_t *,
	               uint32_t *, uint64_t, Agg *,
	               uint32_t);
	if (valid32 != nullptr || in32 != nullptr) {
		kernel = is_signed ? filter_words_kernel<S, true, true> : filter_words_kernel<S, false, true>;
	} else {
		kernel = is_signed ? filter_words_kernel<S, true, false> : filter_words_kernel<S, false, false>;
	}
	cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	if (e != cudaSuccess) {
		return e;
	}
	int per_sm = 0;
	e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kFWarps * 32, smem);
	if (e != cudaSuccess) {
		return e;
	}
	uint64_t grid = (uint64_t)sm_count * (uint64_t)(per_sm > 0 ? per_sm : 1);
	const uint64_t need = ((n_tiles + kFRun - 1) / kFRun + kFWarps - 1) / kFWarps;
	if (grid > need) {
		grid = need;
	}
	kernel<<<(uint32_t)grid, kFWarps * 32, smem, stream>>>(desc, tiles, n_tiles, raw, literal, range, valid32, in32, bitmap32, out_row0,
	                                                       agg, window);
	return cudaGetLastError();
}

template <int S>
cudaError_t launch_s(bool is_signed, const SegDesc *desc, uint32_t seg_b, uint32_t n_segs, uint32_t cps, const uint8_t *raw,
                     int literal, uint32_t op_mask, uint64_t constant, uint32_t *bitmap32, uint64_t out_row0, Agg *agg,
                     int sm_count, cudaStream_t stream) {
	uint32_t cps_shift = 0;
	while ((1u << cps_shift) < cps) {
		cps_shift++;
	}
	const uint64_t jobs = (uint64_t)n_segs << cps_shift;
	if (jobs >= (1ull << 32)) {
		return cudaErrorInvalidValue;
	}
	const uint32_t resident = (uint32_t)sm_count * 8u;
	const uint32_t grid = (uint32_t)(jobs < resident ? jobs : resident);
	if (is_signed) {
		filter_fields_kernel<S, true><<<grid, kFiltThreads, 0, stream>>>(desc, seg_b, n_segs, cps_shift, raw, literal, op_mask,
		                                                                 constant, bitmap32, out_row0, agg);
	} else {
		filter_fields_kernel<S, false><<<grid, kFiltThreads, 0, stream>>>(desc, seg_b, n_segs, cps_shift, raw, literal, op_mask,
		                                                                  constant, bitmap32, out_row0, agg);
	}
	return cudaGetLastError();
}

}  // namespace

cudaError_t launch_filter_words(TypeInfo t, const SegDesc *desc, const Job *tiles, uint32_t n_tiles, const uint8_t *raw,
                                int literal, const FilterRange &range, const uint32_t *valid32, const uint32_t *in32,
                                uint32_t *bitmap32, uint64_t out_row0, Agg *agg, uint32_t wmax, int sm_count,
                                cudaStream_t stream) {
	if (n_tiles == 0) {
		return cudaSuccess;
	}
	switch (t.size) {
	case 1: return launch_w<1>(t.is_signed, desc, tiles, n_tiles, raw, literal, range, valid32, in32, bitmap32, out_row0, agg, wmax, sm_count, stream);
	case 2: return launch_w<2>(t.is_signed, desc, tiles, n_tiles, raw, literal, range, valid32, in32, bitmap32, out_row0, agg, wmax, sm_count, stream);
	case 4: return launch_w<4>(t.is_signed, desc, tiles, n_tiles, raw, literal, range, valid32, in32, bitmap32, out_row0, agg, wmax, sm_count, stream);
	case 8: return launch_w<8>(t.is_signed, desc, tiles, n_tiles, raw, literal, range, valid32, in32, bitmap32, out_row0, agg, wmax, sm_count, stream);
	default: return cudaErrorInvalidValue;
	}
}

cudaError_t launch_filter_fields(TypeInfo t, const SegDesc *desc, uint32_t seg_b, uint32_t n_segs, uint32_t seg_capacity,
                                 const uint8_t *raw, int literal, uint32_t op_mask, uint64_t constant, uint32_t *bitmap32,
                                 uint64_t out_row0, Agg *agg, int sm_count, cudaStream_t stream) {
	if (n_segs == 0) {
		return cudaSuccess;
	}
	const uint32_t cps = (seg_capacity + kFiltChunk - 1) / kFiltChunk;
	switch (t.size) {
	case 1: return launch_s<1>(t.is_signed, desc, seg_b, n_segs, cps, raw, literal, op_mask, constant, bitmap32, out_row0, agg, sm_count, stream);
	case 2: return launch_s<2>(t.is_signed, desc, seg_b, n_segs, cps, raw, literal, op_mask, constant, bitmap32, out_row0, agg, sm_count, stream);
	case 4: return launch_s<4>(t.is_signed, desc, seg_b, n_segs, cps, raw, literal, op_mask, constant, bitmap32, out_row0, agg, sm_count, stream);
	case 8: return launch_s<8>(t.is_signed, desc, seg_b, n_segs, cps, raw, literal, op_mask, constant, bitmap32, out_row0, agg, sm_count, stream);
	default: return cudaErrorInvalidValue;
	}
}

}  // namespace adacom
