// kernels_query.cu -- fused multi-column scan: conjunctive range filters + SUM of affine products, one pass.
//
// What DuckDB does above the codec for a query like TPC-H Q6 (extension/tpch/dbgen/queries/q06.sql) is, per
// 2048-row vector of a row group: ColumnData::Scan every referenced column (SuccinctScan / SuccinctScanPartial,
// src/storage/compression/succinct.cpp:123-144,232-240 -- a vector may straddle two segments because segment
// capacities are 262136/S values, not multiples of 2048), refine a selection vector filter by filter
// (ColumnSegment::FilterSelection, src/storage/table/column_segment.cpp:575-658; conjunctions :685-844), then
// evaluate the aggregate expression on the surviving rows (callers: RowGroup::TemplatedScan,
// src/storage/table/row_group.cpp:391-447).  Here the whole pipeline is one kernel over row blocks:
//
//   * a row block = R rows (2048 | 4096 | 8192, aligned to 2048) of the TABLE; for every column the block is
//     covered by <= 3 "pieces" (segment, first value, count), found through a static vector->segment table
//     and the live descriptors (widths / frames of reference differ per piece);
//   * producer warp: lane = (column, piece); resolves the pieces of the NEXT block while the current one is in
//     flight, lays the pieces out back to back in one stage and issues one cp.async.bulk (TMA) per piece, all
//     completing on the stage's mbarrier;
//   * 4 consumer warps, column at a time like DuckDB's vectorised filters: a warp owns a contiguous quarter of the
//     block's 32-row groups and keeps their selection words in shared memory.  Predicate columns are evaluated in
//     the FIELD domain (the producer turns lo <= min + f <= hi into f - flo <= span per piece); a LANE owns a whole
//     32-row group = w whole words of the piece, realigned once, so every field sits at a compile-time position
//     (pred32<W>: shift, mask, subtract, compare, select per row; no address arithmetic, no ballot).  Surviving rows are compacted into a per-warp list and only those rows decode
//     the factor columns and fold prod_k(add_k + mul_k * value_k) into 128-bit per-thread accumulators.
//     Optional selection bitmap in (AND) / out.  No barrier between the consumer warps.
//
// Algorithmic HBM bytes per row: sum over the columns of w_c / 8 (+ 1/8 per bitmap).
#include "adacom_kernels.cuh"
#include "adacom_pred32.cuh"

namespace adacom {

namespace {

constexpr int kQConsumerWarps = kQueryConsumerWarps;  // 4: a warp owns 32 consecutive 32-row groups of a 4096-row block
constexpr int kQConsumers = kQConsumerWarps * 32;
constexpr int kQThreads = kQConsumers + 32;
constexpr int kQMaxStages = 3;
constexpr int kQLanesPerCol = 4;  // producer lanes per column: piece 0..2, lane 3 idle
constexpr uint32_t kQMaxBlockRows = 8192;
constexpr uint32_t kQListCap = 256;

struct QPiece {
	uint64_t min_add;   // frame of reference to add (0 when none)
	uint64_t flo;       // predicate in the field domain: passes iff (f - flo) <= span (unsigned); linear pieces only
	uint64_t span;
	uint32_t smem_off;  // byte offset of the staged range inside the stage (16-byte aligned)
	uint32_t bit_base;  // bit position of the piece's first field inside the staged range (0..127)
	uint32_t cnt;       // values of this piece inside the block (0: piece unused)
	uint32_t w;
	uint32_t linear;    // value == extend(min) + f without wrap-around for every f < 2^w
	uint32_t pad;
	uint64_t vbase;     // extend(min) of a linear piece: value = vbase + f
};

struct QStage {
	uint64_t row0;      // first table row of the block
	uint32_t rows;      // rows in the block
	uint32_t done;      // sentinel
	QPiece piece[kQueryMaxCols][kQueryMaxPieces];
};

__device__ __forceinline__ uint64_t extend_rt(uint64_t x, uint32_t type_size, uint32_t is_signed) {
	const uint32_t shift = 64u - 8u * type_size;
	return is_signed ? (uint64_t)(((int64_t)(x << shift)) >> shift) : ((x << shift) >> shift);
}

// value of block row i of one column, from the staged pieces
__device__ __forceinline__ uint64_t column_value(const QPiece *pc, const uint8_t *stage, uint32_t i, uint32_t type_size,
                                                 uint32_t is_signed) {
	uint32_t p = 0, off = i;
	const uint32_t c0 = pc[0].cnt;
	if (off >= c0) {
		off -= c0;
		p = 1;
		const uint32_t c1 = pc[1].cnt;
		if (off >= c1) {
			off -= c1;
			p = 2;
		}
	}
	const QPiece &q = pc[p];
	const uint64_t *words = (const uint64_t *)(stage + q.smem_off);
	const uint32_t bit = q.bit_base + off * q.w;
	if (q.linear && q.w <= 32u) {
		return q.vbase + (uint64_t)(read_bits32((const uint32_t *)words, bit) & (0xffffffffu >> (32u - q.w)));
	}
	const uint64_t mask = lo_set64(q.w);
	const uint64_t f = q.w <= 32u ? (uint64_t)(read_bits32((const uint32_t *)words, bit) & (uint32_t)mask)
	                              : (read_bits64(words, bit) & mask);
	return extend_rt(f + q.min_add, type_size, is_signed);
}

// the same for a row of the 32-row group starting at block row g32, when the whole group lies inside one piece (the
// piece search and the piece record are then warp-uniform); groups that straddle a piece boundary take column_value
__device__ __forceinline__ uint64_t group_value(const QPiece *pc, const uint8_t *stage, uint32_t g32, uint32_t i,
                                                uint32_t type_size, uint32_t is_signed) {
	uint32_t p = 0, start = 0;
	const uint32_t c0 = pc[0].cnt;
	if (g32 >= c0) {
		p = 1;
		start = c0;
		const uint32_t c1 = pc[1].cnt;
		if (g32 - c0 >= c1) {
			p = 2;
			start = c0 + c1;
		}
	}
	if (g32 + 31u - start >= pc[p].cnt) {
		return column_value(pc, stage, i, type_size, is_signed);
	}
	const uint32_t w = pc[p].w;
	const uint32_t bit = pc[p].bit_base + (i - start) * w;
	const uint8_t *base = stage + pc[p].smem_off;
	if (pc[p].linear && w <= 32u) {
		// the common case: a narrow field on a frame of reference that cannot wrap
		return pc[p].vbase + (uint64_t)(read_bits32((const uint32_t *)base, bit) & (0xffffffffu >> (32u - w)));
	}
	const uint64_t mask = lo_set64(w);
	const uint64_t f = w <= 32u ? (uint64_t)(read_bits32((const uint32_t *)base, bit) & (uint32_t)mask)
	                            : (read_bits64((const uint64_t *)base, bit) & mask);
	return extend_rt(f + pc[p].min_add, type_size, is_signed);
}

// predicate of column qc on block row i: field-domain compare on linear narrow pieces, decoded compare otherwise
__device__ __forceinline__ bool row_passes(const QPiece *pc, const uint8_t *stage, uint32_t i, const QueryCol &qc) {
	uint32_t p = 0, off = i;
	const uint32_t c0 = pc[0].cnt;
	if (off >= c0) {
		off -= c0;
		p = 1;
		const uint32_t c1 = pc[1].cnt;
		if (off >= c1) {
			off -= c1;
			p = 2;
		}
	}
	const QPiece &q = pc[p];
	const uint32_t w = q.w;
	const uint32_t bit = q.bit_base + off * w;
	if (q.linear && w <= 32u && (q.flo >> 32) == 0) {
		const uint32_t f = read_bits32((const uint32_t *)(stage + q.smem_off), bit) & (0xffffffffu >> (32u - w));
		return (f - (uint32_t)q.flo) <= (uint32_t)q.span;
	}
	const uint64_t v = column_value(pc, stage, i, qc.type_size, qc.is_signed);
	return qc.is_signed ? ((int64_t)v >= qc.lo && (int64_t)v <= qc.hi) : (v >= (uint64_t)qc.lo && v <= (uint64_t)qc.hi);
}

struct PieceRes {
	uint64_t src16;
	uint64_t min_add;
	uint64_t flo, span;
	uint32_t bytes;  // 16-byte rounded staged range (0: no piece)
	uint32_t bit_base;
	uint32_t cnt;
	uint32_t w;
	uint32_t linear;
};

// [lo, hi] on T values -> [flo, flo + span] on fields for a piece whose values are extend(min) + f, f < 2^w
// (no wrap: the piece's whole interval lies inside one sign / truncation zone of T).  An empty intersection
// becomes flo = 2^64 - 1 (no field of width < 64 equals it).
__device__ __forceinline__ void field_range(const QueryCol &c, uint64_t min_add, uint32_t w, uint64_t &flo, uint64_t &span) {
	const __int128 e = c.is_signed ? (__int128)(int64_t)extend_rt(min_add, c.type_size, 1)
	                               : (__int128)extend_rt(min_add, c.type_size, 0);
	const __int128 lo = c.is_signed ? (__int128)c.lo : (__int128)(uint64_t)c.lo;
	const __int128 hi = c.is_signed ? (__int128)c.hi : (__int128)(uint64_t)c.hi;
	const __int128 fmax = (__int128)lo_set64(w);
	__int128 a = lo - e, b = hi - e;
	if (a < 0) {
		a = 0;
	}
	if (b > fmax) {
		b = fmax;
	}
	if (a > b) {
		flo = ~0ull;
		span = 0;
	} else {
		flo = (uint64_t)a;
		span = (uint64_t)(b - a);
	}
}

__device__ __forceinline__ bool piece_is_linear(const QueryCol &c, uint64_t min_add, uint32_t w, bool repacked) {
	if (w >= 64) {
		return false;
	}
	if (!repacked && (c.is_signed || min_add != 0)) {
		return false;  // raw signed slots: pattern order != value order across the sign
	}
	const uint64_t range = (1ull << w) - 1;
	const uint64_t top = min_add + range;
	if (top < min_add) {
		return false;
	}
	const uint64_t e_lo = extend_rt(min_add, c.type_size, c.is_signed), e_hi = extend_rt(top, c.type_size, c.is_signed);
	if (e_hi - e_lo != range) {
		return false;
	}
	return c.is_signed ? (int64_t)e_lo <= (int64_t)e_hi : e_lo <= e_hi;
}

// piece `dq` (a segment of column c) restricted to the block rows [r0, r0 + rows)
__device__ __forceinline__ PieceRes resolve_piece(const QueryCol &c, const DescQ &dq, bool valid, uint64_t r0, uint32_t rows,
                                                  int literal) {
	PieceRes r;
	r.bytes = 0;
	r.cnt = 0;
	r.src16 = 0;
	r.min_add = 0;
	r.bit_base = 0;
	r.w = 8u * c.type_size;
	r.flo = ~0ull;
	r.span = 0;
	r.linear = 0;
	if (!valid) {
		return r;
	}
	const uint64_t rs = dq.row_start();
	const uint64_t re = rs + dq.count();
	const uint64_t lo = rs > r0 ? rs : r0;
	const uint64_t hi = re < r0 + rows ? re : r0 + rows;
	if (hi <= lo) {
		return r;
	}
	const uint32_t first = (uint32_t)(lo - rs);
	const uint32_t cnt = (uint32_t)(hi - lo);
	const uint32_t w = dq.width();
	const bool repacked = dq.state() & kSegRepacked;
	const uint64_t mn = dq.min_factor();
	const uint64_t bit = (uint64_t)first * w;
	const uint64_t src = repacked ? dq.packed_ptr() + (bit >> 3) : (uint64_t)c.raw + (rs + first) * (uint64_t)c.type_size;
	r.src16 = src & ~15ull;
	const uint32_t byte_base = (uint32_t)(src - r.src16);
	r.bit_base = byte_base * 8u + (repacked ? (uint32_t)(bit & 7u) : 0u);
	r.bytes = (byte_base + (((uint32_t)(bit & 7u) + cnt * w + 7u) >> 3) + 15u) & ~15u;
	r.cnt = cnt;
	r.w = w;
	r.min_add = (mn != ~0ull && (repacked || literal)) ? mn : 0ull;
	if (piece_is_linear(c, r.min_add, w, repacked)) {
		r.linear = 1;
		if (c.has_pred) {
			field_range(c, r.min_add, w, r.flo, r.span);
		}
	}
	return r;
}

// Aggregates over the rows a warp has collected: lane e takes list entries e, e + 32, ...; it decodes the
// factor columns of its row (generic per-value path) and folds every sum expression into its accumulators.
__device__ __forceinline__ void aggregate_list(const QueryArgs &a, const QStage &st, const uint8_t *stage,
                                               const uint16_t *my_list, uint32_t n_list,
                                               uint64_t (*vals)[kQConsumers], uint32_t tid, uint32_t lane,
                                               unsigned __int128 *acc) {
	for (uint32_t e = lane; e < n_list; e += 32u) {
		const uint32_t i = my_list[e];
#pragma unroll 1
		for (uint32_t c = 0; c < a.n_cols; c++) {
			if ((a.factor_mask >> c) & 1u) {
				const QueryCol &qc = a.cols[c];
				vals[c][tid] = column_value(st.piece[c], stage, i, qc.type_size, qc.is_signed);
			}
		}
		if (a.narrow_products) {
			// the planner bounded every product below 2^63 from the segments' min / max: plain 64-bit multiplies
			// (a 128 x 128-bit product costs ~40 instructions, Q1 has 16 of them per row)
#pragma unroll
			for (int g = 0; g < kQueryMaxAggs; g++) {
				if (g < (int)a.n_aggs) {
					const QueryAgg &qa = a.aggs[g];
					int64_t term = 1;
#pragma unroll
					for (int k = 0; k < kQueryMaxFactors; k++) {
						if (k < (int)qa.n_factors) {
							term *= qa.add[k] + qa.mul[k] * (int64_t)vals[qa.col[k]][tid];
						}
					}
					acc[g] += (unsigned __int128)(__int128)term;
				}
			}
			continue;
		}
#pragma unroll
		for (int g = 0; g < kQueryMaxAggs; g++) {
			if (g < (int)a.n_aggs) {
				const QueryAgg &qa = a.aggs[g];
				__int128 term = 1;
#pragma unroll
				for (int k = 0; k < kQueryMaxFactors; k++) {
					if (k < (int)qa.n_factors) {
						const int64_t v = (int64_t)vals[qa.col[k]][tid];
						term *= (__int128)qa.add[k] + (__int128)qa.mul[k] * (__int128)v;
					}
				}
				acc[g] += (unsigned __int128)term;
			}
		}
	}
}

// GROUP BY a small-domain column: the warp's survivor list is split by group key into dense sub-lists (ballot
// compaction), every sub-list is folded with the per-thread register accumulators exactly like the ungrouped
// case, and the warp's partial sums are then reduced and added to its private row of shared accumulators -- no
// atomics, no divergent per-row group dispatch.
struct GroupAcc {
	unsigned long long sum[kQConsumerWarps][kQueryMaxGroups][kQueryMaxAggs][2];
	unsigned long long cnt[kQConsumerWarps][kQueryMaxGroups];
};

struct GroupScratch {
	GroupAcc acc;
	uint16_t list[kQConsumerWarps][kQListCap];  // the warp's surviving rows, sorted by group
	uint8_t keys[kQConsumerWarps][kQListCap];
};

__device__ __forceinline__ void aggregate_grouped(const QueryArgs &a, const QStage &st, const uint8_t *stage,
                                                  const uint16_t *my_list, uint32_t n_list, uint16_t *sub_list,
                                                  uint8_t *keys, GroupAcc &ga, uint64_t (*vals)[kQConsumers],
                                                  uint32_t tid, uint32_t lane, uint32_t warp, unsigned __int128 *acc) {
	if (n_list == 0) {
		return;
	}
	const QueryCol &gc = a.cols[a.group_col];
	uint32_t count[kQueryMaxGroups];
#pragma unroll
	for (int g = 0; g < kQueryMaxGroups; g++) {
		count[g] = 0;
	}
	for (uint32_t e0 = 0; e0 < n_list; e0 += 32u) {  // pass 1: keys and group sizes
		const uint32_t e = e0 + lane;
		uint32_t key = 0xffu;
		if (e < n_list) {
			const uint64_t v = column_value(st.piece[a.group_col], stage, my_list[e], gc.type_size, gc.is_signed);
			key = v < (uint64_t)a.n_groups ? (uint32_t)v : 0xffu;
			keys[e] = (uint8_t)key;
		}
#pragma unroll
		for (int g = 0; g < kQueryMaxGroups; g++) {
			if (g < (int)a.n_groups) {
				count[g] += __popc(__ballot_sync(0xffffffffu, key == (uint32_t)g));
			}
		}
	}
	uint32_t offset[kQueryMaxGroups];
	uint32_t run = 0;
#pragma unroll
	for (int g = 0; g < kQueryMaxGroups; g++) {
		offset[g] = run;
		run += count[g];
	}
	__syncwarp();
	for (uint32_t e0 = 0; e0 < n_list; e0 += 32u) {  // pass 2: scatter into the group's slice
		const uint32_t e = e0 + lane;
		const uint32_t key = e < n_list ? keys[e] : 0xffu;
		const uint16_t row = e < n_list ? my_list[e] : 0;
#pragma unroll
		for (int g = 0; g < kQueryMaxGroups; g++) {
			if (g < (int)a.n_groups) {
				const uint32_t m = __ballot_sync(0xffffffffu, key == (uint32_t)g);
				if (key == (uint32_t)g) {
					sub_list[offset[g] + __popc(m & ((1u << lane) - 1u))] = row;
				}
				offset[g] += __popc(m);
			}
		}
	}
	__syncwarp();
	uint32_t begin = 0;
#pragma unroll 1
	for (uint32_t g = 0; g < a.n_groups; g++) {
		const uint32_t n = count[g];
		if (n == 0) {
			continue;
		}
#pragma unroll
		for (int k = 0; k < kQueryMaxAggs; k++) {
			acc[k] = 0;
		}
		aggregate_list(a, st, stage, sub_list + begin, n, vals, tid, lane, acc);
		begin += n;
#pragma unroll
		for (int k = 0; k < kQueryMaxAggs; k++) {
			if (k < (int)a.n_aggs) {
				uint64_t lo = (uint64_t)acc[k], hi = (uint64_t)(acc[k] >> 64);
#pragma unroll
				for (int d = 16; d > 0; d >>= 1) {
					const uint64_t olo = __shfl_down_sync(0xffffffffu, lo, d);
					const uint64_t ohi = __shfl_down_sync(0xffffffffu, hi, d);
					add128(lo, hi, olo, ohi);
				}
				if (lane == 0) {
					uint64_t slo = ga.sum[warp][g][k][0], shi = ga.sum[warp][g][k][1];
					add128(slo, shi, lo, hi);
					ga.sum[warp][g][k][0] = slo;
					ga.sum[warp][g][k][1] = shi;
				}
			}
		}
		if (lane == 0) {
			ga.cnt[warp][g] += n;
		}
	}
	__syncwarp();
}

__global__ void __launch_bounds__(kQThreads, 4) scan_columns_kernel(const QueryArgs a) {
	extern __shared__ __align__(128) uint8_t q_stage_mem[];
	__shared__ __align__(8) uint64_t full_bar[kQMaxStages];
	__shared__ __align__(8) uint64_t empty_bar[kQMaxStages];
	__shared__ QStage stages[kQMaxStages];
	__shared__ uint64_t vals[kQueryMaxCols][kQConsumers];  // decoded factors of the row a thread is working on
	__shared__ uint32_t sel[kQMaxBlockRows / 32];           // selection word per 32-row group (owned by g % 8)
	__shared__ uint16_t list[kQConsumerWarps][kQListCap];   // surviving block rows (< 8192), per warp
	__shared__ uint64_t red[96];
	const uint32_t tid = threadIdx.x;
	const uint32_t lane = tid & 31u;
	const uint32_t warp = tid >> 5;
	const uint32_t n_stages = a.stages;
	const uint32_t R = a.block_rows;
	const uint64_t n_rows = a.row_end - a.row_begin;
	const uint32_t n_blocks = (uint32_t)((n_rows + R - 1) / R);

	if (tid == 0) {
		for (uint32_t s = 0; s < n_stages; s++) {
			mbar_init(&full_bar[s], 1);
			mbar_init(&empty_bar[s], kQConsumerWarps);
		}
		mbar_fence_init();
	}
	// GROUP BY scratch lives behind the stages in dynamic shared memory (only allocated for grouped launches)
	GroupScratch &gs = *(GroupScratch *)(q_stage_mem + (size_t)n_stages * a.stage_bytes);
	GroupAcc &gacc = gs.acc;
	// dense grouped launches: int64 accumulator per (group, aggregate, consumer thread) + a row counter per (group,
	// thread) behind the scratch: a row costs one shared read-modify-write per aggregate, no list, no reduction
	long long *dacc = (long long *)(q_stage_mem + (size_t)n_stages * a.stage_bytes + ((sizeof(GroupScratch) + 127u) & ~127u));
	uint32_t *dcnt = (uint32_t *)(dacc + (size_t)a.n_groups * a.n_aggs * kQConsumers);
	if (a.n_groups) {
		unsigned long long *z = (unsigned long long *)&gacc;
		for (uint32_t i = tid; i < sizeof(GroupAcc) / 8; i += kQThreads) {
			z[i] = 0ull;
		}
		if (a.dense_groups) {
			for (uint32_t i = tid; i < a.n_groups * a.n_aggs * kQConsumers; i += kQThreads) {
				dacc[i] = 0ll;
			}
			for (uint32_t i = tid; i < a.n_groups * kQConsumers; i += kQThreads) {
				dcnt[i] = 0u;
			}
		}
	}
	__syncthreads();

	unsigned __int128 acc[kQueryMaxAggs];
#pragma unroll
	for (int g = 0; g < kQueryMaxAggs; g++) {
		acc[g] = 0;
	}
	uint64_t n_pass = 0;

	if (warp == kQConsumerWarps) {
		// ---------------- producer warp: lane = (column, piece) ----------------
		const uint32_t ci = lane / kQLanesPerCol, p = lane % kQLanesPerCol;
		const bool active = ci < a.n_cols && p < (uint32_t)kQueryMaxPieces;
		const QueryCol &col = a.cols[active ? ci : 0];
		auto block_seg = [&](uint32_t b) -> uint32_t {  // segment holding the first row of block b
			return __ldg(col.vec_seg + ((a.row_begin + (uint64_t)b * R) >> 11));
		};
		uint32_t b = blockIdx.x;
		uint32_t seg_cur = 0, seg_next = 0;
		DescQ dq_next;
		bool valid_next = false;
		if (active && b < n_blocks) {
			seg_cur = block_seg(b);
			valid_next = seg_cur + p < col.n_segs;
			if (valid_next) {
				dq_next = load_desc(col.desc + seg_cur + p);
			}
			if (b + gridDim.x < n_blocks) {
				seg_next = block_seg(b + gridDim.x);
			}
		}
		uint32_t seq = 0;
		for (; b < n_blocks; b += gridDim.x) {
			const DescQ dq = dq_next;
			const bool valid = valid_next;
			// fetch ahead: descriptors of the next block, vector->segment entry of the one after
			if (active && b + gridDim.x < n_blocks) {
				valid_next = seg_next + p < col.n_segs;
				if (valid_next) {
					dq_next = load_desc(col.desc + seg_next + p);
				}
				if (b + 2u * gridDim.x < n_blocks) {
					seg_next = block_seg(b + 2u * gridDim.x);
				}
			}
			const uint64_t r0 = a.row_begin + (uint64_t)b * R;
			const uint32_t rows = (uint32_t)(a.row_end - r0 < R ? a.row_end - r0 : R);
			const PieceRes pr = resolve_piece(col, dq, active && valid, r0, rows, a.literal);
			// back-to-back layout of the staged ranges: exclusive scan of the byte counts over the lanes
			uint32_t incl = pr.bytes;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
				if ((int)lane >= d) {
					incl += v;
				}
			}
			const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
			const uint32_t s = seq % n_stages;
			if (lane == 0) {
				mbar_wait_backoff(&empty_bar[s], ((seq / n_stages) & 1u) ^ 1u, 200);
				stages[s].row0 = r0;
				stages[s].rows = rows;
				stages[s].done = 0;
			}
			__syncwarp();
			if (active) {
				QPiece q;
				q.min_add = pr.min_add;
				q.flo = pr.flo;
				q.span = pr.span;
				q.linear = pr.linear;
				q.pad = 0;
				q.vbase = extend_rt(pr.min_add, col.type_size, col.is_signed);
				q.smem_off = incl - pr.bytes;
				q.bit_base = pr.bit_base;
				q.cnt = pr.cnt;
				q.w = pr.w;
				stages[s].piece[ci][p] = q;
			}
			__syncwarp();
			if (lane == 0) {
				mbar_arrive_expect_tx(&full_bar[s], total);
			}
			__syncwarp();
			if (pr.bytes) {
				tma_load_1d(q_stage_mem + (size_t)s * a.stage_bytes + (incl - pr.bytes), (const void *)pr.src16, pr.bytes,
				            &full_bar[s]);
			}
			seq++;
		}
		if (lane == 0) {
			const uint32_t s = seq % n_stages;
			mbar_wait_backoff(&empty_bar[s], ((seq / n_stages) & 1u) ^ 1u, 200);
			stages[s].done = 1;
			mbar_arrive(&full_bar[s]);
		}
	} else {
		// ---------------- consumer warps: a warp owns the 32-row groups g with g % 8 == warp ----------------
		uint16_t *my_list = list[warp];
		for (uint32_t it = 0;; it++) {
			const uint32_t s = it % n_stages;
			mbar_wait_backoff(&full_bar[s], (it / n_stages) & 1u, 64);
			const QStage &st = stages[s];
			if (st.done) {
				break;
			}
			const uint8_t *stage = q_stage_mem + (size_t)s * a.stage_bytes;
			const uint32_t rows = st.rows;
			const uint32_t n_groups = (rows + 31u) >> 5;
			const uint64_t word0 = (st.row0 - a.row_begin) >> 5;  // bitmap word of the block's first group

			// ---- my groups: a contiguous range of 32-row groups (one per lane for blocks of <= 4096 rows) ----
			const uint32_t gpw = (n_groups + kQConsumerWarps - 1u) / kQConsumerWarps;
			const uint32_t g_lo = min(warp * gpw, n_groups), g_hi = min(g_lo + gpw, n_groups);
			// ---- selection words of my groups: rows that exist, AND the optional input bitmap ----
			for (uint32_t g = g_lo + lane; g < g_hi; g += 32u) {
				const uint32_t left = rows - g * 32u;
				uint32_t word = left >= 32u ? 0xffffffffu : ((1u << left) - 1u);
				if (a.in_bitmap != nullptr) {
					word &= __ldg(a.in_bitmap + word0 + g);
				}
				sel[g] = word;
			}
			__syncwarp();

			// ---- predicate columns, one after the other; a column is walked piece by piece ----
			bool list_mode = false;
			uint32_t n_list = 0;
#pragma unroll 1
			for (uint32_t c = 0; c < a.n_cols; c++) {
				const QueryCol &qc = a.cols[c];
				if (!qc.has_pred) {
					continue;
				}
				if (list_mode) {
					// the earlier predicates left few rows: this one is tested on the compacted survivors only
					// (a row costs one field read; a full round costs every row of the block)
					uint32_t kept = 0;
					for (uint32_t e0 = 0; e0 < n_list; e0 += 32u) {
						const uint32_t e = e0 + lane;
						bool keep = false;
						uint16_t row = 0;
						if (e < n_list) {
							row = my_list[e];
							keep = row_passes(st.piece[c], stage, row, qc);
						}
						const uint32_t m = __ballot_sync(0xffffffffu, keep);  // every read of this round is done
						if (keep) {
							my_list[kept + __popc(m & ((1u << lane) - 1u))] = row;
						}
						kept += __popc(m);
					}
					__syncwarp();
					n_list = kept;
					continue;
				}
				uint32_t off_p = 0;
#pragma unroll 1
				for (int p = 0; p < kQueryMaxPieces; p++) {
					if (st.piece[c][p].cnt == 0) {
						continue;
					}
					const QPiece q = st.piece[c][p];
					const uint32_t end_p = off_p + q.cnt;
					const uint8_t *pbase = stage + q.smem_off;
					const uint64_t mask64 = lo_set64(q.w);
					if (q.linear && q.flo == 0 && q.span == mask64) {
						off_p = end_p;  // zone map: every value of the piece passes
						continue;
					}
					const uint32_t *words32 = (const uint32_t *)pbase;
					if (q.linear && q.w <= 32u && (q.flo >> 32) == 0) {
						// fast path: 32-bit fields, predicate in the field domain; lane = one of my 32-row groups
						// (pred32: compile-time field positions), groups cut by a piece boundary lane = row
						const uint32_t mask = (uint32_t)mask64, flo = (uint32_t)q.flo, span = (uint32_t)q.span;
						for (uint32_t gb = g_lo; gb < g_hi; gb += 32u) {
							const uint32_t g = gb + lane;
							const uint32_t r_lo = g * 32u;
							if (g < g_hi && r_lo >= off_p && r_lo + 32u <= end_p) {
								const uint32_t bit = q.bit_base + (r_lo - off_p) * q.w;
								sel[g] &= pred32_any(q.w, words32 + (bit >> 5), bit & 31u, flo, span);
							}
						}
						__syncwarp();
#pragma unroll 1
						for (int e = 0; e < 2; e++) {
							const uint32_t edge = e == 0 ? off_p : end_p;
							const uint32_t g = edge >> 5;
							if ((edge & 31u) == 0 || g < g_lo || g >= g_hi) {
								continue;  // warp-uniform
							}
							const uint32_t i = g * 32u + lane;
							const bool inside = i >= off_p && i < end_p;
							const uint32_t bit = inside ? q.bit_base + (i - off_p) * q.w : q.bit_base;
							const uint32_t f = read_bits32(words32, bit) & mask;
							const uint32_t keep = __ballot_sync(0xffffffffu, !inside || (f - flo) <= span);
							if (lane == 0) {
								sel[g] &= keep;
							}
							__syncwarp();
						}
						off_p = end_p;
						continue;
					}
					uint32_t g = max(off_p >> 5, g_lo);  // first of my groups that touches the piece
					for (; g * 32u < end_p && g < g_hi; g++) {
						const uint32_t word = sel[g];
						if (word == 0) {
							continue;  // nothing left to test in these 32 rows
						}
						const uint32_t i = g * 32u + lane;
						const bool inside = i >= off_p && i < end_p;
						bool pass = true;
						if (inside) {
							const uint32_t bit = q.bit_base + (i - off_p) * q.w;
							const uint64_t f = q.w <= 32u ? (uint64_t)(read_bits32(words32, bit) & (uint32_t)mask64)
							                              : (read_bits64((const uint64_t *)words32, bit) & mask64);
							if (q.linear) {
								pass = (f - q.flo) <= q.span;
							} else {
								const uint64_t v = extend_rt(f + q.min_add, qc.type_size, qc.is_signed);
								pass = qc.is_signed ? ((int64_t)v >= qc.lo && (int64_t)v <= qc.hi)
								                    : (v >= (uint64_t)qc.lo && v <= (uint64_t)qc.hi);
							}
						}
						const uint32_t keep = __ballot_sync(0xffffffffu, pass);
						if (lane == 0) {
							sel[g] = word & keep;
						}
					}
					__syncwarp();
					off_p = end_p;
				}
				if (a.out_bitmap == nullptr && c + 1 < a.n_cols) {
					// survivors so far; at <= 25 % of my rows (and within the list) the remaining predicate columns
					// run on the compacted list
					const uint32_t g = g_lo + lane;  // gpw <= 32: blocks of <= 4096 rows
					const uint32_t word = g < g_hi ? sel[g] : 0u;
					const uint32_t cnt = __popc(word);
					uint32_t incl = cnt;
#pragma unroll
					for (int d = 1; d < 32; d <<= 1) {
						const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
						if ((int)lane >= d) {
							incl += v;
						}
					}
					const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
					const uint32_t my_groups = g_hi - g_lo;
					if (total <= kQListCap && total * 16u <= my_groups * 32u) {
						uint32_t pos = incl - cnt;
						uint32_t rest = word;
						while (rest) {
							const uint32_t bpos = (uint32_t)__ffs((int)rest) - 1u;
							my_list[pos++] = (uint16_t)(g * 32u + bpos);
							rest &= rest - 1u;
						}
						__syncwarp();
						n_list = total;
						list_mode = true;
					}
				}
			}

			// ---- survivors: bitmap out, count, compaction into the warp's list, aggregates on the list ----
			// sparse selections (Q6 keeps ~2 % of the rows): lane j owns my j-th group (8192 rows = 32 groups per warp),
			// an exclusive scan of the survivor counts places every lane's rows, each lane then emits its own set bits.
			// Same list order as the group-at-a-time loop below, which remains for dense selections.
			uint32_t g_first = g_lo;
			if (list_mode) {
				if (lane == 0) {
					n_pass += n_list;
				}
				g_first = g_hi;  // the list already holds the survivors of every predicate
			} else {
				const uint32_t g = g_lo + lane;
				const uint32_t word = g < g_hi ? sel[g] : 0u;
				const uint32_t cnt = __popc(word);
				uint32_t incl = cnt;
#pragma unroll
				for (int d = 1; d < 32; d <<= 1) {
					const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
					if ((int)lane >= d) {
						incl += v;
					}
				}
				const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
				const uint32_t my_groups = g_hi - g_lo;
				if (a.dense_groups && total * 4u >= my_groups * 32u) {
					// dense selection of a grouped scan: group at a time, lane = row, values straight into the thread's
					// shared accumulators of the row's group
					for (uint32_t gd = g_lo; gd < g_hi; gd++) {
						const uint32_t wd = sel[gd];
						if (a.out_bitmap != nullptr && lane == 0) {
							a.out_bitmap[word0 + gd] = wd;
						}
						if (wd == 0) {
							continue;
						}
						const bool on = (wd >> lane) & 1u;
						const uint32_t i = gd * 32u + (on ? lane : (uint32_t)__ffs((int)wd) - 1u);  // idle lanes shadow a live row
#pragma unroll 1
						for (uint32_t c = 0; c < a.n_cols; c++) {
							if (((a.factor_mask >> c) & 1u) || c == a.group_col) {
								const QueryCol &qc = a.cols[c];
								vals[c][tid] = group_value(st.piece[c], stage, gd * 32u, i, qc.type_size, qc.is_signed);
							}
						}
						const uint64_t kv = vals[a.group_col][tid];
						if (on && kv < (uint64_t)a.n_groups) {
							long long *slot = dacc + (size_t)kv * a.n_aggs * kQConsumers + tid;
#pragma unroll 1
							for (uint32_t k = 0; k < a.n_aggs; k++) {
								const QueryAgg &qa = a.aggs[k];
								long long term = 1;
#pragma unroll
								for (int f = 0; f < kQueryMaxFactors; f++) {
									if (f < (int)qa.n_factors) {
										term *= qa.add[f] + qa.mul[f] * (long long)vals[qa.col[f]][tid];
									}
								}
								slot[(size_t)k * kQConsumers] += term;
							}
							dcnt[(uint32_t)kv * kQConsumers + tid] += 1u;
						}
					}
					g_first = g_hi;
				} else if (total <= kQListCap) {
					if (a.out_bitmap != nullptr && g < g_hi) {
						a.out_bitmap[word0 + g] = word;
					}
					uint32_t pos = incl - cnt;
					uint32_t rest = word;
					while (rest) {
						const uint32_t bpos = (uint32_t)__ffs((int)rest) - 1u;
						my_list[pos++] = (uint16_t)(g * 32u + bpos);
						rest &= rest - 1u;
					}
					if (lane == 0) {
						n_pass += total;
					}
					n_list = total;
					g_first = g_hi;  // the loop below has nothing left to do
				}
			}
			for (uint32_t g = g_first; g < g_hi; g++) {
				const uint32_t word = sel[g];
				if (a.out_bitmap != nullptr && lane == 0) {
					a.out_bitmap[word0 + g] = word;
				}
				if (word == 0) {
					continue;
				}
				if (lane == 0) {
					n_pass += __popc(word);
				}
				if ((word >> lane) & 1u) {
					my_list[n_list + __popc(word & ((1u << lane) - 1u))] = (uint16_t)(g * 32u + lane);
				}
				n_list += __popc(word);
				if (n_list > kQListCap - 32u) {
					__syncwarp();
					if (a.n_groups) {
						aggregate_grouped(a, st, stage, my_list, n_list, gs.list[warp], gs.keys[warp], gacc, vals, tid, lane, warp,
						                  acc);
					} else {
						aggregate_list(a, st, stage, my_list, n_list, vals, tid, lane, acc);
					}
					n_list = 0;
					__syncwarp();
				}
			}
			__syncwarp();
			if (a.n_groups) {
				aggregate_grouped(a, st, stage, my_list, n_list, gs.list[warp], gs.keys[warp], gacc, vals, tid, lane, warp, acc);
			} else {
				aggregate_list(a, st, stage, my_list, n_list, vals, tid, lane, acc);
			}
			__syncwarp();
			if (lane == 0) {
				mbar_arrive(&empty_bar[s]);
			}
		}
	}
	if (a.n_groups) {
		__syncthreads();
		if (a.dense_groups && warp < (uint32_t)kQConsumerWarps) {
			// fold the per-thread accumulators: a warp per (group, aggregate) slot, into row 0 of the warp accumulators
			for (uint32_t sidx = warp; sidx < a.n_groups * a.n_aggs; sidx += kQConsumerWarps) {
				const uint32_t g = sidx / a.n_aggs, k = sidx % a.n_aggs;
				uint64_t lo = 0, hi = 0;
				uint32_t cnt = 0;
				for (uint32_t j = lane; j < (uint32_t)kQConsumers; j += 32u) {
					const long long v = dacc[(size_t)sidx * kQConsumers + j];
					add128(lo, hi, (uint64_t)v, v < 0 ? ~0ull : 0ull);
					cnt += dcnt[g * kQConsumers + j];
				}
#pragma unroll
				for (int d = 16; d > 0; d >>= 1) {
					const uint64_t olo = __shfl_down_sync(0xffffffffu, lo, d);
					const uint64_t ohi = __shfl_down_sync(0xffffffffu, hi, d);
					add128(lo, hi, olo, ohi);
					cnt += __shfl_down_sync(0xffffffffu, cnt, d);
				}
				if (lane == 0) {
					// slot (g, k) of warp row 0 is touched by this warp only here: the main loop is behind the barrier
					uint64_t slo = gacc.sum[0][g][k][0], shi = gacc.sum[0][g][k][1];
					add128(slo, shi, lo, hi);
					gacc.sum[0][g][k][0] = slo;
					gacc.sum[0][g][k][1] = shi;
					if (k == 0) {
						gacc.cnt[0][g] += cnt;
					}
				}
			}
		}
		if (a.dense_groups) {
			__syncthreads();
		}
		// one thread per (group, aggregate): add up the 8 warp rows, one atomic triple per CTA
		if (tid < a.n_groups * a.n_aggs) {
			const uint32_t g = tid / a.n_aggs, k = tid % a.n_aggs;
			uint64_t lo = 0, hi = 0, cnt = 0;
			for (int wv = 0; wv < kQConsumerWarps; wv++) {
				add128(lo, hi, gacc.sum[wv][g][k][0], gacc.sum[wv][g][k][1]);
				cnt += gacc.cnt[wv][g];
			}
			if (cnt | lo | hi) {
				Agg *out = a.out + g * a.n_aggs + k;
				const unsigned long long old = atomicAdd((unsigned long long *)&out->sum_lo, (unsigned long long)lo);
				const unsigned long long carry = (old + lo < old) ? 1ull : 0ull;
				atomicAdd((unsigned long long *)&out->sum_hi, (unsigned long long)(hi + carry));
				atomicAdd((unsigned long long *)&out->count, (unsigned long long)cnt);
			}
		}
		return;
	}
#pragma unroll
	for (int g = 0; g < kQueryMaxAggs; g++) {
		if (g < (int)a.n_aggs) {  // uniform
			block_reduce_agg((uint64_t)acc[g], (uint64_t)(acc[g] >> 64), n_pass, red, a.out + g);
		}
	}
}


// =====================================================================================================================
// scan_columns_sparse_kernel -- the ungrouped form (TPC-H Q6 shape: conjunctive ranges, few survivors, SUM of products)
// without the producer / consumer pipeline.  Every WARP works alone on blocks of 2048 table rows (runs of four
// consecutive blocks, so the segment state of a column is mostly reused):
//   A  lane = (column, piece) resolves the <= 3 segments of its column that overlap the block; the per-segment part
//      (frame of reference, field-domain predicate range) is cached in the lane's registers until the segment changes;
//   B  the packed bytes of the predicate pieces are copied into the warp's private shared-memory window with 16-byte
//      cp.async, all columns at once (one wait per block); piece parameters travel by shuffle, not through memory;
//   C  a lane owns two 32-row groups: pred32<W> per predicate piece (compile-time field positions); groups cut by a
//      piece boundary and pieces that are not field-comparable (wrap-around frames, > 32 bits, raw signed slots) are
//      evaluated lane = row;
//   D  the few survivors are compacted into a list (in the window) and their factor values are read from global memory.
// (Testing the later predicate columns on the compacted survivors of the first one -- what the pipelined kernel does --
// was measured slower here: 0.23 against 0.16 ms on TPC-H Q6; a listed row costs ~17 shuffles for its piece record.)
// No CTA barrier, no mbarrier, no shared selection state: latency is hidden by the 32 independent warps an SM holds.
// The integer pipe bounds this kernel (ncu: math-pipe throttle), so the design minimises instructions per row.
// =====================================================================================================================
constexpr int kSWarps = 8;
constexpr uint32_t kSRows = 2048;
constexpr int kSGroupsPerLane = (int)(kSRows / 1024u);
constexpr uint32_t kSRun = 4;  // consecutive blocks a warp takes at a time

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
	asm volatile("cp.async.wait_all;" ::: "memory");
}

__device__ __forceinline__ uint64_t load_slot(const uint8_t *p, uint32_t type_size) {
	return type_size == 8 ? *(const uint64_t *)p
	                      : (type_size == 4 ? (uint64_t) * (const uint32_t *)p
	                                        : (type_size == 2 ? (uint64_t) * (const uint16_t *)p : (uint64_t)*p));
}

// field k of a segment payload in global memory (packed words) or slot k of the raw arena
__device__ __forceinline__ uint64_t field_global(uint64_t words, uint32_t repacked, uint32_t k, uint32_t w, uint32_t type_size) {
	if (!repacked) {
		return load_slot((const uint8_t *)words + (uint64_t)k * type_size, type_size);
	}
	const uint64_t *gw = (const uint64_t *)words;
	const uint64_t bit = (uint64_t)k * w;
	const uint32_t off = (uint32_t)(bit & 63u);
	uint64_t f = __ldg(gw + (bit >> 6)) >> off;
	if (off + w > 64u) {
		f |= __ldg(gw + (bit >> 6) + 1) << (64u - off);
	}
	return f & lo_set64(w);
}

// decoded value of table row `row` of column c, straight from global memory (descriptor walk + <= 2 words)
__device__ __forceinline__ uint64_t row_value_global(const QueryCol &c, uint64_t row, int literal) {
	uint32_t s = __ldg(c.vec_seg + (row >> 11));
	DescQ dq = load_desc(c.desc + s);
	while (dq.row_start() + dq.count() <= row && s + 1 < c.n_segs) {
		s++;
		dq = load_desc(c.desc + s);
	}
	const uint32_t i = (uint32_t)(row - dq.row_start());
	const bool repacked = dq.state() & kSegRepacked;
	const uint64_t mn = dq.min_factor();
	const uint64_t min_add = (mn != ~0ull && (repacked || literal)) ? mn : 0ull;
	const uint64_t words = repacked ? dq.packed_ptr() : (uint64_t)c.raw + dq.row_start() * (uint64_t)c.type_size;
	return extend_rt(field_global(words, repacked, i, dq.width(), c.type_size) + min_add, c.type_size, c.is_signed);
}

__device__ __forceinline__ bool in_range(const QueryCol &c, uint64_t v) {
	return c.is_signed ? ((int64_t)v >= c.lo && (int64_t)v <= c.hi) : (v >= (uint64_t)c.lo && v <= (uint64_t)c.hi);
}

// value of block row i of column c for a surviving row: the column's <= 3 piece records sit in the registers of lanes
// 4c .. 4c+2 and are fetched by shuffle (every lane of the warp takes part; `live` lanes read the field)
__device__ __noinline__ uint64_t sparse_factor_value(uint32_t type_size, uint32_t is_signed, uint32_t c, uint32_t i, bool live, uint32_t row_lo,
                                                     uint32_t cnt, uint32_t first, uint32_t c_w, uint32_t c_repacked,
                                                     uint64_t c_words, uint64_t c_min_add) {
	const uint32_t lo1 = __shfl_sync(0xffffffffu, row_lo, c * kQLanesPerCol + 1);
	const uint32_t n1 = __shfl_sync(0xffffffffu, cnt, c * kQLanesPerCol + 1);
	const uint32_t lo2 = __shfl_sync(0xffffffffu, row_lo, c * kQLanesPerCol + 2);
	const uint32_t n2 = __shfl_sync(0xffffffffu, cnt, c * kQLanesPerCol + 2);
	const int p = (n2 && i >= lo2) ? 2 : ((n1 && i >= lo1) ? 1 : 0);
	const int src_lane = c * kQLanesPerCol + p;
	const uint32_t w = __shfl_sync(0xffffffffu, c_w, src_lane);
	const uint64_t words = __shfl_sync(0xffffffffu, c_words, src_lane);
	const uint64_t min_add = __shfl_sync(0xffffffffu, c_min_add, src_lane);
	const uint32_t rep = __shfl_sync(0xffffffffu, c_repacked, src_lane);
	const uint32_t k0 = __shfl_sync(0xffffffffu, first, src_lane);
	const uint32_t off_p = __shfl_sync(0xffffffffu, row_lo, src_lane);
	if (!live) {
		return 0;
	}
	return extend_rt(field_global(words, rep, k0 + (i - off_p), w, type_size) + min_add, type_size, is_signed);
}

template <int NA>  // upper bound of a.n_aggs (1, 2, 4, 8): the 128-bit accumulators live in registers
__global__ void __launch_bounds__(kSWarps * 32, 4) scan_columns_sparse_kernel(const QueryArgs a, uint32_t window_bytes) {
	extern __shared__ __align__(128) uint8_t s_mem[];
	__shared__ uint64_t red[96];
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	uint8_t *window = s_mem + (size_t)warp * window_bytes;
	uint16_t *my_list = (uint16_t *)window;  // the survivor list reuses the window once the predicates are done (>= 4 KiB)
	const uint64_t n_rows = a.row_end - a.row_begin;
	const uint32_t n_blocks = (uint32_t)((n_rows + kSRows - 1) / kSRows);
	const uint32_t ci = lane / kQLanesPerCol, pi = lane % kQLanesPerCol;
	const bool active = ci < a.n_cols && pi < (uint32_t)kQueryMaxPieces;
	const QueryCol &mycol = a.cols[active ? ci : 0];
	const uint32_t my_has_pred = active ? mycol.has_pred : 0u;

	// ---- per-lane segment cache ----
	uint32_t c_seg = ~0u, c_w = 0, c_linear = 0, c_repacked = 0, c_count = 0;
	uint64_t c_rs = 0, c_words = 0, c_min_add = 0, c_flo = ~0ull, c_span = 0;

	unsigned __int128 acc[NA];
#pragma unroll
	for (int g = 0; g < NA; g++) {
		acc[g] = 0;
	}
	uint64_t n_pass = 0;

	const uint32_t total_warps = gridDim.x * kSWarps, wg = blockIdx.x * kSWarps + warp;
	for (uint32_t run = wg; run * kSRun < n_blocks; run += total_warps) {
		for (uint32_t b = run * kSRun; b < min(n_blocks, (run + 1u) * kSRun); b++) {
			const uint64_t r0 = a.row_begin + (uint64_t)b * kSRows;  // multiple of 2048: vec_seg is exact
			const uint32_t rows = (uint32_t)(a.row_end - r0 < kSRows ? a.row_end - r0 : kSRows);
			const uint32_t n_groups = (rows + 31u) >> 5;
			// ---- A: my piece ----
			uint32_t cnt = 0, row_lo = 0, first = 0, bit_base = 0, bytes = 0;
			uint64_t src16 = 0;
			if (active) {
				const uint32_t s = __ldg(mycol.vec_seg + (r0 >> 11)) + pi;
				if (s < mycol.n_segs) {
					if (s != c_seg) {
						const DescQ dq = load_desc(mycol.desc + s);
						c_seg = s;
						c_rs = dq.row_start();
						c_count = dq.count();
						c_w = dq.width();
						c_repacked = (dq.state() & kSegRepacked) ? 1u : 0u;
						const uint64_t mn = dq.min_factor();
						c_min_add = (mn != ~0ull && (c_repacked || a.literal)) ? mn : 0ull;
						c_words = c_repacked ? dq.packed_ptr() : (uint64_t)mycol.raw + c_rs * (uint64_t)mycol.type_size;
						c_linear = piece_is_linear(mycol, c_min_add, c_w, c_repacked) ? 1u : 0u;
						c_flo = ~0ull;
						c_span = 0;
						if (c_linear && my_has_pred) {
							field_range(mycol, c_min_add, c_w, c_flo, c_span);
						}
					}
					const uint64_t re = c_rs + c_count;
					const uint64_t lo = c_rs > r0 ? c_rs : r0;
					const uint64_t hi = re < r0 + rows ? re : r0 + rows;
					if (hi > lo) {
						cnt = (uint32_t)(hi - lo);
						row_lo = (uint32_t)(lo - r0);
						first = (uint32_t)(lo - c_rs);
						const uint64_t bit = (uint64_t)first * c_w;
						const uint64_t src = c_repacked ? c_words + (bit >> 3) : c_words + (uint64_t)first * mycol.type_size;
						src16 = src & ~15ull;
						const uint32_t byte_base = (uint32_t)(src - src16);
						bit_base = byte_base * 8u + (c_repacked ? (uint32_t)(bit & 7u) : 0u);
						bytes = (byte_base + (((uint32_t)(bit & 7u) + cnt * c_w + 7u) >> 3) + 15u) & ~15u;
					}
				}
			}
			const uint64_t mask64 = lo_set64(c_w);
			const bool pred_piece = cnt && my_has_pred;
			const bool all_pass = pred_piece && c_linear && c_flo == 0 && c_span == mask64;  // zone map
			const bool fast = pred_piece && !all_pass && c_linear && c_w <= 32u && (c_flo >> 32) == 0;
			const bool generic = pred_piece && !all_pass && !fast;
			const uint32_t my_bytes = fast ? bytes : 0u;
			uint32_t incl = my_bytes;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
				if ((int)lane >= d) {
					incl += v;
				}
			}
			const uint32_t smem_off = incl - my_bytes;
			// every column must be covered by its three pieces, else the block takes the row-at-a-time path
			uint32_t covered = cnt;
			covered += __shfl_xor_sync(0xffffffffu, covered, 1);
			covered += __shfl_xor_sync(0xffffffffu, covered, 2);
			const bool slow_block = __any_sync(0xffffffffu, active && covered != rows);
			const uint32_t fast_mask = slow_block ? 0u : __ballot_sync(0xffffffffu, fast);
			const uint32_t generic_mask = slow_block ? 0u : __ballot_sync(0xffffffffu, generic);
			__syncwarp();  // the previous block's survivor list (in the window) has been consumed
			// ---- B: stage the predicate pieces ----
			for (uint32_t m = fast_mask; m; m &= m - 1u) {
				const int src_lane = __ffs((int)m) - 1;
				const uint8_t *src = (const uint8_t *)__shfl_sync(0xffffffffu, src16, src_lane);
				const uint32_t nb = __shfl_sync(0xffffffffu, my_bytes, src_lane);
				uint8_t *dst = window + __shfl_sync(0xffffffffu, smem_off, src_lane);
				for (uint32_t off = lane * 16u; off < nb; off += 512u) {
					cp_async16(dst + off, src + off);
				}
			}
			// ---- selection words of my groups (lane + 32 j) ----
			uint32_t sel[kSGroupsPerLane];
#pragma unroll
			for (int j = 0; j < kSGroupsPerLane; j++) {
				const uint32_t g = lane + 32u * j;
				sel[j] = 0;
				if (g < n_groups) {
					const uint32_t left = rows - g * 32u;
					sel[j] = left >= 32u ? 0xffffffffu : ((1u << left) - 1u);
					if (a.in_bitmap != nullptr) {
						sel[j] &= __ldg(a.in_bitmap + ((r0 - a.row_begin) >> 5) + g);
					}
				}
			}
			if (fast_mask) {
				cp_async_wait_all();
				__syncwarp();
			}
			// ---- C: predicates (pieces in column order) ----
			{
			for (uint32_t m = fast_mask; m; m &= m - 1u) {
				const int src_lane = __ffs((int)m) - 1;
				const uint32_t w = __shfl_sync(0xffffffffu, c_w, src_lane);
				const uint32_t flo = __shfl_sync(0xffffffffu, (uint32_t)c_flo, src_lane);
				const uint32_t span = __shfl_sync(0xffffffffu, (uint32_t)c_span, src_lane);
				const uint32_t bb = __shfl_sync(0xffffffffu, bit_base, src_lane);
				const uint32_t off_p = __shfl_sync(0xffffffffu, row_lo, src_lane);
				const uint32_t end_p = off_p + __shfl_sync(0xffffffffu, cnt, src_lane);
				const uint32_t *words32 = (const uint32_t *)(window + __shfl_sync(0xffffffffu, smem_off, src_lane));
#pragma unroll 1
				for (int j = 0; j < kSGroupsPerLane; j++) {
					const uint32_t r_lo = (lane + 32u * j) * 32u;
					if (r_lo >= off_p && r_lo + 32u <= end_p) {  // (groups beyond the block fail this: end_p <= rows)
						const uint32_t bit = bb + (r_lo - off_p) * w;
						const uint32_t keep = pred32_any(w, words32 + (bit >> 5), bit & 31u, flo, span);
						sel[0] &= j == 0 ? keep : 0xffffffffu;
						sel[1] &= j == 1 ? keep : 0xffffffffu;
					}
				}
				if ((off_p | end_p) & 31u) {  // a piece boundary inside a group: lane = row
					const uint32_t mask = w >= 32u ? 0xffffffffu : ((1u << w) - 1u);
#pragma unroll 1
					for (int e = 0; e < 2; e++) {
						const uint32_t edge = e == 0 ? off_p : end_p;
						if ((edge & 31u) == 0) {
							continue;
						}
						const uint32_t ge = edge >> 5;
						const uint32_t i = ge * 32u + lane;
						const bool inside = i >= off_p && i < end_p;
						const uint32_t bit = inside ? bb + (i - off_p) * w : bb;
						const uint32_t f = read_bits32(words32, bit) & mask;
						const uint32_t keep = __ballot_sync(0xffffffffu, !inside || (f - flo) <= span);
#pragma unroll
						for (int j = 0; j < kSGroupsPerLane; j++) {
							if (lane + 32u * j == ge) {
								sel[j] &= keep;
							}
						}
					}
				}
			}
			for (uint32_t m = generic_mask; m; m &= m - 1u) {
				// not field-comparable: decode the rows of the piece from global memory, lane = row
				const int src_lane = __ffs((int)m) - 1;
				const QueryCol &qc = a.cols[src_lane / kQLanesPerCol];
				const uint32_t w = __shfl_sync(0xffffffffu, c_w, src_lane);
				const uint64_t words = __shfl_sync(0xffffffffu, c_words, src_lane);
				const uint64_t min_add = __shfl_sync(0xffffffffu, c_min_add, src_lane);
				const uint32_t rep = __shfl_sync(0xffffffffu, c_repacked, src_lane);
				const uint32_t k0 = __shfl_sync(0xffffffffu, first, src_lane);
				const uint32_t off_p = __shfl_sync(0xffffffffu, row_lo, src_lane);
				const uint32_t end_p = off_p + __shfl_sync(0xffffffffu, cnt, src_lane);
				for (uint32_t gg = off_p >> 5; gg * 32u < end_p; gg++) {
					const uint32_t i = gg * 32u + lane;
					bool pass = true;
					if (i >= off_p && i < end_p) {
						const uint64_t f = field_global(words, rep, k0 + (i - off_p), w, qc.type_size);
						pass = in_range(qc, extend_rt(f + min_add, qc.type_size, qc.is_signed));
					}
					const uint32_t keep = __ballot_sync(0xffffffffu, pass);
#pragma unroll
					for (int j = 0; j < kSGroupsPerLane; j++) {
						if (lane + 32u * j == gg) {
							sel[j] &= keep;
						}
					}
				}
			}
			}
			if (slow_block) {
#pragma unroll 1
				for (uint32_t c = 0; c < a.n_cols; c++) {
					const QueryCol &qc = a.cols[c];
					if (!qc.has_pred) {
						continue;
					}
					for (uint32_t gg = 0; gg < n_groups; gg++) {
						const uint32_t i = gg * 32u + lane;
						const bool pass = i >= rows || in_range(qc, row_value_global(qc, r0 + i, a.literal));
						const uint32_t keep = __ballot_sync(0xffffffffu, pass);
#pragma unroll
						for (int j = 0; j < kSGroupsPerLane; j++) {
							if (lane + 32u * j == gg) {
								sel[j] &= keep;
							}
						}
					}
				}
			}
			// ---- D: survivors ----
			uint32_t my_cnt = 0;
#pragma unroll
			for (int j = 0; j < kSGroupsPerLane; j++) {
				const uint32_t g = lane + 32u * j;
				if (a.out_bitmap != nullptr && g < n_groups) {
					a.out_bitmap[((r0 - a.row_begin) >> 5) + g] = sel[j];
				}
				my_cnt += __popc(sel[j]);
			}
			uint32_t incl_c = my_cnt;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t v = __shfl_up_sync(0xffffffffu, incl_c, d);
				if ((int)lane >= d) {
					incl_c += v;
				}
			}
			const uint32_t total = __shfl_sync(0xffffffffu, incl_c, 31);
			if (total == 0) {
				continue;
			}
			if (lane == 0) {
				n_pass += total;
			}
			__syncwarp();  // every lane is past its last read of the staged columns: the list may overwrite them
			uint32_t pos = incl_c - my_cnt;
#pragma unroll
			for (int j = 0; j < kSGroupsPerLane; j++) {
				uint32_t rest = sel[j];
				while (rest) {
					const uint32_t bpos = (uint32_t)__ffs((int)rest) - 1u;
					my_list[pos++] = (uint16_t)((lane + 32u * j) * 32u + bpos);
					rest &= rest - 1u;
				}
			}
			__syncwarp();
			for (uint32_t e0 = 0; e0 < total; e0 += 32u) {
				const uint32_t e = e0 + lane;
				const bool live = e < total;
				const uint32_t i = live ? my_list[e] : 0u;  // block row
#pragma unroll
				for (int k = 0; k < NA; k++) {
					if (k < (int)a.n_aggs) {  // uniform
						const QueryAgg &qa = a.aggs[k];
						__int128 term = 1;
						int64_t term64 = 1;
#pragma unroll 1
						for (uint32_t f = 0; f < qa.n_factors; f++) {
							const uint32_t c = qa.col[f];
							const QueryCol &qc = a.cols[c];
							int64_t x;
							if (slow_block) {
								x = live ? (int64_t)row_value_global(qc, r0 + i, a.literal) : 0;
							} else {
								x = (int64_t)sparse_factor_value(qc.type_size, qc.is_signed, c, i, live, row_lo, cnt, first, c_w, c_repacked, c_words,
								                                 c_min_add);
							}
							if (a.narrow_products) {
								term64 *= qa.add[f] + qa.mul[f] * x;
							} else {
								term *= (__int128)qa.add[f] + (__int128)qa.mul[f] * (__int128)x;
							}
						}
						if (live) {
							acc[k] += a.narrow_products ? (unsigned __int128)(__int128)term64 : (unsigned __int128)term;
						}
					}
				}
			}
		}
	}
#pragma unroll
	for (int g = 0; g < NA; g++) {
		if (g < (int)a.n_aggs) {  // uniform
			block_reduce_agg((uint64_t)acc[g], (uint64_t)(acc[g] >> 64), n_pass, red, a.out + g);
		}
	}
}


// =====================================================================================================================
// scan_columns_dense_kernel -- GROUP BY a small-domain column with most rows selected (TPC-H Q1 shape), warp per block.
// The host cuts the table rows at every segment boundary of every referenced column ("uniform ranges": inside one, each
// column is ONE segment), so a block of <= 1024 rows of a range has exactly one piece per column -- no piece search, no
// boundary groups.  Per block: lane c resolves column c (cached per segment), all columns are staged with cp.async,
// predicates run lane = group (pred32), then the 32-row groups are walked lane = row: every needed column is decoded
// ONCE into the thread's slice of shared memory (value = extend(min) + field on a linear frame of reference), the sum
// expressions read their factors from there, and every aggregate is one read-modify-write of the thread's own int64
// accumulator of the row's group.  The planner guarantees (from the analyzed min / max) that every product and every
// per-thread partial sum fits int64; the 128-bit fold happens once at the end.
// =====================================================================================================================
constexpr int kDWarps = 4;
constexpr int kDThreads = kDWarps * 32;
constexpr int kDIlp = 8;  // 32-row groups a lane works on at a time

// one step of the factor program resolved against the current block (per warp, in shared memory)
struct __align__(16) DenseStep {
	uint32_t win_off;    // byte offset of the column's staged piece inside the warp's window
	uint32_t bit_base;   // bit position of the block's first row inside it
	uint32_t w_flags;    // width | fresh << 8 | emit << 16
	uint32_t mask;
	long long base;      // add + mul * extend(min)
	long long mul;
};
static_assert(sizeof(DenseStep) == 32, "two 16-byte broadcast loads per step");
constexpr uint32_t kDRows = kDenseBlockRows;

template <int NA, bool N32>
__global__ void __launch_bounds__(kDThreads, 4) scan_columns_dense_kernel(const QueryArgs a, const DenseRange *ranges,
                                                                          uint32_t n_ranges, uint32_t n_blocks,
                                                                          uint32_t window_bytes) {
	extern __shared__ __align__(128) uint8_t d_mem[];
	__shared__ uint64_t red[96];
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	// dynamic shared memory: windows | per-warp step tables | accumulators [group][agg][thread] | counts
	uint8_t *window = d_mem + (size_t)warp * window_bytes;
	DenseStep *tab = (DenseStep *)(d_mem + (size_t)kDWarps * window_bytes) + (size_t)warp * kQueryMaxSteps;
	long long *dacc = (long long *)(d_mem + (size_t)kDWarps * window_bytes + (size_t)kDWarps * kQueryMaxSteps * sizeof(DenseStep));
	uint32_t *dcnt = (uint32_t *)(dacc + (size_t)a.n_groups * a.n_aggs * kDThreads);
	for (uint32_t i = tid; i < a.n_groups * a.n_aggs * kDThreads; i += kDThreads) {
		dacc[i] = 0ll;
	}
	for (uint32_t i = tid; i < a.n_groups * kDThreads; i += kDThreads) {
		dcnt[i] = 0u;
	}
	__syncthreads();
	const bool active = lane < a.n_cols;
	const QueryCol &mycol = a.cols[active ? lane : 0];
	// my step of the factor program (lane = step) and the column it reads
	const QueryStep &mystep = a.steps[lane < a.n_steps ? lane : 0];
	const uint32_t step_col = mystep.col;
	// per-lane segment cache (lane = column)
	uint32_t c_seg = ~0u, c_w = 0;
	uint64_t c_rs = 0, c_words = 0, c_vbase = 0, c_flo = ~0ull, c_span = 0;

	const uint32_t total_warps = gridDim.x * kDWarps, wg = blockIdx.x * kDWarps + warp;
	for (uint32_t run = wg; run * kSRun < n_blocks; run += total_warps) {
		for (uint32_t b = run * kSRun; b < min(n_blocks, (run + 1u) * kSRun); b++) {
			// range of block b: last range whose first_block <= b
			uint32_t lo = 0, hi = n_ranges;
			while (hi - lo > 1) {
				const uint32_t mid = (lo + hi) >> 1;
				if (__ldg(&ranges[mid].first_block) <= b) {
					lo = mid;
				} else {
					hi = mid;
				}
			}
			const DenseRange *rg = ranges + lo;
			const uint32_t k = b - __ldg(&rg->first_block);
			const uint64_t r0 = __ldg(&rg->row_begin) + (uint64_t)k * kDRows;
			const uint32_t left = __ldg(&rg->rows) - k * kDRows;
			const uint32_t rows = left < kDRows ? left : kDRows;
			const uint32_t n_groups32 = (rows + 31u) >> 5;
			// ---- my column's piece ----
			uint32_t bit_base = 0, bytes = 0;
			uint64_t src16 = 0;
			if (active) {
				const uint32_t sg = __ldg(&rg->seg[lane]);
				if (sg != c_seg) {
					const DescQ dq = load_desc(mycol.desc + sg);
					c_seg = sg;
					c_rs = dq.row_start();
					c_w = dq.width();
					c_words = dq.packed_ptr();
					const uint64_t mn = dq.min_factor();
					const uint64_t min_add = mn != ~0ull ? mn : 0ull;  // (every segment here is re-packed: host check)
					c_vbase = extend_rt(min_add, mycol.type_size, mycol.is_signed);
					c_flo = ~0ull;
					c_span = 0;
					if (mycol.has_pred) {
						field_range(mycol, min_add, c_w, c_flo, c_span);
					}
				}
				const uint64_t bit = (r0 - c_rs) * (uint64_t)c_w;
				const uint64_t src = c_words + (bit >> 3);
				src16 = src & ~15ull;
				const uint32_t byte_base = (uint32_t)(src - src16);
				bit_base = byte_base * 8u + (uint32_t)(bit & 7u);
				bytes = (byte_base + (((uint32_t)(bit & 7u) + rows * c_w + 7u) >> 3) + 15u) & ~15u;
			}
			uint32_t incl = bytes;
#pragma unroll
			for (int d = 1; d < 32; d <<= 1) {
				const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
				if ((int)lane >= d) {
					incl += v;
				}
			}
			const uint32_t smem_off = incl - bytes;
			__syncwarp();  // the previous block's window has been consumed
			for (uint32_t c = 0; c < a.n_cols; c++) {
				const uint8_t *src = (const uint8_t *)__shfl_sync(0xffffffffu, src16, c);
				const uint32_t nb = __shfl_sync(0xffffffffu, bytes, c);
				uint8_t *dst = window + __shfl_sync(0xffffffffu, smem_off, c);
				for (uint32_t off = lane * 16u; off < nb; off += 512u) {
					cp_async16(dst + off, src + off);
				}
			}
			uint32_t sel = 0;
			if (lane < n_groups32) {
				const uint32_t l = rows - lane * 32u;
				sel = l >= 32u ? 0xffffffffu : ((1u << l) - 1u);
			}
			cp_async_wait_all();
			__syncwarp();
			// ---- predicates: lane = group ----
#pragma unroll 1
			for (uint32_t c = 0; c < a.n_cols; c++) {
				if (!a.cols[c].has_pred) {
					continue;
				}
				const uint32_t w = __shfl_sync(0xffffffffu, c_w, c);
				const uint64_t flo = __shfl_sync(0xffffffffu, c_flo, c);
				const uint64_t span = __shfl_sync(0xffffffffu, c_span, c);
				const uint32_t bb = __shfl_sync(0xffffffffu, bit_base, c);
				const uint32_t *words32 = (const uint32_t *)(window + __shfl_sync(0xffffffffu, smem_off, c));
				if (flo == 0 && span == lo_set64(w)) {
					continue;  // zone map
				}
				if (flo >> 32) {
					sel = 0;  // empty intersection (flo = 2^64 - 1)
					continue;
				}
				const uint32_t bit = bb + lane * 32u * w;
				const uint32_t keep = pred32_any(w, words32 + (bit >> 5), bit & 31u, (uint32_t)flo, (uint32_t)span);
				sel &= keep;
			}
			if (!__any_sync(0xffffffffu, sel != 0)) {
				continue;
			}
			// ---- the block's step table: lane s resolves step s against the piece of the column it reads.  A value is
			// extend(min) + field, so x = add + mul * value = (add + mul * extend(min)) + mul * field: the row loop never
			// materialises the value (each column of TPC-H Q1 is read by exactly one step) ----
			{
				const uint32_t sw = __shfl_sync(0xffffffffu, c_w, step_col);
				const uint32_t sbb = __shfl_sync(0xffffffffu, bit_base, step_col);
				const uint32_t soff = __shfl_sync(0xffffffffu, smem_off, step_col);
				const uint64_t svb = __shfl_sync(0xffffffffu, c_vbase, step_col);
				__syncwarp();  // the previous block's rows are done with the table
				if (lane < a.n_steps) {
					DenseStep e;
					e.win_off = soff;
					e.bit_base = sbb;
					e.w_flags = sw | ((uint32_t)mystep.fresh << 8) | ((uint32_t)mystep.emit << 16);
					e.mask = 0xffffffffu >> (32u - sw);
					e.base = (long long)((unsigned long long)mystep.add + (unsigned long long)mystep.mul * svb);
					e.mul = mystep.mul;
					tab[lane] = e;
				}
				__syncwarp();
			}
			const uint32_t gw = __shfl_sync(0xffffffffu, c_w, a.group_col);
			const uint32_t gbb = __shfl_sync(0xffffffffu, bit_base, a.group_col);
			const uint32_t *gwin = (const uint32_t *)(window + __shfl_sync(0xffffffffu, smem_off, a.group_col));
			const uint64_t gvb = __shfl_sync(0xffffffffu, c_vbase, a.group_col);
			const uint32_t gmask = 0xffffffffu >> (32u - gw);
			// ---- rows: lane = row, kDIlp 32-row groups per iteration (independent dependency chains per lane: the
			// kernel is latency-bound at 12 warps per SM) ----
#pragma unroll 1
			for (uint32_t gg = 0; gg < n_groups32; gg += kDIlp) {
				uint32_t word[kDIlp], idx[kDIlp];
				uint32_t any = 0;
#pragma unroll
				for (int r = 0; r < kDIlp; r++) {
					const bool in = gg + r < n_groups32;
					word[r] = in ? __shfl_sync(0xffffffffu, sel, (gg + r) & 31u) : 0u;
					idx[r] = (in ? gg + r : gg) * 32u + lane;
					any |= word[r];
				}
				if (any == 0) {
					continue;
				}
				bool ok[kDIlp];
				uint32_t kvi[kDIlp];
				long long *slot[kDIlp];
				long long term[kDIlp];
#pragma unroll
				for (int r = 0; r < kDIlp; r++) {
					const uint32_t f = read_bits32(gwin, gbb + idx[r] * gw) & gmask;
					const uint64_t kv = N32 ? (uint64_t)(long long)(int)((uint32_t)gvb + f) : gvb + (uint64_t)f;
					ok[r] = ((word[r] >> lane) & 1u) && kv < (uint64_t)a.n_groups;
					kvi[r] = ok[r] ? (uint32_t)kv : 0u;
					slot[r] = dacc + (size_t)kvi[r] * a.n_aggs * kDThreads + tid;
					term[r] = 0;
				}
#pragma unroll 1
				for (uint32_t st = 0; st < a.n_steps; st++) {
					const DenseStep e = tab[st];  // broadcast
					const uint32_t *win32 = (const uint32_t *)(window + e.win_off);
					const uint32_t w = e.w_flags & 0xffu;
					const bool fresh = (e.w_flags >> 8) & 1u;
#pragma unroll
					for (int r = 0; r < kDIlp; r++) {
						const uint32_t g = read_bits32(win32, e.bit_base + idx[r] * w) & e.mask;
						long long x;
						if (N32) {
							// every value and every factor fits int32 (planner: narrow_values); wrap-around arithmetic
							x = (long long)(int)((uint32_t)e.base + (uint32_t)e.mul * g);
						} else {
							x = e.base + e.mul * (long long)g;
						}
						term[r] = fresh ? x : term[r] * x;
					}
					const uint32_t emit = e.w_flags >> 16;
					if (emit) {  // uniform
#pragma unroll
						for (int r = 0; r < kDIlp; r++) {
							if (ok[r]) {
								slot[r][(size_t)(emit - 1u) * kDThreads] += term[r];
							}
						}
					}
				}
#pragma unroll
				for (int r = 0; r < kDIlp; r++) {
					if (ok[r]) {
						dcnt[kvi[r] * kDThreads + tid] += 1u;
					}
				}
			}
		}
	}
	__syncthreads();
	// fold the per-thread accumulators: one warp per (group, aggregate) slot at a time, one atomic triple each
	for (uint32_t sidx = warp; sidx < a.n_groups * a.n_aggs; sidx += kDWarps) {
		const uint32_t g = sidx / a.n_aggs, kk = sidx % a.n_aggs;
		uint64_t lo = 0, hi = 0;
		uint64_t cnt = 0;
		for (uint32_t j = lane; j < (uint32_t)kDThreads; j += 32u) {
			const long long v = dacc[(size_t)sidx * kDThreads + j];
			add128(lo, hi, (uint64_t)v, v < 0 ? ~0ull : 0ull);
			cnt += dcnt[g * kDThreads + j];
		}
#pragma unroll
		for (int d = 16; d > 0; d >>= 1) {
			const uint64_t olo = __shfl_down_sync(0xffffffffu, lo, d);
			const uint64_t ohi = __shfl_down_sync(0xffffffffu, hi, d);
			add128(lo, hi, olo, ohi);
			cnt += __shfl_down_sync(0xffffffffu, cnt, d);
		}
		if (lane == 0 && (cnt | lo | hi)) {
			Agg *out = a.out + g * a.n_aggs + kk;
			const unsigned long long old = atomicAdd((unsigned long long *)&out->sum_lo, (unsigned long long)lo);
			const unsigned long long carry = (old + lo < old) ? 1ull : 0ull;
			atomicAdd((unsigned long long *)&out->sum_hi, (unsigned long long)(hi + carry));
			atomicAdd((unsigned long long *)&out->count, (unsigned long long)cnt);
		}
	}
	(void)red;
}

}  // namespace

uint32_t scan_columns_group_smem(uint32_t n_groups, uint32_t n_aggs, uint32_t dense) {
	uint32_t bytes = (uint32_t)((sizeof(GroupScratch) + 127u) & ~127u);
	if (dense) {
		bytes += n_groups * n_aggs * (uint32_t)kQConsumers * 8u + n_groups * (uint32_t)kQConsumers * 4u;
	}
	return bytes;
}

cudaError_t launch_scan_columns(const QueryArgs &args, int sm_count, cudaStream_t stream) {
	if (args.row_end <= args.row_begin || args.n_aggs == 0) {
		return cudaSuccess;
	}
	int dev = 0;
	cudaGetDevice(&dev);
	if (args.n_groups && args.dense_ranges != nullptr && args.in_bitmap == nullptr && args.out_bitmap == nullptr) {
		const size_t smem = (size_t)kDWarps * args.dense_window + (size_t)kDWarps * kQueryMaxSteps * sizeof(DenseStep) +
		                    (size_t)args.n_groups * args.n_aggs * kDThreads * 8 + (size_t)args.n_groups * kDThreads * 4;
		void (*kernel)(const QueryArgs, const DenseRange *, uint32_t, uint32_t, uint32_t);
		if (args.narrow_values) {
			kernel = args.n_aggs <= 2 ? scan_columns_dense_kernel<2, true>
			                          : (args.n_aggs <= 4 ? scan_columns_dense_kernel<4, true> : scan_columns_dense_kernel<8, true>);
		} else {
			kernel = args.n_aggs <= 2 ? scan_columns_dense_kernel<2, false>
			                          : (args.n_aggs <= 4 ? scan_columns_dense_kernel<4, false> : scan_columns_dense_kernel<8, false>);
		}
		cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		int per_sm = 0;
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kDThreads, smem);
		if (e != cudaSuccess) {
			return e;
		}
		uint64_t grid = (uint64_t)sm_count * (uint64_t)(per_sm > 0 ? per_sm : 1);
		const uint64_t need = ((args.dense_blocks + kSRun - 1) / kSRun + kDWarps - 1) / kDWarps;
		if (grid > need) {
			grid = need;
		}
		kernel<<<(uint32_t)grid, kDThreads, smem, stream>>>(args, args.dense_ranges, args.dense_n_ranges, args.dense_blocks,
		                                                     args.dense_window);
		return cudaGetLastError();
	}
	if (args.n_groups == 0 && args.sparse_window) {
		// ungrouped: the warp-per-block kernel; window = staged bytes of the predicate columns of 1024 rows
		const size_t smem = (size_t)kSWarps * args.sparse_window;
		void (*kernel)(const QueryArgs, uint32_t) =
		    args.n_aggs <= 1 ? scan_columns_sparse_kernel<1>
		                     : (args.n_aggs <= 2 ? scan_columns_sparse_kernel<2>
		                                         : (args.n_aggs <= 4 ? scan_columns_sparse_kernel<4> : scan_columns_sparse_kernel<8>));
		cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		int per_sm = 0;
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kSWarps * 32, smem);
		if (e != cudaSuccess) {
			return e;
		}
		const uint64_t n_blocks = (args.row_end - args.row_begin + kSRows - 1) / kSRows;
		uint64_t grid = (uint64_t)sm_count * (uint64_t)(per_sm > 0 ? per_sm : 1);
		const uint64_t need = ((n_blocks + kSRun - 1) / kSRun + kSWarps - 1) / kSWarps;  // a warp takes runs of kSRun blocks
		if (grid > need) {
			grid = need;
		}
		kernel<<<(uint32_t)grid, kSWarps * 32, smem, stream>>>(args, args.sparse_window);
		return cudaGetLastError();
	}
	static size_t configured[64] = {};
	const size_t smem = (size_t)args.stages * args.stage_bytes +
	                    (args.n_groups ? scan_columns_group_smem(args.n_groups, args.n_aggs, args.dense_groups) : 0);
	if (configured[dev & 63] < smem) {
		cudaError_t e = cudaFuncSetAttribute(scan_columns_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		configured[dev & 63] = smem;
	}
	const uint64_t n_blocks = (args.row_end - args.row_begin + args.block_rows - 1) / args.block_rows;
	int per_sm = 0;
	cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, scan_columns_kernel, kQThreads, smem);
	if (e != cudaSuccess) {
		return e;
	}
	uint64_t grid = (uint64_t)sm_count * (uint64_t)(per_sm > 0 ? per_sm : 1);
	if (grid > n_blocks) {
		grid = n_blocks;
	}
	scan_columns_kernel<<<(uint32_t)grid, kQThreads, smem, stream>>>(args);
	return cudaGetLastError();
}

}  // namespace adacom
