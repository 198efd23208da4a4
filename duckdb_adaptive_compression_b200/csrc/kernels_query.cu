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
//   * 8 consumer warps: thread = row; unpack the referenced columns straight from shared memory (funnel
//     shift at bit_base + i*w), apply the range predicates (AND), and fold prod_k(add_k + mul_k * value_k)
//     into 128-bit per-thread accumulators; optional selection bitmap in (AND) / out.
//
// Algorithmic HBM bytes per row: sum over the columns of w_c / 8 (+ 1/8 per bitmap).
#include "adacom_kernels.cuh"

namespace adacom {

namespace {

constexpr int kQConsumerWarps = 8;
constexpr int kQConsumers = kQConsumerWarps * 32;
constexpr int kQThreads = kQConsumers + 32;
constexpr int kQMaxStages = 3;
constexpr int kQLanesPerCol = 4;  // producer lanes per column: piece 0..2, lane 3 idle

struct QPiece {
	uint64_t min_add;   // frame of reference to add (0 when none)
	uint32_t smem_off;  // byte offset of the staged range inside the stage (16-byte aligned)
	uint32_t bit_base;  // bit position of the piece's first field inside the staged range (0..127)
	uint32_t cnt;       // values of this piece inside the block (0: piece unused)
	uint32_t w;
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
	const QPiece q = pc[p];
	const uint64_t *words = (const uint64_t *)(stage + q.smem_off);
	const uint32_t bit = q.bit_base + off * q.w;
	const uint64_t mask = lo_set64(q.w);
	const uint64_t f = q.w <= 32u ? (uint64_t)(read_bits32((const uint32_t *)words, bit) & (uint32_t)mask)
	                              : (read_bits64(words, bit) & mask);
	return extend_rt(f + q.min_add, type_size, is_signed);
}

struct PieceRes {
	uint64_t src16;
	uint64_t min_add;
	uint32_t bytes;  // 16-byte rounded staged range (0: no piece)
	uint32_t bit_base;
	uint32_t cnt;
	uint32_t w;
};

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
	return r;
}

__global__ void __launch_bounds__(kQThreads, 2) scan_columns_kernel(const QueryArgs a) {
	extern __shared__ __align__(128) uint8_t q_stage_mem[];
	__shared__ __align__(8) uint64_t full_bar[kQMaxStages];
	__shared__ __align__(8) uint64_t empty_bar[kQMaxStages];
	__shared__ QStage stages[kQMaxStages];
	__shared__ uint64_t vals[kQueryMaxCols][kQConsumers];  // decoded factors of the row a thread is working on
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
				mbar_wait(&empty_bar[s], ((seq / n_stages) & 1u) ^ 1u);
				stages[s].row0 = r0;
				stages[s].rows = rows;
				stages[s].done = 0;
			}
			__syncwarp();
			if (active) {
				QPiece q;
				q.min_add = pr.min_add;
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
			mbar_wait(&empty_bar[s], ((seq / n_stages) & 1u) ^ 1u);
			stages[s].done = 1;
			mbar_arrive(&full_bar[s]);
		}
	} else {
		// ---------------- consumer warps: thread = row ----------------
		for (uint32_t it = 0;; it++) {
			const uint32_t s = it % n_stages;
			mbar_wait(&full_bar[s], (it / n_stages) & 1u);
			const QStage &st = stages[s];
			if (st.done) {
				break;
			}
			const uint8_t *stage = q_stage_mem + (size_t)s * a.stage_bytes;
			const uint32_t rows = st.rows;
			const uint64_t rel0 = st.row0 - a.row_begin;  // multiple of 2048
			for (uint32_t base = 0; base < rows; base += kQConsumers) {
				const uint32_t i = base + tid;
				bool pass = i < rows;
				if (a.in_bitmap != nullptr) {
					const uint32_t word = __ldg(a.in_bitmap + ((rel0 + base + (tid & ~31u)) >> 5));
					pass = pass && ((word >> lane) & 1u);
				}
#pragma unroll
				for (int c = 0; c < kQueryMaxCols; c++) {
					if (c < (int)a.n_cols && pass) {
						const QueryCol &qc = a.cols[c];
						const uint64_t v = column_value(st.piece[c], stage, i, qc.type_size, qc.is_signed);
						if (qc.has_pred) {
							pass = qc.is_signed ? ((int64_t)v >= qc.lo && (int64_t)v <= qc.hi)
							                    : (v >= (uint64_t)qc.lo && v <= (uint64_t)qc.hi);
						}
						vals[c][tid] = v;
					}
				}
				if (a.out_bitmap != nullptr) {
					const uint32_t ballot = __ballot_sync(0xffffffffu, pass);
					if (lane == 0 && base + (tid & ~31u) < rows) {
						a.out_bitmap[(rel0 + base + (tid & ~31u)) >> 5] = ballot;
					}
				}
				if (pass) {
					n_pass++;
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
			__syncwarp();
			if (lane == 0) {
				mbar_arrive(&empty_bar[s]);
			}
		}
	}
#pragma unroll
	for (int g = 0; g < kQueryMaxAggs; g++) {
		if (g < (int)a.n_aggs) {  // uniform
			block_reduce_agg((uint64_t)acc[g], (uint64_t)(acc[g] >> 64), n_pass, red, a.out + g);
		}
	}
}

}  // namespace

cudaError_t launch_scan_columns(const QueryArgs &args, int sm_count, cudaStream_t stream) {
	static size_t configured[64] = {};
	const size_t smem = (size_t)args.stages * args.stage_bytes;
	int dev = 0;
	cudaGetDevice(&dev);
	if (configured[dev & 63] < smem) {
		cudaError_t e = cudaFuncSetAttribute(scan_columns_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) {
			return e;
		}
		configured[dev & 63] = smem;
	}
	if (args.row_end <= args.row_begin || args.n_aggs == 0) {
		return cudaSuccess;
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
