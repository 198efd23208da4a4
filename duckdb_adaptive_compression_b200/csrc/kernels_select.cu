// kernels_select.cu -- the general form of a table filter pushed to ONE column, fused with the decode:
// ColumnSegment::FilterSelection (src/storage/table/column_segment.cpp:685-844) over the conjunction of
// CONSTANT_COMPARISON (TemplatedFilterSelection, :575-586: `(!HAS_NULL || mask.RowIsValid(idx)) && OP(vec[idx], c)`),
// IS_NULL / IS_NOT_NULL (TemplatedNullSelection, :660-683) children, combined by CONJUNCTION_AND (:717-723) or
// CONJUNCTION_OR (:688-716), refining an incoming selection (the `sel` vector) and followed by the aggregate DuckDB
// runs above the scan (SUM / COUNT skip NULL rows through the validity column).
//
// This is the catch-all path: any width 1..64, raw slots, validity mask, selection in / out.  Conjunctions that amount
// to one value range over the valid rows (adacom_capi.cu: canonical_filter_range) run on the field-domain kernel
// (kernels_filter.cu, filter_words_kernel<.., ALLOW>); this kernel decodes per value and serves the rest (IS NULL
// children, two separate ranges, 8-byte columns with slot-width segments).  Job = (segment, 2048-row tile) from the column's static tile table; lane = consecutive rows, so a ballot is
// 32 consecutive selection bits.  Algorithmic bytes: packed payload + 1 bit per row per bitmap / validity word read.
#include "adacom_kernels.cuh"

namespace adacom {

namespace {

constexpr int kSelThreads = 256;

template <int S, bool SIGNED>
__global__ void __launch_bounds__(kSelThreads) select_kernel(SelectArgs a) {
	using T = typename UIntOf<S>::type;
	__shared__ uint64_t red[96];
	const uint32_t tid = threadIdx.x, lane = tid & 31u;
	Acc96 acc;
	uint32_t n_sel = 0, n_valid = 0;
	for (uint32_t j = blockIdx.x; j < a.n_jobs; j += gridDim.x) {
		const Job job = a.jobs[j];
		const SegDesc *d = a.desc + job.seg;
		const uint32_t count = d->count;
		const uint32_t first = job.k * kVectorSize;
		if (first >= count) {
			continue;
		}
		const uint32_t cnt = min(kVectorSize, count - first);
		const bool repacked = d->state & kSegRepacked;
		const uint32_t w = repacked ? d->width : 8u * S;
		const uint64_t mn = d->min_factor;
		const uint64_t min_add = (mn != ~0ull && (repacked || a.literal)) ? mn : 0ull;
		const uint64_t mask = lo_set64(w);
		const uint64_t row0 = d->row_start + first;  // column row of the tile's first value
		const uint64_t *words = (const uint64_t *)d->packed_ptr;
		const T *slots = (const T *)a.raw + row0;
		for (uint32_t base = 0; base < cnt; base += kSelThreads) {
			const uint32_t i = base + tid;
			const bool in = i < cnt;
			bool pass = false, valid = false;
			uint64_t v = 0;
			if (in) {
				const uint64_t row = row0 + i;
				uint64_t f;
				if (repacked) {
					const uint64_t bit = (uint64_t)(first + i) * w;
					const uint32_t off = (uint32_t)(bit & 63u);
					f = __ldg(words + (bit >> 6)) >> off;
					if (off + w > 64u) {
						f |= __ldg(words + (bit >> 6) + 1) << (64u - off);
					}
					f &= mask;
				} else {
					f = (uint64_t)__ldg(slots + i);
				}
				v = extend_low<S, SIGNED>(f + min_add);
				valid = a.validity == nullptr || ((__ldg(a.validity + (row >> 6)) >> (row & 63u)) & 1ull);
				// conjunction over the children; zero children: every row passes
				bool any = false, all = true;
				for (uint32_t p = 0; p < a.n_preds; p++) {
					const SelectPred &pr = a.preds[p];
					bool r;
					if (pr.kind == 0) {
						r = valid && cmp_passes<SIGNED>(v, pr.constant, pr.op_mask);
					} else {
						r = valid != (pr.kind == 1);
					}
					any |= r;
					all &= r;
				}
				pass = a.conj_or ? (a.n_preds == 0 || any) : all;
				if (pass && a.in_bitmap != nullptr) {
					const uint64_t rel = row - a.out_row0;
					pass = (__ldg(a.in_bitmap + (rel >> 6)) >> (rel & 63u)) & 1ull;
				}
			}
			if (pass) {
				n_sel++;
				if (valid) {
					acc.add<SIGNED>(v);
					n_valid++;
				}
			}
			if (a.out_bitmap32 != nullptr) {
				const uint32_t ballot = __ballot_sync(0xffffffffu, pass);
				if (lane == 0 && ballot) {
					const uint64_t rel = row0 + (base + (tid & ~31u)) - a.out_row0;
					const uint32_t sh = (uint32_t)(rel & 31u);
					atomicOr(a.out_bitmap32 + (rel >> 5), ballot << sh);
					if (sh) {
						atomicOr(a.out_bitmap32 + (rel >> 5) + 1, ballot >> (32u - sh));
					}
				}
			}
		}
	}
	const uint32_t wv = __reduce_add_sync(0xffffffffu, n_valid);
	if (lane == 0 && wv) {
		atomicAdd((unsigned long long *)&a.agg->reserved, (unsigned long long)wv);
	}
	block_reduce_agg(acc.lo64(), acc.hi64(), (uint64_t)n_sel, red, a.agg);
}

template <int S>
cudaError_t launch_s(bool is_signed, const SelectArgs &a, uint32_t grid, cudaStream_t stream) {
	if (is_signed) {
		select_kernel<S, true><<<grid, kSelThreads, 0, stream>>>(a);
	} else {
		select_kernel<S, false><<<grid, kSelThreads, 0, stream>>>(a);
	}
	return cudaGetLastError();
}

}  // namespace

cudaError_t launch_select(TypeInfo t, const SelectArgs &a, int sm_count, cudaStream_t stream) {
	if (a.n_jobs == 0) {
		return cudaSuccess;
	}
	const uint32_t resident = (uint32_t)sm_count * 8u;
	const uint32_t grid = a.n_jobs < resident ? a.n_jobs : resident;
	switch (t.size) {
	case 1: return launch_s<1>(t.is_signed, a, grid, stream);
	case 2: return launch_s<2>(t.is_signed, a, grid, stream);
	case 4: return launch_s<4>(t.is_signed, a, grid, stream);
	case 8: return launch_s<8>(t.is_signed, a, grid, stream);
	default: return cudaErrorInvalidValue;
	}
}

}  // namespace adacom
