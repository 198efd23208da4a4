// adacom_capi.cu -- host implementation of the C ABI declared in include/adacom_gpu.h.
//
// Owns the HBM layout of a column (raw arena, packed extents, descriptor table), builds the work-unit
// tables and enqueues the kernels of kernels_{encode,decode,misc}.cu on the context's compute stream.
// There is no CPU compute path in this file: every entry point either launches kernels or fails.
#include "../../include/adacom_gpu.h"
#include "adacom_kernels.cuh"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <cmath>
#include <map>
#include <memory>
#include <mutex>
#include <new>
#include <string>
#include <tuple>
#include <vector>

using namespace adacom;

static_assert(sizeof(adacom_segment_meta) == 64, "adacom_segment_meta must stay 64 bytes");
static_assert(sizeof(adacom_agg) == sizeof(Agg), "adacom_agg layout");

namespace {
std::string g_process_error;
constexpr size_t kArenaPad = 256;          // slack behind every arena for the 16-byte rounded bulk copies
constexpr size_t kSmallFetch = 64;         // fetch batches up to this size go through the host-mapped page
}  // namespace

struct adacom_ctx {
	int device = 0;
	int sm_count = 148;
	cudaStream_t stream = nullptr;
	std::string err;
	uint64_t launches = 0;
	std::mutex mu;
	void *flush_buf = nullptr;
	size_t flush_bytes = 0;
	Agg *d_agg = nullptr;
	unsigned long long *d_totals = nullptr;  // device alias of h_totals (host-mapped)
	unsigned long long ticket = 0;           // sequence number of the last layout launch
	unsigned long long *d_layout_scratch = nullptr;  // chained scan of layout_kernel (2 words per CTA, <= 128 CTAs)
	Agg *d_aggs = nullptr;                // adacom_scan_columns: kQueryMaxAggs aggregates (lazily allocated)
	Agg *h_aggs = nullptr;                // pinned
	Agg *h_agg = nullptr;                 // pinned
	unsigned long long *h_totals = nullptr;  // pinned
	// uniform ranges of the last grouped scan (scan_columns_dense_kernel): rebuilt when the columns / epochs / rows change
	std::vector<uint64_t> dense_key;
	DenseRange *d_dense_ranges = nullptr;
	uint32_t dense_n_ranges = 0, dense_blocks = 0;
	size_t dense_capacity = 0;
	// staging pair of the batched look-ups (pinned host + device, grown on demand): one H2D, one kernel, one D2H
	uint8_t *h_stage = nullptr, *d_stage = nullptr;
	size_t stage_bytes = 0;
	// host-mapped page for the smallest fetch batches (fetch_row of one row id): the kernel reads the ids and writes
	// the answers through the mapping, so the call is one launch + one synchronisation instead of copy, launch, copy
	uint8_t *h_small = nullptr, *d_small = nullptr;
	// window rings of closed scan streams (pinned host + device buffer + copy stream + events), reused by the next
	// adacom_scan_stream_open: allocating 24 MiB of pinned memory costs milliseconds, a scan opens one stream per
	// device column it crosses
	struct ScanRing {
		uint8_t *d_ring = nullptr, *h_ring = nullptr;
		size_t bytes = 0;
		cudaStream_t copy_stream = nullptr;
		std::vector<cudaEvent_t> events;
	};
	std::vector<ScanRing> ring_cache;
};

// Every entry point that touches the device holds the context mutex AND makes the context's device current: one
// process may drive several GPUs (one context each, row group g -> device g mod G), and a host thread's current device
// is whatever its last call left behind.
struct CtxLock {
	explicit CtxLock(adacom_ctx *ctx) : guard(ctx->mu) {
		cudaSetDevice(ctx->device);
	}
	std::lock_guard<std::mutex> guard;
};

struct Extent {
	uint64_t *base = nullptr;
	uint64_t words = 0;      // payload words in use (incl. alignment gaps)
	uint64_t capacity = 0;   // allocated words
	uint64_t live = 0;       // repacked segments still pointing into this extent (valid after gc_extents)
	uint64_t seg_begin = 0, seg_end = 0;  // the encode call that created it covered [seg_begin, seg_end)
};

struct adacom_column {
	adacom_ctx *ctx = nullptr;
	adacom_ptype ptype = ADACOM_INT64;
	TypeInfo t {8, true};
	uint64_t n_values = 0;
	uint64_t n_segs = 0;
	uint8_t *raw = nullptr;
	SegDesc *d_desc = nullptr;
	std::vector<SegDesc> h_desc;
	bool mirror_dirty = false;
	std::vector<Extent> extents;  // indexed by SegDesc::extent; a dead slot (base == nullptr) is reused by the next encode
	size_t latest = 0;            // slot of the most recent encode: "the packed arena" of the accessors
	std::vector<Extent> spares;  // dead extents kept for the next encode (no allocator call on the hot path)
	// 2048-value tiles of every segment, (segment, k) pairs in segment order: depends on the segment counts only,
	// so it is built once; the tiles of the segments [b, e) are d_tiles[tile_first[b] .. tile_first[e])
	Job *d_tiles = nullptr;
	std::vector<uint32_t> tile_first;
	// segment holding row 2048*v (static, like the tile table): entry point of the multi-column row-block scans
	uint32_t *d_vec_seg = nullptr;
	std::vector<uint32_t> vec_seg;
	uint32_t *d_parts = nullptr;       // device word behind d_vec_seg: scan chunks per segment currently needed
	uint32_t raw_parts = 1;            // that number for slot-width (raw) segments: static
	uint32_t max_span[3] = {0, 0, 0};  // most segments any aligned block of 2048 / 4096 / 8192 rows touches (static)
	uint32_t wmax = 0;                 // widest live payload, refreshed with the host mirror
	bool any_raw = false;              // some non-empty segment still lives in the raw arena (mirror)
	uint64_t epoch = 0;
};

namespace adacom {
void set_process_error(const std::string &msg) {
	g_process_error = msg;
}
}  // namespace adacom

namespace {

adacom_status fail(adacom_ctx *ctx, adacom_status st, const std::string &msg) {
	static const bool verbose = getenv("ADACOM_VERBOSE_ERRORS") != nullptr;  // hosts that only forward the status code
	if (verbose) {
		fprintf(stderr, "[adacom] status %d: %s\n", (int)st, msg.c_str());
	}
	if (ctx) {
		ctx->err = msg;
	} else {
		g_process_error = msg;
	}
	return st;
}

#define CK(ctx, call)                                                                                        \
	do {                                                                                                     \
		cudaError_t e__ = (call);                                                                            \
		if (e__ != cudaSuccess) {                                                                            \
			return fail((ctx), e__ == cudaErrorMemoryAllocation ? ADACOM_ERR_NOMEM : ADACOM_ERR_CUDA,       \
			            std::string(#call) + ": " + cudaGetErrorString(e__));                               \
		}                                                                                                    \
	} while (0)

#define LAUNCH(ctx, call)  \
	do {                   \
		CK((ctx), (call)); \
		(ctx)->launches++; \
	} while (0)

bool type_info(adacom_ptype p, TypeInfo &t) {
	switch (p) {
	case ADACOM_UINT8: t = {1, false}; return true;
	case ADACOM_INT8: t = {1, true}; return true;
	case ADACOM_UINT16: t = {2, false}; return true;
	case ADACOM_INT16: t = {2, true}; return true;
	case ADACOM_UINT32: t = {4, false}; return true;
	case ADACOM_INT32: t = {4, true}; return true;
	case ADACOM_UINT64: t = {8, false}; return true;
	case ADACOM_INT64: t = {8, true}; return true;
	default: return false;
	}
}

// A dead extent becomes a spare for the next encode; at most two spares are kept per column.
adacom_status retire_extent(adacom_column *col, Extent &x) {
	adacom_ctx *ctx = col->ctx;
	if (x.base) {
		Extent spare;
		spare.base = x.base;
		spare.capacity = x.capacity;
		col->spares.push_back(spare);
		if (col->spares.size() > 2) {
			size_t smallest = 0;
			for (size_t i = 1; i < col->spares.size(); i++) {
				if (col->spares[i].capacity < col->spares[smallest].capacity) {
					smallest = i;
				}
			}
			CK(ctx, cudaFreeAsync(col->spares[smallest].base, ctx->stream));
			col->spares.erase(col->spares.begin() + smallest);
		}
	}
	x.base = nullptr;
	x.words = 0;
	x.capacity = 0;
	x.live = 0;
	return ADACOM_OK;
}

// free every packed extent that no repacked segment references any more (needs a fresh mirror)
adacom_status gc_extents(adacom_column *col) {
	adacom_ctx *ctx = col->ctx;
	for (auto &x : col->extents) {
		x.live = 0;
	}
	for (const SegDesc &d : col->h_desc) {
		if ((d.state & kSegRepacked) && d.extent < col->extents.size()) {
			col->extents[d.extent].live++;
		}
	}
	for (auto &x : col->extents) {
		if (x.base && x.live == 0) {
			adacom_status st = retire_extent(col, x);
			if (st != ADACOM_OK) {
				return st;
			}
		}
	}
	return ADACOM_OK;
}

adacom_status sync_mirror(adacom_column *col) {
	adacom_ctx *ctx = col->ctx;
	if (col->mirror_dirty) {
		CK(ctx, cudaMemcpyAsync(col->h_desc.data(), col->d_desc, col->n_segs * sizeof(SegDesc), cudaMemcpyDeviceToHost,
		                        ctx->stream));
		CK(ctx, cudaStreamSynchronize(ctx->stream));
		col->mirror_dirty = false;
		col->wmax = 0;
		return gc_extents(col);
	}
	return ADACOM_OK;
}

// widest live payload / raw residency of the column, cached per mirror refresh (query planning)
void refresh_widths(adacom_column *col) {
	if (col->wmax) {
		return;
	}
	uint32_t wmax = 1;
	bool any_raw = false;
	for (const SegDesc &d : col->h_desc) {
		wmax = std::max<uint32_t>(wmax, d.width);
		any_raw = any_raw || (d.count && !(d.state & kSegRepacked));
	}
	col->wmax = wmax;
	col->any_raw = any_raw;
}

adacom_status build_tiles(adacom_column *col) {
	adacom_ctx *ctx = col->ctx;
	std::vector<Job> jobs;
	col->tile_first.assign(col->n_segs + 1, 0);
	for (uint64_t g = 0; g < col->n_segs; g++) {
		col->tile_first[g] = (uint32_t)jobs.size();
		const uint32_t n = (col->h_desc[g].count + kVectorSize - 1) / kVectorSize;
		for (uint32_t k = 0; k < n; k++) {
			jobs.push_back({(uint32_t)g, k});
		}
	}
	col->tile_first[col->n_segs] = (uint32_t)jobs.size();
	{
		const uint64_t n_vec = (col->n_values + kVectorSize - 1) / kVectorSize;
		col->vec_seg.assign(n_vec + 1, 0);
		uint64_t g = 0;
		for (uint64_t v = 0; v < n_vec; v++) {
			const uint64_t row = v * kVectorSize;
			while (g + 1 < col->n_segs && row >= col->h_desc[g].row_start + col->h_desc[g].count) {
				g++;
			}
			col->vec_seg[v] = (uint32_t)g;
		}
		for (int ri = 0; ri < 3; ri++) {
			const uint64_t R = (uint64_t)kVectorSize << ri;
			uint32_t worst = 0;
			uint64_t gl = 0;
			for (uint64_t r0 = 0; r0 < col->n_values; r0 += R) {
				const uint64_t last = std::min<uint64_t>(r0 + R, col->n_values) - 1;
				gl = std::max<uint64_t>(gl, col->vec_seg[last / kVectorSize]);
				while (gl + 1 < col->n_segs && last >= col->h_desc[gl].row_start + col->h_desc[gl].count) {
					gl++;
				}
				worst = std::max<uint32_t>(worst, (uint32_t)(gl - col->vec_seg[r0 / kVectorSize] + 1));
			}
			col->max_span[ri] = worst;
		}
		// the word behind the table: scan chunks per segment (starts at the live widths; kept current on the device)
		uint32_t parts = 1;
		col->raw_parts = 1;
		for (const SegDesc &d : col->h_desc) {
			if (d.count) {
				const uint32_t raw_gran = scan_chunk_values(8u * (uint32_t)col->t.size);
				col->raw_parts = std::max<uint32_t>(col->raw_parts, (d.count + raw_gran - 1) / raw_gran);
				const uint32_t gran = scan_chunk_values(d.width ? d.width : 8u * (uint32_t)col->t.size);
				parts = std::max<uint32_t>(parts, (d.count + gran - 1) / gran);
			}
		}
		col->vec_seg[n_vec] = parts;
		CK(ctx, cudaMallocAsync((void **)&col->d_vec_seg, (n_vec + 1) * sizeof(uint32_t), ctx->stream));
		CK(ctx, cudaMemcpyAsync(col->d_vec_seg, col->vec_seg.data(), (n_vec + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice,
		                        ctx->stream));
		col->d_parts = col->d_vec_seg + n_vec;
	}
	CK(ctx, cudaMallocAsync((void **)&col->d_tiles, (jobs.size() + 1) * sizeof(Job), ctx->stream));
	if (!jobs.empty()) {
		// pageable source: staged before the call returns
		CK(ctx, cudaMemcpyAsync(col->d_tiles, jobs.data(), jobs.size() * sizeof(Job), cudaMemcpyHostToDevice, ctx->stream));
	}
	return ADACOM_OK;
}

bool range_ok(const adacom_column *col, uint64_t b, uint64_t e) {
	return col && b <= e && e <= col->n_segs;
}

uint64_t seg_bits(const SegDesc &d, int type_size) {
	return (d.state & kSegRepacked) ? (uint64_t)d.count * d.width : (uint64_t)d.capacity * 8u * (uint64_t)type_size;
}

void fill_meta(const adacom_column *col, uint64_t g, adacom_segment_meta *m) {
	const SegDesc &d = col->h_desc[g];
	memset(m, 0, sizeof(*m));
	m->row_start = d.row_start;
	m->count = d.count;
	m->word_offset = d.word_offset;
	m->words = (d.state & kSegRepacked) ? ((uint64_t)d.count * d.width + 63) >> 6 : 0;
	m->min_factor = d.min_factor;
	m->max_factor = d.max_factor;
	if (d.state & kSegUncompressedFn) {
		m->size_in_bytes = (uint64_t)d.capacity * col->t.size;  // ColumnSegment::segment_size
	} else {
		m->size_in_bytes = 9 + 8 * ((seg_bits(d, col->t.size) + 63) >> 6);  // sdsl::size_in_bytes
	}
	m->width = d.width;
	m->state = d.state;
	m->type_size = (uint8_t)col->t.size;
	m->is_signed = col->t.is_signed;
	m->capacity = d.capacity;
}

// pinned host + device staging pair of the batched look-up paths, grown on demand (the caller holds the lock)
adacom_status ensure_stage(adacom_ctx *ctx, size_t need) {
	if (ctx->stage_bytes >= need) {
		return ADACOM_OK;
	}
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	cudaFreeHost(ctx->h_stage);
	cudaFree(ctx->d_stage);
	ctx->h_stage = ctx->d_stage = nullptr;
	ctx->stage_bytes = 0;
	const size_t cap = need + need / 2 + 4096;
	CK(ctx, cudaHostAlloc((void **)&ctx->h_stage, cap, cudaHostAllocDefault));
	CK(ctx, cudaMalloc((void **)&ctx->d_stage, cap));
	ctx->stage_bytes = cap;
	return ADACOM_OK;
}

adacom_status ensure_raw(adacom_column *col) {
	if (!col->raw) {
		CK(col->ctx, cudaMallocAsync((void **)&col->raw, col->n_values * col->t.size + kArenaPad, col->ctx->stream));
	}
	return ADACOM_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------------

extern "C" int adacom_abi_version(void) {
	return ADACOM_ABI_VERSION;
}

extern "C" const char *adacom_last_error(const adacom_ctx *ctx) {
	return ctx ? ctx->err.c_str() : g_process_error.c_str();
}

extern "C" adacom_status adacom_ctx_create(int device, adacom_ctx **out) {
	if (!out) {
		return fail(nullptr, ADACOM_ERR_INVALID, "adacom_ctx_create: out is NULL");
	}
	*out = nullptr;
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) {
		return fail(nullptr, ADACOM_ERR_NO_DEVICE,
		            std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
		                " (this codec has no CPU fallback)");
	}
	if (device < 0 || device >= n) {
		return fail(nullptr, ADACOM_ERR_INVALID, "adacom_ctx_create: device ordinal out of range");
	}
	cudaDeviceProp prop;
	CK(nullptr, cudaGetDeviceProperties(&prop, device));
	if (prop.major != 10) {
		return fail(nullptr, ADACOM_ERR_NO_DEVICE,
		            std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major * 10 + prop.minor) +
		                "; the kernels are built for sm_100a only");
	}
	CK(nullptr, cudaSetDevice(device));
	adacom_ctx *ctx = new (std::nothrow) adacom_ctx();
	if (!ctx) {
		return fail(nullptr, ADACOM_ERR_NOMEM, "out of host memory");
	}
	ctx->device = device;
	ctx->sm_count = prop.multiProcessorCount;
	CK(nullptr, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
	CK(nullptr, cudaMalloc((void **)&ctx->d_agg, sizeof(Agg)));
	CK(nullptr, cudaHostAlloc((void **)&ctx->h_agg, sizeof(Agg), cudaHostAllocDefault));
	CK(nullptr, cudaMalloc((void **)&ctx->d_layout_scratch, 256 * sizeof(unsigned long long)));
	CK(nullptr, cudaMemset(ctx->d_layout_scratch, 0, 256 * sizeof(unsigned long long)));
	// {payload words, ticket} written by layout_kernel straight into host memory; the encode path spins on the
	// ticket (a few microseconds) instead of a copy + stream synchronisation
	CK(nullptr, cudaHostAlloc((void **)&ctx->h_totals, 4 * sizeof(unsigned long long), cudaHostAllocMapped));
	memset(ctx->h_totals, 0, 4 * sizeof(unsigned long long));
	CK(nullptr, cudaHostGetDevicePointer((void **)&ctx->d_totals, ctx->h_totals, 0));
	CK(nullptr, cudaHostAlloc((void **)&ctx->h_small, kSmallFetch * 16, cudaHostAllocMapped));
	CK(nullptr, cudaHostGetDevicePointer((void **)&ctx->d_small, ctx->h_small, 0));
	// keep freed extents cached in the stream-ordered pool instead of returning them to the driver
	cudaMemPool_t pool;
	if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
		uint64_t threshold = ~0ull;
		cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
	}
	*out = ctx;
	return ADACOM_OK;
}

extern "C" void adacom_ctx_destroy(adacom_ctx *ctx) {
	if (!ctx) {
		return;
	}
	cudaSetDevice(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	cudaFree(ctx->flush_buf);
	cudaFree(ctx->d_agg);
	cudaFree(ctx->d_layout_scratch);
	if (ctx->d_dense_ranges) {
		cudaFree(ctx->d_dense_ranges);
	}
	cudaFree(ctx->d_aggs);
	cudaFreeHost(ctx->h_aggs);
	cudaFreeHost(ctx->h_agg);
	cudaFreeHost(ctx->h_totals);
	cudaFreeHost(ctx->h_small);
	cudaFreeHost(ctx->h_stage);
	cudaFree(ctx->d_stage);
	for (auto &r : ctx->ring_cache) {
		cudaStreamSynchronize(r.copy_stream);
		for (auto e : r.events) {
			cudaEventDestroy(e);
		}
		cudaFreeHost(r.h_ring);
		cudaFree(r.d_ring);
		cudaStreamDestroy(r.copy_stream);
	}
	cudaStreamDestroy(ctx->stream);
	delete ctx;
}

extern "C" adacom_status adacom_ctx_synchronize(adacom_ctx *ctx) {
	if (!ctx) {
		return ADACOM_ERR_INVALID;
	}
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	return ADACOM_OK;
}

extern "C" void *adacom_ctx_stream(adacom_ctx *ctx) {
	return ctx ? (void *)ctx->stream : nullptr;
}

extern "C" uint64_t adacom_ctx_launch_count(const adacom_ctx *ctx) {
	return ctx ? ctx->launches : 0;
}

extern "C" adacom_status adacom_event_create(adacom_ctx *ctx, void **event_out) {
	if (!ctx || !event_out) {
		return ADACOM_ERR_INVALID;
	}
	cudaEvent_t ev;
	CK(ctx, cudaSetDevice(ctx->device));  // an event belongs to the device that was current when it was created
	CK(ctx, cudaEventCreate(&ev));
	*event_out = (void *)ev;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_event_record(adacom_ctx *ctx, void *event) {
	if (!ctx || !event) {
		return ADACOM_ERR_INVALID;
	}
	CK(ctx, cudaEventRecord((cudaEvent_t)event, ctx->stream));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_event_elapsed_ms(adacom_ctx *ctx, void *start, void *stop, float *ms_out) {
	if (!ctx || !start || !stop || !ms_out) {
		return ADACOM_ERR_INVALID;
	}
	CK(ctx, cudaEventSynchronize((cudaEvent_t)stop));
	CK(ctx, cudaEventElapsedTime(ms_out, (cudaEvent_t)start, (cudaEvent_t)stop));
	return ADACOM_OK;
}

extern "C" void adacom_event_destroy(adacom_ctx *ctx, void *event) {
	(void)ctx;
	if (event) {
		cudaEventDestroy((cudaEvent_t)event);
	}
}

extern "C" adacom_status adacom_host_alloc(adacom_ctx *ctx, size_t bytes, void **out) {
	if (!ctx || !out) {
		return ADACOM_ERR_INVALID;
	}
	cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
	if (e != cudaSuccess) {
		*out = nullptr;
		return fail(ctx, ADACOM_ERR_NOMEM, std::string("cudaHostAlloc: ") + cudaGetErrorString(e));
	}
	return ADACOM_OK;
}

extern "C" void adacom_host_free(adacom_ctx *ctx, void *ptr) {
	(void)ctx;
	if (ptr) {
		cudaFreeHost(ptr);
	}
}

extern "C" adacom_status adacom_device_alloc(adacom_ctx *ctx, size_t bytes, void **out) {
	if (!ctx || !out) {
		return ADACOM_ERR_INVALID;
	}
	CtxLock lock(ctx);
	cudaError_t e = cudaMalloc(out, bytes ? bytes : 1);
	if (e != cudaSuccess) {
		*out = nullptr;
		return fail(ctx, ADACOM_ERR_NOMEM, std::string("cudaMalloc: ") + cudaGetErrorString(e));
	}
	return ADACOM_OK;
}

extern "C" void adacom_device_free(adacom_ctx *ctx, void *ptr) {
	if (ctx && ptr) {
		CtxLock lock(ctx);
		cudaStreamSynchronize(ctx->stream);  // kernels of this context may still read the buffer
		cudaFree(ptr);
	}
}

extern "C" adacom_status adacom_flush_l2(adacom_ctx *ctx) {
	if (!ctx) {
		return ADACOM_ERR_INVALID;
	}
	CtxLock lock(ctx);
	if (!ctx->flush_buf) {
		ctx->flush_bytes = (size_t)256 << 20;  // 2x the 126 MB L2
		CK(ctx, cudaMalloc(&ctx->flush_buf, ctx->flush_bytes));
	}
	CK(ctx, cudaMemsetAsync(ctx->flush_buf, 0, ctx->flush_bytes, ctx->stream));
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// column life cycle
// ---------------------------------------------------------------------------------------------------

static adacom_status column_alloc(adacom_ctx *ctx, adacom_ptype type, uint64_t n_segments, adacom_column **out) {
	TypeInfo t;
	if (!type_info(type, t)) {
		return fail(ctx, ADACOM_ERR_UNSUPPORTED, "physical type not supported by the succinct codec");
	}
	if (n_segments >= (1ull << 32)) {
		return fail(ctx, ADACOM_ERR_INVALID, "too many segments");
	}
	adacom_column *col = new (std::nothrow) adacom_column();
	if (!col) {
		return fail(ctx, ADACOM_ERR_NOMEM, "out of host memory");
	}
	col->ctx = ctx;
	col->ptype = type;
	col->t = t;
	col->n_segs = n_segments;
	col->h_desc.assign(n_segments, SegDesc());
	*out = col;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_column_create(adacom_ctx *ctx, adacom_ptype type, uint64_t n_segments,
                                              const uint64_t *seg_counts, uint32_t seg_capacity, adacom_column **out) {
	if (!ctx || !out || (n_segments && !seg_counts)) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_column_create: NULL argument");
	}
	*out = nullptr;
	CtxLock lock(ctx);
	CK(ctx, cudaSetDevice(ctx->device));
	adacom_column *col = nullptr;
	adacom_status st = column_alloc(ctx, type, n_segments, &col);
	if (st != ADACOM_OK) {
		return st;
	}
	const uint32_t cap = seg_capacity ? seg_capacity : kBlockSize / (uint32_t)col->t.size;
	if ((uint64_t)cap * col->t.size > kBlockSize) {
		delete col;
		return fail(ctx, ADACOM_ERR_INVALID, "segment capacity exceeds one block (Storage::BLOCK_SIZE)");
	}
	uint64_t row = 0;
	for (uint64_t g = 0; g < n_segments; g++) {
		if (seg_counts[g] > cap) {
			delete col;
			return fail(ctx, ADACOM_ERR_INVALID, "segment count exceeds its capacity");
		}
		SegDesc &d = col->h_desc[g];
		memset(&d, 0, sizeof(d));
		d.row_start = row;
		d.min_factor = ~0ull;  // column_segment.cpp:96
		d.max_factor = 0;
		d.count = (uint32_t)seg_counts[g];
		d.capacity = cap;
		d.width = (uint8_t)(8 * col->t.size);  // succinct_vec.width(type_size * 8), column_segment.cpp:102
		row += seg_counts[g];
	}
	col->n_values = row;
	cudaError_t e = cudaMallocAsync((void **)&col->raw, row * col->t.size + kArenaPad, ctx->stream);
	if (e == cudaSuccess) {
		e = cudaMallocAsync((void **)&col->d_desc, std::max<uint64_t>(n_segments, 1) * sizeof(SegDesc), ctx->stream);
	}
	if (e == cudaSuccess && n_segments) {
		e = cudaMemcpyAsync(col->d_desc, col->h_desc.data(), n_segments * sizeof(SegDesc), cudaMemcpyHostToDevice,
		                    ctx->stream);
	}
	if (e != cudaSuccess) {
		if (col->raw) {
			cudaFreeAsync(col->raw, ctx->stream);
		}
		if (col->d_desc) {
			cudaFreeAsync(col->d_desc, ctx->stream);
		}
		delete col;
		return fail(ctx, e == cudaErrorMemoryAllocation ? ADACOM_ERR_NOMEM : ADACOM_ERR_CUDA,
		            std::string("adacom_column_create: ") + cudaGetErrorString(e));
	}
	st = build_tiles(col);
	if (st != ADACOM_OK) {
		if (col->raw) {
			cudaFreeAsync(col->raw, ctx->stream);
		}
		if (col->d_desc) {
			cudaFreeAsync(col->d_desc, ctx->stream);
		}
		delete col;
		return st;
	}
	*out = col;
	return ADACOM_OK;
}

extern "C" void adacom_column_destroy(adacom_column *col) {
	if (!col) {
		return;
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	cudaSetDevice(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	for (auto &x : col->extents) {
		if (x.base) {
			cudaFreeAsync(x.base, ctx->stream);
		}
	}
	for (auto &x : col->spares) {
		if (x.base) {
			cudaFreeAsync(x.base, ctx->stream);
		}
	}

	// stream-ordered frees: no device-wide synchronisation (other contexts keep their copies in flight)
	for (void *p : {(void *)col->raw, (void *)col->d_desc, (void *)col->d_tiles, (void *)col->d_vec_seg}) {
		if (p) {
			cudaFreeAsync(p, ctx->stream);
		}
	}
	cudaStreamSynchronize(ctx->stream);
	delete col;
}

extern "C" uint64_t adacom_column_num_values(const adacom_column *col) {
	return col ? col->n_values : 0;
}

extern "C" uint64_t adacom_column_num_segments(const adacom_column *col) {
	return col ? col->n_segs : 0;
}

extern "C" void *adacom_column_raw_device_ptr(adacom_column *col) {
	return col ? (void *)col->raw : nullptr;
}

extern "C" void *adacom_column_packed_device_ptr(adacom_column *col) {
	if (!col || col->extents.empty()) {
		return nullptr;
	}
	return (void *)col->extents[col->latest].base;
}

extern "C" adacom_status adacom_column_upload_raw(adacom_column *col, const void *host_values, uint64_t first_value,
                                                  uint64_t n) {
	if (!col || (!host_values && n) || first_value + n > col->n_values) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_column_upload_raw: bad range");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	adacom_status st = ensure_raw(col);
	if (st != ADACOM_OK) {
		return st;
	}
	CK(ctx, cudaMemcpyAsync(col->raw + first_value * col->t.size, host_values, n * col->t.size, cudaMemcpyHostToDevice,
	                        ctx->stream));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_column_release_raw(adacom_column *col) {
	if (!col) {
		return ADACOM_ERR_INVALID;
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	adacom_status st = sync_mirror(col);
	if (st != ADACOM_OK) {
		return st;
	}
	for (const SegDesc &d : col->h_desc) {
		if (d.count && !(d.state & kSegRepacked)) {
			return fail(ctx, ADACOM_ERR_STATE, "adacom_column_release_raw: a segment still lives in the raw arena");
		}
	}
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	CK(ctx, cudaFreeAsync(col->raw, ctx->stream));
	col->raw = nullptr;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_column_from_packed(adacom_ctx *ctx, adacom_ptype type, uint64_t n_segments,
                                                   const adacom_segment_meta *metas, const uint64_t *host_words,
                                                   uint64_t total_words, adacom_column **out) {
	if (!ctx || !out || (n_segments && !metas) || (total_words && !host_words)) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_column_from_packed: NULL argument");
	}
	*out = nullptr;
	CtxLock lock(ctx);
	CK(ctx, cudaSetDevice(ctx->device));
	adacom_column *col = nullptr;
	adacom_status st = column_alloc(ctx, type, n_segments, &col);
	if (st != ADACOM_OK) {
		return st;
	}
	Extent ext;
	ext.words = total_words;
	ext.capacity = total_words;
	ext.seg_end = n_segments;
	cudaError_t e = cudaMallocAsync((void **)&ext.base, total_words * 8 + kArenaPad, ctx->stream);
	if (e == cudaSuccess) {
		e = cudaMallocAsync((void **)&col->d_desc, std::max<uint64_t>(n_segments, 1) * sizeof(SegDesc), ctx->stream);
	}
	if (e != cudaSuccess) {
		delete col;
		return fail(ctx, ADACOM_ERR_NOMEM, std::string("adacom_column_from_packed: ") + cudaGetErrorString(e));
	}
	// the caller's metas become descriptors the kernels trust: same checks as a checkpoint image gets (restore_locked)
	auto bad = [&](uint64_t g, const char *what) -> adacom_status {
		cudaFreeAsync(ext.base, ctx->stream);
		cudaFreeAsync(col->d_desc, ctx->stream);
		col->d_desc = nullptr;
		delete col;
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_column_from_packed: segment " + std::to_string(g) + ": " + what);
	};
	const uint32_t slot_w = 8u * (uint32_t)col->t.size;
	uint64_t row = 0;
	for (uint64_t g = 0; g < n_segments; g++) {
		const adacom_segment_meta &m = metas[g];
		SegDesc &d = col->h_desc[g];
		memset(&d, 0, sizeof(d));
		bool repacked = m.state & ADACOM_SEG_REPACKED;
		if (m.state & ~(ADACOM_SEG_COMPACTED | ADACOM_SEG_REPACKED | ADACOM_SEG_ANALYZED)) {
			return bad(g, "unknown state bits");
		}
		if (m.count && !repacked) {
			return bad(g, "not repacked (no packed image)");
		}
		if (repacked && !(m.state & ADACOM_SEG_COMPACTED)) {
			return bad(g, "repacked but not compacted");
		}
		if (repacked && (m.width == 0 || m.width >= slot_w)) {
			return bad(g, "width of a repacked payload must be in [1, 8 * sizeof(T))");
		}
		if (!repacked && m.width != slot_w && m.width != 0) {
			return bad(g, "width of an empty segment must be the slot width");
		}
		if (m.count > m.capacity) {
			return bad(g, "count exceeds capacity");
		}
		if (repacked && (m.word_offset % kSegAlignWords != 0 ||
		                 m.word_offset + (((uint64_t)m.count * m.width + 63) >> 6) > total_words)) {
			return bad(g, "bad word_offset");
		}
		if (m.count * (uint64_t)col->t.size > kBlockSize || (uint64_t)m.capacity * col->t.size > kBlockSize) {
			return bad(g, "exceeds one block");
		}
		d.row_start = row;
		d.word_offset = m.word_offset;
		d.packed_ptr = repacked ? (uint64_t)(ext.base + m.word_offset) : 0;
		d.min_factor = m.min_factor;
		d.max_factor = m.max_factor;
		d.count = (uint32_t)m.count;
		d.capacity = m.capacity;
		d.extent = 0;
		d.width = repacked ? m.width : (uint8_t)slot_w;
		d.state = m.state;
		ext.live += repacked ? 1 : 0;
		row += m.count;
	}
	col->n_values = row;
	col->extents.push_back(ext);
	col->latest = col->extents.size() - 1;
	if (total_words) {
		CK(ctx, cudaMemcpyAsync(ext.base, host_words, total_words * 8, cudaMemcpyHostToDevice, ctx->stream));
	}
	if (n_segments) {
		CK(ctx, cudaMemcpyAsync(col->d_desc, col->h_desc.data(), n_segments * sizeof(SegDesc), cudaMemcpyHostToDevice,
		                        ctx->stream));
	}
	st = build_tiles(col);
	if (st != ADACOM_OK) {
		return st;
	}
	*out = col;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_column_fill_splitmix(adacom_column *col, uint64_t seed, uint64_t first_index,
                                                     uint64_t mask) {
	if (!col) {
		return ADACOM_ERR_INVALID;
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	adacom_status st = ensure_raw(col);
	if (st != ADACOM_OK) {
		return st;
	}
	LAUNCH(ctx, launch_fill_splitmix(col->t, col->raw, col->n_values, seed, first_index, mask, ctx->stream));
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// analyze / encode
// ---------------------------------------------------------------------------------------------------

extern "C" adacom_status adacom_analyze(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                        const uint64_t *validity, int validity_on_device) {
	if (!range_ok(col, seg_begin, seg_end)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_analyze: bad segment range");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (!col->raw) {
		return fail(ctx, ADACOM_ERR_STATE, "adacom_analyze: the raw arena was released");
	}
	uint64_t *d_validity = nullptr;
	bool temp = false;
	if (validity) {
		if (validity_on_device) {
			d_validity = const_cast<uint64_t *>(validity);
		} else {
			size_t bytes = ((col->n_values + 63) / 64) * 8;
			CK(ctx, cudaMallocAsync((void **)&d_validity, bytes ? bytes : 8, ctx->stream));
			CK(ctx, cudaMemcpyAsync(d_validity, validity, bytes, cudaMemcpyHostToDevice, ctx->stream));
			temp = true;
		}
	}
	LAUNCH(ctx, launch_analyze(col->t, col->d_desc, (uint32_t)seg_begin, (uint32_t)(seg_end - seg_begin), col->raw,
	                           d_validity, (flags & ADACOM_ZERO_EXT_MINMAX) ? 1 : 0, ctx->stream));
	if (temp) {
		CK(ctx, cudaFreeAsync(d_validity, ctx->stream));
	}
	col->mirror_dirty = true;
	return ADACOM_OK;
}

// the encode of the segments [seg_begin, seg_begin + n) or, with d_list, of the n listed segments whose 2048-value
// tiles are d_jobs[0 .. n_jobs) (the caller holds the context lock)
static adacom_status encode_locked(adacom_column *col, uint64_t seg_begin, uint32_t n, int flags, const uint32_t *d_list,
                                   const Job *d_jobs, uint32_t n_jobs) {
	adacom_ctx *ctx = col->ctx;
	const uint64_t seg_end = d_list ? seg_begin : seg_begin + n;  // a list owns no segment range (extent retirement)
	// extent id = the first dead slot (retired, or created without any payload): no Repacked descriptor carries its id
	// any more (retire follows the descriptor reset in stream order; gc_extents counted none), so the table stays as
	// small as the number of live extents however many encodes a column sees
	uint32_t extent_id = (uint32_t)col->extents.size();
	for (size_t i = 0; i < col->extents.size(); i++) {
		if (col->extents[i].base == nullptr) {
			extent_id = (uint32_t)i;
			break;
		}
	}
	// 1. widths, repack decisions and payload offsets, all on the device
	const unsigned long long ticket = ++ctx->ticket;
	LAUNCH(ctx, launch_layout(col->t, col->d_desc, (uint32_t)seg_begin, n, flags, extent_id, ctx->d_totals, ticket,
	                          col->d_parts, (!d_list && seg_begin == 0 && seg_end == col->n_segs) ? 1 : 0,
	                          ctx->d_layout_scratch, ctx->stream, d_list));
	// From here on the descriptors of the call are rewritten (Compacted | Repacked, packed_ptr = 0, extent = extent_id).
	// If anything below fails -- the extent allocation running out of memory is an expected path -- they are put back
	// (launch_revert_layout) before the error is returned: no segment is left "repacked" without a payload, and the next
	// encode, which reuses extent_id, cannot bind such a segment into an unrelated extent.
	auto rollback = [&](adacom_status st, uint64_t *allocated) -> adacom_status {
		const std::string msg = ctx->err;
		cudaGetLastError();  // clear the non-sticky error of the failed call
		cudaError_t r = launch_revert_layout(col->d_desc, (uint32_t)seg_begin, n, extent_id, col->d_parts, col->raw_parts,
		                                     ctx->stream, d_list);
		ctx->launches++;
		if (allocated) {
			cudaFreeAsync(allocated, ctx->stream);
		}
		col->mirror_dirty = true;
		ctx->err = msg + (r == cudaSuccess ? " (descriptors rolled back)" : std::string(" (rollback failed: ") + cudaGetErrorString(r) + ")");
		return st;
	};
	// the only host round trip of an encode: 8 bytes that size the new extent, polled from mapped memory
	{
		volatile unsigned long long *t = ctx->h_totals;
		uint32_t spins = 0;
		while (t[1] != ticket) {
			if ((++spins & 0x3ffu) == 0) {
				cudaError_t q = cudaStreamQuery(ctx->stream);
				if (q != cudaSuccess && q != cudaErrorNotReady) {
					return rollback(fail(ctx, ADACOM_ERR_CUDA, std::string("adacom_encode: layout: ") + cudaGetErrorString(q)), nullptr);
				}
				if (q == cudaSuccess && t[1] != ticket) {
					return rollback(fail(ctx, ADACOM_ERR_CUDA, "adacom_encode: layout kernel finished without publishing its total"),
					                nullptr);
				}
			}
		}
		std::atomic_thread_fence(std::memory_order_acquire);
	}
	Extent ext;
	ext.words = ctx->h_totals[0];
	ext.seg_begin = d_list ? ~0ull : seg_begin;  // a list extent is never retired by range (only by gc_extents)
	ext.seg_end = d_list ? ~0ull : seg_end;
	if (ext.words) {
		// 2. a fresh extent: scans keep reading the old payloads until the descriptors are re-bound.
		//    Prefer a retired extent of this column (smallest that fits) over an allocator call.
		int pick = -1;
		for (size_t i = 0; i < col->spares.size(); i++) {
			if (col->spares[i].capacity >= ext.words &&
			    (pick < 0 || col->spares[i].capacity < col->spares[(size_t)pick].capacity)) {
				pick = (int)i;
			}
		}
		uint64_t *fresh = nullptr;
		if (pick >= 0) {
			ext.base = col->spares[(size_t)pick].base;
			ext.capacity = col->spares[(size_t)pick].capacity;
		} else {
			ext.capacity = ext.words + ext.words / 32 + 64;
			// ADACOM_FAULT_ENCODE_ALLOC (test hook, tests/test_gpu_edges.py): behave as if the allocator had run dry
			const char *fault = getenv("ADACOM_FAULT_ENCODE_ALLOC");
			cudaError_t e = (fault && fault[0] == '1') ? cudaErrorMemoryAllocation
			                                           : cudaMallocAsync((void **)&ext.base, ext.capacity * 8 + kArenaPad, ctx->stream);
			if (e != cudaSuccess) {
				return rollback(fail(ctx, e == cudaErrorMemoryAllocation ? ADACOM_ERR_NOMEM : ADACOM_ERR_CUDA,
				                     std::string("adacom_encode: packed extent of ") + std::to_string(ext.capacity * 8) +
				                         " bytes: " + cudaGetErrorString(e)),
				                nullptr);
			}
			fresh = ext.base;
		}
		// 3. pack, then publish packed_ptr = base + word_offset (scans launched later on this stream see it)
		cudaError_t e = launch_encode(col->t, col->d_desc, d_jobs, n_jobs, col->raw, extent_id, ext.base, ctx->sm_count, ctx->stream);
		ctx->launches++;
		if (e == cudaSuccess) {
			e = launch_bind_extent(col->d_desc, (uint32_t)seg_begin, n, ext.base, extent_id, ctx->stream, d_list);
			ctx->launches++;
		}
		if (e != cudaSuccess) {
			// a spare stays a spare; a fresh allocation goes back to the pool
			return rollback(fail(ctx, ADACOM_ERR_CUDA, std::string("adacom_encode: pack / bind launch: ") + cudaGetErrorString(e)), fresh);
		}
		if (pick >= 0) {
			col->spares.erase(col->spares.begin() + pick);
		}
	}
	if (extent_id == col->extents.size()) {
		col->extents.push_back(ext);
	} else {
		col->extents[extent_id] = ext;
	}
	col->latest = extent_id;
	col->mirror_dirty = true;
	col->epoch++;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_encode(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags) {
	if (!range_ok(col, seg_begin, seg_end)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_encode: bad segment range");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (!col->raw) {
		return fail(ctx, ADACOM_ERR_STATE, "adacom_encode: the raw arena was released");
	}
	const uint32_t n = (uint32_t)(seg_end - seg_begin);
	if (n == 0) {
		return ADACOM_OK;
	}
	return encode_locked(col, seg_begin, n, flags, nullptr, col->d_tiles + col->tile_first[seg_begin],
	                     col->tile_first[seg_end] - col->tile_first[seg_begin]);
}

// ---- segment-list forms: what the access-frequency policy needs (ColumnSegmentCatalog::CompactSegments,
// src/catalog/catalog_entry/column_segment_catalog.cpp:64-116: the coldest 90 % of the segments, in read-count order,
// are Compact()ed one by one and the rest Uncompact()ed -- arbitrary subsets, not ranges).  One launch set per call.

namespace {

struct SegList {
	uint32_t *d_list = nullptr;  // device copy of the segment ids, the tile jobs behind it
	Job *d_jobs = nullptr;
	uint32_t n_jobs = 0;
};

adacom_status stage_seg_list(adacom_column *col, uint64_t n, const uint32_t *segs, const char *who, SegList &out) {
	adacom_ctx *ctx = col->ctx;
	std::vector<Job> jobs;
	for (uint64_t i = 0; i < n; i++) {
		if (segs[i] >= col->n_segs || (i && segs[i] <= segs[i - 1])) {
			return fail(ctx, ADACOM_ERR_INVALID, std::string(who) + ": segment ids must be strictly ascending and in range");
		}
		for (uint32_t t = col->tile_first[segs[i]]; t < col->tile_first[segs[i] + 1]; t++) {
			jobs.push_back({segs[i], t - col->tile_first[segs[i]]});
		}
	}
	const size_t list_bytes = ((size_t)n * sizeof(uint32_t) + 15) & ~(size_t)15;
	uint8_t *buf = nullptr;
	CK(ctx, cudaMallocAsync((void **)&buf, list_bytes + (jobs.size() + 1) * sizeof(Job), ctx->stream));
	CK(ctx, cudaMemcpyAsync(buf, segs, n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
	if (!jobs.empty()) {
		CK(ctx, cudaMemcpyAsync(buf + list_bytes, jobs.data(), jobs.size() * sizeof(Job), cudaMemcpyHostToDevice, ctx->stream));
	}
	// both sources are pageable: staged before the copies return
	out.d_list = (uint32_t *)buf;
	out.d_jobs = (Job *)(buf + list_bytes);
	out.n_jobs = (uint32_t)jobs.size();
	return ADACOM_OK;
}

}  // namespace

extern "C" adacom_status adacom_analyze_list(adacom_column *col, uint64_t n, const uint32_t *segs, int flags,
                                             const uint64_t *device_validity) {
	if (!col || (n && !segs) || n >= (1ull << 32)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_analyze_list: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (!col->raw) {
		return fail(ctx, ADACOM_ERR_STATE, "adacom_analyze_list: the raw arena was released");
	}
	if (n == 0) {
		return ADACOM_OK;
	}
	SegList sl;
	adacom_status st = stage_seg_list(col, n, segs, "adacom_analyze_list", sl);
	if (st != ADACOM_OK) {
		return st;
	}
	LAUNCH(ctx, launch_analyze(col->t, col->d_desc, 0, (uint32_t)n, col->raw, device_validity,
	                           (flags & ADACOM_ZERO_EXT_MINMAX) ? 1 : 0, ctx->stream, sl.d_list));
	CK(ctx, cudaFreeAsync(sl.d_list, ctx->stream));
	col->mirror_dirty = true;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_encode_list(adacom_column *col, uint64_t n, const uint32_t *segs, int flags) {
	if (!col || (n && !segs) || n >= (1ull << 32)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_encode_list: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (!col->raw) {
		return fail(ctx, ADACOM_ERR_STATE, "adacom_encode_list: the raw arena was released");
	}
	if (n == 0) {
		return ADACOM_OK;
	}
	SegList sl;
	adacom_status st = stage_seg_list(col, n, segs, "adacom_encode_list", sl);
	if (st != ADACOM_OK) {
		return st;
	}
	st = encode_locked(col, 0, (uint32_t)n, flags, sl.d_list, sl.d_jobs, sl.n_jobs);
	cudaFreeAsync(sl.d_list, ctx->stream);
	return st;
}

extern "C" adacom_status adacom_uncompact_list(adacom_column *col, uint64_t n, const uint32_t *segs, int flags) {
	if (!col || (n && !segs) || n >= (1ull << 32)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_uncompact_list: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (n == 0) {
		return ADACOM_OK;
	}
	adacom_status st = ensure_raw(col);
	if (st != ADACOM_OK) {
		return st;
	}
	SegList sl;
	st = stage_seg_list(col, n, segs, "adacom_uncompact_list", sl);
	if (st != ADACOM_OK) {
		return st;
	}
	DecodeArgs a;
	a.desc = col->d_desc;
	a.jobs = sl.d_jobs;
	a.n_jobs = sl.n_jobs;
	a.raw = col->raw;
	a.literal = (flags & ADACOM_LITERAL) ? 1 : 0;
	a.uncompact = 1;
	a.out = col->raw;
	a.out_row0 = 0;
	LAUNCH(ctx, launch_decode(col->t, kModeFlat, a, ctx->sm_count, ctx->stream));
	LAUNCH(ctx, launch_reset_uncompacted(col->d_desc, 0, (uint32_t)n, col->t.size, col->d_parts, col->raw_parts, ctx->stream,
	                                     sl.d_list));
	CK(ctx, cudaFreeAsync(sl.d_list, ctx->stream));
	// dead extents are reclaimed when the host mirror is refreshed (gc_extents)
	col->mirror_dirty = true;
	col->epoch++;
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// metadata / accounting
// ---------------------------------------------------------------------------------------------------

extern "C" adacom_status adacom_column_get_meta(adacom_column *col, uint64_t seg_begin, uint64_t seg_end,
                                                adacom_segment_meta *out) {
	if (!range_ok(col, seg_begin, seg_end) || (!out && seg_end > seg_begin)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_column_get_meta: bad arguments");
	}
	CtxLock lock(col->ctx);
	adacom_status st = sync_mirror(col);
	if (st != ADACOM_OK) {
		return st;
	}
	for (uint64_t g = seg_begin; g < seg_end; g++) {
		fill_meta(col, g, out + (g - seg_begin));
	}
	return ADACOM_OK;
}

extern "C" adacom_status adacom_column_data_size(adacom_column *col, uint64_t seg_begin, uint64_t seg_end,
                                                 uint64_t *out) {
	if (!range_ok(col, seg_begin, seg_end) || !out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_column_data_size: bad arguments");
	}
	CtxLock lock(col->ctx);
	adacom_status st = sync_mirror(col);
	if (st != ADACOM_OK) {
		return st;
	}
	uint64_t total = 0;
	adacom_segment_meta m;
	for (uint64_t g = seg_begin; g < seg_end; g++) {
		fill_meta(col, g, &m);
		total += m.size_in_bytes;
	}
	*out = total;
	return ADACOM_OK;
}

extern "C" uint64_t adacom_column_packed_words(const adacom_column *col) {
	if (!col || col->extents.empty()) {
		return 0;
	}
	return col->extents[col->latest].words;
}

extern "C" adacom_status adacom_column_download_packed(adacom_column *col, uint64_t first_word, uint64_t n_words,
                                                       uint64_t *host_out) {
	if (!col || col->extents.empty() || (!host_out && n_words) || first_word + n_words > col->extents[col->latest].words) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_column_download_packed: bad range");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (n_words) {
		CK(ctx, cudaMemcpyAsync(host_out, col->extents[col->latest].base + first_word, n_words * 8, cudaMemcpyDeviceToHost,
		                        ctx->stream));
	}
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_segment_serialize(adacom_column *col, uint64_t seg, uint8_t *host_out,
                                                  uint64_t capacity, uint64_t *bytes_out) {
	if (!col || seg >= col->n_segs || !bytes_out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_segment_serialize: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	adacom_status st = sync_mirror(col);
	if (st != ADACOM_OK) {
		return st;
	}
	const SegDesc &d = col->h_desc[seg];
	const uint64_t bits = seg_bits(d, col->t.size);
	const uint64_t data_bytes = 8 * ((bits + 63) >> 6);
	*bytes_out = 9 + data_bytes;
	if (!host_out || capacity < 9 + data_bytes) {
		return ADACOM_OK;  // size query
	}
	memcpy(host_out, &bits, 8);  // int_vector<0>::write_header: m_size, then the width byte
	host_out[8] = d.width;
	memset(host_out + 9, 0, data_bytes);
	if (d.state & kSegRepacked) {
		CK(ctx, cudaMemcpyAsync(host_out + 9, (const void *)d.packed_ptr, data_bytes, cudaMemcpyDeviceToHost, ctx->stream));
	} else if (col->raw && d.count) {
		// slot-width vector: the first `count` slots are the raw values, the rest was never written
		CK(ctx, cudaMemcpyAsync(host_out + 9, col->raw + d.row_start * col->t.size, (size_t)d.count * col->t.size,
		                        cudaMemcpyDeviceToHost, ctx->stream));
	}
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// decode / scan
// ---------------------------------------------------------------------------------------------------

static uint64_t range_rows(const adacom_column *col, uint64_t b, uint64_t e) {
	if (b == e) {
		return 0;
	}
	const SegDesc &last = col->h_desc[e - 1];
	return last.row_start + last.count - col->h_desc[b].row_start;
}

extern "C" adacom_status adacom_decode(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags, void *out,
                                       int out_on_device) {
	if (!range_ok(col, seg_begin, seg_end) || !out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_decode: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	const uint64_t rows = range_rows(col, seg_begin, seg_end);
	if (rows == 0) {
		return ADACOM_OK;
	}
	void *d_out = out;
	if (!out_on_device) {
		CK(ctx, cudaMallocAsync(&d_out, rows * col->t.size, ctx->stream));
	}
	DecodeArgs a;
	a.desc = col->d_desc;
	a.jobs = col->d_tiles + col->tile_first[seg_begin];
	a.n_jobs = col->tile_first[seg_end] - col->tile_first[seg_begin];
	a.raw = col->raw;
	a.literal = (flags & ADACOM_LITERAL) ? 1 : 0;
	a.out = d_out;
	a.out_row0 = col->h_desc[seg_begin].row_start;
	LAUNCH(ctx, launch_decode(col->t, kModeFlat, a, ctx->sm_count, ctx->stream));
	if (!out_on_device) {
		CK(ctx, cudaMemcpyAsync(out, d_out, rows * col->t.size, cudaMemcpyDeviceToHost, ctx->stream));
		CK(ctx, cudaFreeAsync(d_out, ctx->stream));
		CK(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_partial(adacom_column *col, uint64_t seg, uint64_t start, uint64_t scan_count,
                                             int flags, void *result, uint64_t result_offset, int result_on_device) {
	if (!col || seg >= col->n_segs || (!result && scan_count)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_scan_partial: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	if (start + scan_count > col->h_desc[seg].count) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_partial: scan runs past the end of the segment");
	}
	CtxLock lock(ctx);
	if (scan_count == 0) {
		return ADACOM_OK;
	}
	const int literal = (flags & ADACOM_LITERAL) ? 1 : 0;
	if (result_on_device) {
		LAUNCH(ctx, launch_scan_range(col->t, col->d_desc, (uint32_t)seg, col->raw, start, scan_count, literal, result,
		                              result_offset, ctx->stream));
		return ADACOM_OK;
	}
	const size_t bytes = scan_count * col->t.size;
	if (bytes <= ((size_t)1 << 20)) {
		// the per-vector pull of the drop-in path: no allocator call, the values come back through the context's pinned
		// staging pair (a D2H copy into pageable memory is staged by the driver anyway, and slower)
		adacom_status st = ensure_stage(ctx, bytes);
		if (st != ADACOM_OK) {
			return st;
		}
		LAUNCH(ctx, launch_scan_range(col->t, col->d_desc, (uint32_t)seg, col->raw, start, scan_count, literal, ctx->d_stage, 0,
		                              ctx->stream));
		CK(ctx, cudaMemcpyAsync(ctx->h_stage, ctx->d_stage, bytes, cudaMemcpyDeviceToHost, ctx->stream));
		CK(ctx, cudaStreamSynchronize(ctx->stream));
		memcpy((uint8_t *)result + result_offset * col->t.size, ctx->h_stage, bytes);
		return ADACOM_OK;
	}
	void *d_out = nullptr;
	CK(ctx, cudaMallocAsync(&d_out, bytes, ctx->stream));
	LAUNCH(ctx, launch_scan_range(col->t, col->d_desc, (uint32_t)seg, col->raw, start, scan_count, literal, d_out, 0,
	                              ctx->stream));
	CK(ctx, cudaMemcpyAsync((uint8_t *)result + result_offset * col->t.size, d_out, bytes, cudaMemcpyDeviceToHost,
	                        ctx->stream));
	CK(ctx, cudaFreeAsync(d_out, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_uncompact(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags) {
	if (!range_ok(col, seg_begin, seg_end)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_uncompact: bad segment range");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (seg_begin == seg_end) {
		return ADACOM_OK;
	}
	adacom_status st = ensure_raw(col);
	if (st != ADACOM_OK) {
		return st;
	}
	// fully asynchronous: decode the compacted segments of the range back into the raw arena (the kernel
	// skips the others), then flip their descriptors on the device.  Dead extents are reclaimed the next time
	// the host mirror is refreshed.
	DecodeArgs a;
	a.desc = col->d_desc;
	a.jobs = col->d_tiles + col->tile_first[seg_begin];
	a.n_jobs = col->tile_first[seg_end] - col->tile_first[seg_begin];
	a.raw = col->raw;
	a.literal = (flags & ADACOM_LITERAL) ? 1 : 0;
	a.uncompact = 1;
	a.out = col->raw;
	a.out_row0 = 0;
	LAUNCH(ctx, launch_decode(col->t, kModeFlat, a, ctx->sm_count, ctx->stream));
	LAUNCH(ctx, launch_reset_uncompacted(col->d_desc, (uint32_t)seg_begin, (uint32_t)(seg_end - seg_begin), col->t.size,
	                                     col->d_parts, col->raw_parts, ctx->stream));
	// extents created by encode calls whose range lies inside [seg_begin, seg_end) are dead now (stream order
	// makes the reuse safe); partially covered ones are reclaimed when the mirror is refreshed
	for (auto &x : col->extents) {
		if (x.base && x.seg_begin >= seg_begin && x.seg_end <= seg_end) {
			st = retire_extent(col, x);
			if (st != ADACOM_OK) {
				return st;
			}
		}
	}
	col->mirror_dirty = true;
	col->epoch++;
	return ADACOM_OK;
}

// The filter-words kernel (warp per 2048-row tile, field domain) serves a segment range when every segment is
// field-comparable or cheap per value: always for the 1/2/4-byte types; for the 8-byte types every live segment must be
// re-packed, and a payload wider than 32 bits must sit on a frame of reference that cannot wrap inside T.
static adacom_status filter_words_serves(adacom_column *col, uint64_t b, uint64_t e, bool *serves) {
	*serves = !getenv("ADACOM_FILTER_LEGACY");
	if (!*serves || col->t.size != 8) {
		return ADACOM_OK;
	}
	adacom_status st = sync_mirror(col);
	if (st != ADACOM_OK) {
		return st;
	}
	static const bool wide_legacy = getenv("ADACOM_FILTER_WIDE_LEGACY") != nullptr;  // A/B: wide payloads on decode_kernel
	const bool sg = col->t.is_signed;
	bool ok = true;
	for (uint64_t g = b; g < e && ok; g++) {
		const SegDesc &d = col->h_desc[g];
		if (d.count == 0) {
			continue;
		}
		ok = (d.state & kSegRepacked) != 0;
		if (ok && d.width > 32) {
			// (the kernel's per-value path for a wrapping frame reads global memory lane by lane: slower than
			// decode_kernel<filter>)
			const uint64_t mn = (d.min_factor != ~0ull) ? d.min_factor : 0ull;
			const uint64_t range = d.width < 64 ? (1ull << d.width) - 1 : ~0ull, top = mn + range;
			const bool linear = top >= mn && (sg ? (int64_t)mn <= (int64_t)top : true);
			ok = !wide_legacy && linear;
		}
	}
	*serves = ok;
	return ADACOM_OK;
}

static adacom_status run_agg(adacom_column *col, uint64_t b, uint64_t e, int flags, DecodeMode mode, uint32_t op_mask,
                             uint64_t constant, uint32_t *bitmap32, Agg *device_out) {
	adacom_ctx *ctx = col->ctx;
	CK(ctx, cudaMemsetAsync(device_out, 0, sizeof(Agg), ctx->stream));
	if (b == e) {
		return ADACOM_OK;
	}
	if (mode == kModeSum) {
		ScanSumArgs f;
		f.desc = col->d_desc;
		f.parts = col->d_parts;
		f.seg_b = (uint32_t)b;
		f.seg_e = (uint32_t)e;
		f.raw = col->raw;
		f.literal = (flags & ADACOM_LITERAL) ? 1 : 0;
		f.type_size = col->t.size;
		f.is_signed = col->t.is_signed ? 1 : 0;
		f.agg = device_out;
		LAUNCH(ctx, launch_scan_sum_fields(f, ctx->sm_count, ctx->stream));
		return ADACOM_OK;
	}
	bool words_kernel = false;
	if (mode == kModeFilter) {
		// 8-byte types: slot-width segments are not field-comparable; a range that holds any of them stays on the
		// TMA-staged per-value kernel (decode_kernel<filter>).  Re-packed payloads of any width run in the field domain
		// (<= 32 bits: pred32 on the staged tile; wider: 64-bit fields, lane = row)
		adacom_status st = filter_words_serves(col, b, e, &words_kernel);
		if (st != ADACOM_OK) {
			return st;
		}
	}
	if (words_kernel) {
		// warp-per-tile filter: field domain (pred32) for re-packed payloads <= 32 bits on a frame of reference that cannot
		// wrap, per value otherwise -- decided per segment on the device
		LAUNCH(ctx, launch_filter_words(col->t, col->d_desc, col->d_tiles + col->tile_first[b],
		                                col->tile_first[e] - col->tile_first[b], col->raw, (flags & ADACOM_LITERAL) ? 1 : 0,
		                                filter_range_for_cmp(op_mask, constant, col->t.is_signed), nullptr, nullptr, bitmap32,
		                                col->h_desc[b].row_start, device_out,
		                                (col->mirror_dirty || !col->wmax) ? 8u * (uint32_t)col->t.size : col->wmax,
		                                ctx->sm_count, ctx->stream));
		return ADACOM_OK;
	}
	if (mode == kModeFilter && getenv("ADACOM_FILTER_LEGACY")) {
		// (previous generation, kept for A/B runs) field-domain filter kernel when every re-packed payload of the range is <= 32 bits wide (always true for the
		// 1/2/4-byte types; for the 8-byte types the host mirror decides, raw 8-byte segments stay on decode_kernel)
		bool narrow = col->t.size <= 4;
		if (!narrow) {
			adacom_status st = sync_mirror(col);
			if (st != ADACOM_OK) {
				return st;
			}
			narrow = true;
			for (uint64_t g = b; g < e && narrow; g++) {
				const SegDesc &d = col->h_desc[g];
				narrow = d.count == 0 || ((d.state & kSegRepacked) && d.width <= 32);
			}
		}
		if (narrow) {
			LAUNCH(ctx, launch_filter_fields(col->t, col->d_desc, (uint32_t)b, (uint32_t)(e - b), kBlockSize / (uint32_t)col->t.size,
			                                 col->raw, (flags & ADACOM_LITERAL) ? 1 : 0, op_mask, constant, bitmap32,
			                                 col->h_desc[b].row_start, device_out, ctx->sm_count, ctx->stream));
			return ADACOM_OK;
		}
	}
	DecodeArgs a;
	a.desc = col->d_desc;
	a.parts = col->d_parts;
	a.seg_b = (uint32_t)b;
	a.n_units = (uint32_t)(e - b);
	a.raw = col->raw;
	a.literal = (flags & ADACOM_LITERAL) ? 1 : 0;
	a.agg = device_out;
	a.op_mask = op_mask;
	a.constant = constant;
	a.bitmap32 = bitmap32;
	a.out_row0 = col->h_desc[b].row_start;
	LAUNCH(ctx, launch_decode(col->t, mode, a, ctx->sm_count, ctx->stream));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_sum_async(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                               adacom_agg *device_out) {
	if (!range_ok(col, seg_begin, seg_end) || !device_out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_scan_sum_async: bad arguments");
	}
	CtxLock lock(col->ctx);
	return run_agg(col, seg_begin, seg_end, flags, kModeSum, 0, 0, nullptr, (Agg *)device_out);
}

extern "C" adacom_status adacom_scan_sum(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                         adacom_agg *out) {
	if (!range_ok(col, seg_begin, seg_end) || !out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_scan_sum: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	adacom_status st = run_agg(col, seg_begin, seg_end, flags, kModeSum, 0, 0, nullptr, ctx->d_agg);
	if (st != ADACOM_OK) {
		return st;
	}
	CK(ctx, cudaMemcpyAsync(ctx->h_agg, ctx->d_agg, sizeof(Agg), cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	memcpy(out, ctx->h_agg, sizeof(Agg));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_filter(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                            adacom_cmp op, int64_t constant, adacom_agg *out, uint64_t *device_bitmap) {
	if (!range_ok(col, seg_begin, seg_end) || !out || (int)op < 0 || (int)op > 5) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_scan_filter: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (device_bitmap) {
		uint64_t rows = range_rows(col, seg_begin, seg_end);
		CK(ctx, cudaMemsetAsync(device_bitmap, 0, ((rows + 63) / 64) * 8, ctx->stream));
	}
	adacom_status st = run_agg(col, seg_begin, seg_end, flags, kModeFilter, cmp_mask_for_op((int)op), (uint64_t)constant,
	                           (uint32_t *)device_bitmap, ctx->d_agg);
	if (st != ADACOM_OK) {
		return st;
	}
	CK(ctx, cudaMemcpyAsync(ctx->h_agg, ctx->d_agg, sizeof(Agg), cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	memcpy(out, ctx->h_agg, sizeof(Agg));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_select(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                            const adacom_pred *preds, uint32_t n_preds, adacom_conj conj,
                                            const uint64_t *validity, int validity_on_device,
                                            const uint64_t *device_bitmap_in, uint64_t *device_bitmap_out,
                                            adacom_agg *out) {
	if (!range_ok(col, seg_begin, seg_end) || !out || n_preds > ADACOM_MAX_PREDS || (n_preds && !preds) ||
	    ((int)conj != ADACOM_CONJ_AND && (int)conj != ADACOM_CONJ_OR)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_scan_select: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	SelectArgs a;
	for (uint32_t p = 0; p < n_preds; p++) {
		const adacom_pred &pr = preds[p];
		if (pr.kind < ADACOM_PRED_COMPARE || pr.kind > ADACOM_PRED_IS_NOT_NULL ||
		    (pr.kind == ADACOM_PRED_COMPARE && (pr.op < 0 || pr.op > 5))) {
			return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_select: unknown predicate kind / comparison");
		}
		a.preds[p].kind = (uint32_t)pr.kind;
		a.preds[p].op_mask = pr.kind == ADACOM_PRED_COMPARE ? cmp_mask_for_op(pr.op) : 0u;
		a.preds[p].constant = (uint64_t)pr.constant;
	}
	CtxLock lock(ctx);
	const uint64_t rows = range_rows(col, seg_begin, seg_end);
	if (device_bitmap_out && rows) {
		CK(ctx, cudaMemsetAsync(device_bitmap_out, 0, ((rows + 63) / 64) * 8, ctx->stream));
	}
	CK(ctx, cudaMemsetAsync(ctx->d_agg, 0, sizeof(Agg), ctx->stream));
	uint64_t *d_validity = nullptr;
	bool on_fast = false;
	uint64_t fast_rows = ~0ull;
	if (rows) {
		if (validity) {
			if (validity_on_device) {
				d_validity = const_cast<uint64_t *>(validity);
			} else {
				// only the words the range touches travel; the kernel indexes by column row
				const uint64_t r0 = col->h_desc[seg_begin].row_start;
				const uint64_t w0 = r0 >> 6, w1 = (r0 + rows + 63) >> 6;
				uint64_t *d_part = nullptr;
				CK(ctx, cudaMallocAsync((void **)&d_part, (w1 - w0) * 8, ctx->stream));
				CK(ctx, cudaMemcpyAsync(d_part, validity + w0, (w1 - w0) * 8, cudaMemcpyHostToDevice, ctx->stream));
				d_validity = d_part - w0;
			}
		}
		a.desc = col->d_desc;
		a.jobs = col->d_tiles + col->tile_first[seg_begin];
		a.n_jobs = col->tile_first[seg_end] - col->tile_first[seg_begin];
		a.raw = col->raw;
		a.literal = (flags & ADACOM_LITERAL) ? 1 : 0;
		a.n_preds = n_preds;
		a.conj_or = conj == ADACOM_CONJ_OR ? 1u : 0u;
		a.validity = d_validity;
		a.in_bitmap = device_bitmap_in;
		a.out_bitmap32 = (uint32_t *)device_bitmap_out;
		a.out_row0 = col->h_desc[seg_begin].row_start;
		a.agg = ctx->d_agg;
		// One value range (or the complement of one) over valid rows is the filter-words kernel's shape: field-domain
		// test on the staged tile, validity and incoming selection AND-ed in word by word.  Zero children select every
		// row, NULL or not: only the SUM skips NULLs, so that form takes the fast kernel when no selection is exchanged
		// (count = rows of the range).  Everything else (IS NULL, two separate ranges) decodes per value.
		FilterRange fr;
		bool fast = !getenv("ADACOM_SELECT_LEGACY") && canonical_filter_range(a.preds, a.n_preds, a.conj_or != 0, col->t.is_signed, &fr) &&
		            (n_preds != 0 || (!device_bitmap_in && !device_bitmap_out));
		if (fast) {
			adacom_status st = filter_words_serves(col, seg_begin, seg_end, &fast);
			if (st != ADACOM_OK) {
				return st;
			}
		}
		if (fast) {
			LAUNCH(ctx, launch_filter_words(col->t, col->d_desc, a.jobs, a.n_jobs, col->raw, a.literal, fr,
			                                (const uint32_t *)d_validity, (const uint32_t *)device_bitmap_in, a.out_bitmap32,
			                                a.out_row0, ctx->d_agg,
			                                (col->mirror_dirty || !col->wmax) ? 8u * (uint32_t)col->t.size : col->wmax,
			                                ctx->sm_count, ctx->stream));
			fast_rows = n_preds == 0 ? rows : ~0ull;
			on_fast = true;
		} else {
			LAUNCH(ctx, launch_select(col->t, a, ctx->sm_count, ctx->stream));
		}
		if (validity && !validity_on_device) {
			CK(ctx, cudaFreeAsync(d_validity + (col->h_desc[seg_begin].row_start >> 6), ctx->stream));
		}
	}
	CK(ctx, cudaMemcpyAsync(ctx->h_agg, ctx->d_agg, sizeof(Agg), cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	memcpy(out, ctx->h_agg, sizeof(Agg));
	if (on_fast) {
		// every row the fast kernel counted is a valid one
		out->reserved = out->count;
		if (fast_rows != ~0ull) {
			out->count = fast_rows;
		}
	}
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_sum_valid(adacom_column *col, uint64_t seg_begin, uint64_t seg_end, int flags,
                                               const uint64_t *validity, int validity_on_device, adacom_agg *out) {
	if (!validity) {
		// mask.AllValid(): the width-specialised kernel
		adacom_status st = adacom_scan_sum(col, seg_begin, seg_end, flags, out);
		if (st == ADACOM_OK) {
			out->reserved = out->count;
		}
		return st;
	}
	return adacom_scan_select(col, seg_begin, seg_end, flags, nullptr, 0, ADACOM_CONJ_AND, validity, validity_on_device,
	                          nullptr, nullptr, out);
}

extern "C" adacom_status adacom_point_filter_batch(adacom_column *col, uint64_t n_queries, const uint32_t *seg_idx,
                                                   const int64_t *constants, int flags, uint64_t *out_count,
                                                   uint64_t *out_first_row) {
	if (!col || (n_queries && (!seg_idx || !constants || !out_count || !out_first_row)) || n_queries >= (1ull << 32)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_point_filter_batch: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (n_queries == 0) {
		return ADACOM_OK;
	}
	for (uint64_t q = 0; q < n_queries; q++) {
		if (seg_idx[q] >= col->n_segs) {
			return fail(ctx, ADACOM_ERR_INVALID, "adacom_point_filter_batch: segment index out of range");
		}
	}
	// Shared scan: when every touched segment is re-packed to a byte-aligned width (8 / 16 / 32 bits -- the reference's
	// benchmark columns pack to 16), the queries are counting-sorted by segment and each 16 KiB chunk of a segment is
	// read once per slice of kPointSlice of its queries instead of once per query.
	if (!(flags & ADACOM_LITERAL)) {
		adacom_status st = sync_mirror(col);
		if (st != ADACOM_OK) {
			return st;
		}
		std::vector<uint32_t> start(col->n_segs + 1, 0);
		bool eligible = true;
		for (uint64_t q = 0; q < n_queries && eligible; q++) {
			const SegDesc &d = col->h_desc[seg_idx[q]];
			eligible = (d.state & kSegRepacked) && (d.width == 8 || d.width == 16 || d.width == 32) && d.count > 0;
			start[seg_idx[q] + 1]++;
		}
		if (eligible) {
			for (uint64_t g = 0; g < col->n_segs; g++) {
				start[g + 1] += start[g];
			}
			// job count first, then everything is written straight into the pinned staging buffer:
			// [jobs | sorted constants | counts (0) | first rows (~0)]
			size_t n_jobs = 0;
			for (uint64_t g = 0; g < col->n_segs; g++) {
				const uint32_t k = start[g + 1] - start[g];
				if (k) {
					const SegDesc &d = col->h_desc[g];
					const uint32_t chunk_vals = 16384u * 8u / d.width;
					n_jobs += (size_t)((d.count + chunk_vals - 1) / chunk_vals) * ((k + kPointSlice - 1) / kPointSlice);
				}
			}
			if (n_jobs >= (1ull << 32)) {
				return fail(ctx, ADACOM_ERR_INVALID, "adacom_point_filter_batch: too many look-up jobs");
			}
			const size_t jb = (n_jobs * sizeof(PointJob) + 15) & ~(size_t)15;
			const size_t need = jb + 3 * n_queries * 8;
			st = ensure_stage(ctx, need);
			if (st != ADACOM_OK) {
				return st;
			}
			PointJob *jobs = (PointJob *)ctx->h_stage;
			int64_t *sorted = (int64_t *)(ctx->h_stage + jb);
			uint64_t *res = (uint64_t *)(sorted + n_queries);  // counts, then first rows
			std::vector<uint32_t> perm(n_queries), fill(start.begin(), start.end() - 1);
			for (uint64_t q = 0; q < n_queries; q++) {
				const uint32_t pos = fill[seg_idx[q]]++;
				perm[pos] = (uint32_t)q;
				sorted[pos] = constants[q];
			}
			memset(res, 0, n_queries * 8);
			memset(res + n_queries, 0xff, n_queries * 8);
			size_t j = 0;
			for (uint64_t g = 0; g < col->n_segs; g++) {
				const uint32_t lo = start[g], hi = start[g + 1];
				if (lo == hi) {
					continue;
				}
				const SegDesc &d = col->h_desc[g];
				const uint32_t chunk_vals = 16384u * 8u / d.width;
				const uint32_t chunks = (d.count + chunk_vals - 1) / chunk_vals;
				for (uint32_t c = 0; c < chunks; c++) {
					for (uint32_t q0 = lo; q0 < hi; q0 += kPointSlice) {
						jobs[j++] = {(uint32_t)g, c, q0, std::min(hi, q0 + kPointSlice)};
					}
				}
			}
			CK(ctx, cudaMemcpyAsync(ctx->d_stage, ctx->h_stage, need, cudaMemcpyHostToDevice, ctx->stream));
			int64_t *d_const = (int64_t *)(ctx->d_stage + jb);
			unsigned long long *d_count = (unsigned long long *)(d_const + n_queries);
			LAUNCH(ctx, launch_point_shared(col->d_desc, (const PointJob *)ctx->d_stage, (uint32_t)n_jobs, d_const, d_count,
			                                d_count + n_queries, col->t.size, ctx->sm_count, ctx->stream));
			CK(ctx, cudaMemcpyAsync(res, d_count, 2 * n_queries * 8, cudaMemcpyDeviceToHost, ctx->stream));
			CK(ctx, cudaStreamSynchronize(ctx->stream));
			for (uint64_t pos = 0; pos < n_queries; pos++) {
				out_count[perm[pos]] = res[pos];
				out_first_row[perm[pos]] = res[n_queries + pos];
			}
			return ADACOM_OK;
		}
	}
	// one unit per query; the kernel walks the chunks of seg_idx[q] from the descriptors (no host-built table)
	uint32_t *d_unit_seg = nullptr;
	int64_t *d_const = nullptr;
	unsigned long long *d_count = nullptr, *d_first = nullptr;
	CK(ctx, cudaMallocAsync((void **)&d_unit_seg, n_queries * 4, ctx->stream));
	CK(ctx, cudaMallocAsync((void **)&d_const, n_queries * 8, ctx->stream));
	CK(ctx, cudaMallocAsync((void **)&d_count, n_queries * 8, ctx->stream));
	CK(ctx, cudaMallocAsync((void **)&d_first, n_queries * 8, ctx->stream));
	CK(ctx, cudaMemcpyAsync(d_unit_seg, seg_idx, n_queries * 4, cudaMemcpyHostToDevice, ctx->stream));
	CK(ctx, cudaMemcpyAsync(d_const, constants, n_queries * 8, cudaMemcpyHostToDevice, ctx->stream));
	CK(ctx, cudaMemsetAsync(d_count, 0, n_queries * 8, ctx->stream));
	CK(ctx, cudaMemsetAsync(d_first, 0xff, n_queries * 8, ctx->stream));
	DecodeArgs a;
	a.desc = col->d_desc;
	a.n_units = (uint32_t)n_queries;
	a.unit_seg = d_unit_seg;
	a.parts = col->d_parts;
	a.raw = col->raw;
	a.literal = (flags & ADACOM_LITERAL) ? 1 : 0;
	a.q_const = d_const;
	a.q_count = d_count;
	a.q_first = d_first;
	LAUNCH(ctx, launch_decode(col->t, kModePoint, a, ctx->sm_count, ctx->stream));
	CK(ctx, cudaMemcpyAsync(out_count, d_count, n_queries * 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaMemcpyAsync(out_first_row, d_first, n_queries * 8, cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaFreeAsync(d_unit_seg, ctx->stream));
	CK(ctx, cudaFreeAsync(d_const, ctx->stream));
	CK(ctx, cudaFreeAsync(d_count, ctx->stream));
	CK(ctx, cudaFreeAsync(d_first, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// fused multi-column scan
// ---------------------------------------------------------------------------------------------------

static adacom_status plan_scan_columns(adacom_ctx *ctx, const adacom_scan_term *terms, uint32_t n_terms,
                                       const adacom_sum_expr *sums, uint32_t n_sums, uint64_t row_begin, uint64_t row_end,
                                       int flags, QueryArgs &q, uint32_t group_term = 0, uint32_t n_groups = 0) {
	if (n_groups && (group_term >= n_terms || n_groups > (uint32_t)kQueryMaxGroups)) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_columns_grouped: group term missing or more than 8 groups");
	}
	q.group_col = group_term;
	q.n_groups = n_groups;
	if (!terms || n_terms == 0 || n_terms > (uint32_t)kQueryMaxCols || !sums || n_sums == 0 ||
	    n_sums > (uint32_t)kQueryMaxAggs || row_begin > row_end || (row_begin % kVectorSize) != 0) {
		return fail(ctx, ADACOM_ERR_INVALID,
		            "adacom_scan_columns: 1..8 terms, 1..8 sums, row_begin a multiple of 2048 and <= row_end");
	}
	uint64_t width_bits = 0;
	for (uint32_t c = 0; c < n_terms; c++) {
		adacom_column *col = terms[c].col;
		if (!col || col->ctx != ctx || row_end > col->n_values) {
			return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_columns: column missing, of another context, or too short");
		}
		adacom_status st = sync_mirror(col);  // query planning reads the live widths
		if (st != ADACOM_OK) {
			return st;
		}
		refresh_widths(col);
		if (col->any_raw && !col->raw) {
			return fail(ctx, ADACOM_ERR_STATE, "adacom_scan_columns: the raw arena was released");
		}
		width_bits += col->wmax;
		QueryCol &qc = q.cols[c];
		qc.desc = col->d_desc;
		qc.vec_seg = col->d_vec_seg;
		qc.raw = col->raw;
		qc.n_segs = (uint32_t)col->n_segs;
		qc.type_size = (uint32_t)col->t.size;
		qc.is_signed = col->t.is_signed ? 1u : 0u;
		qc.has_pred = terms[c].has_predicate ? 1u : 0u;
		qc.lo = terms[c].lo;
		qc.hi = terms[c].hi;
	}
	for (uint32_t g = 0; g < n_sums; g++) {
		if (sums[g].n_factors > (uint32_t)kQueryMaxFactors) {
			return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_columns: at most 3 factors per sum");
		}
		QueryAgg &qa = q.aggs[g];
		qa.n_factors = sums[g].n_factors;
		for (uint32_t k = 0; k < sums[g].n_factors; k++) {
			if (sums[g].term[k] >= n_terms) {
				return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_columns: factor refers to a missing term");
			}
			qa.col[k] = sums[g].term[k];
			q.factor_mask |= 1u << sums[g].term[k];
			qa.add[k] = sums[g].add[k];
			qa.mul[k] = sums[g].mul[k];
		}
	}
	// the factor program of the dense kernel: aggregates sorted by their factor lists, so that a list which extends the
	// previous one continues its term (the products are the same numbers, only evaluated once)
	{
		std::vector<uint32_t> order(n_sums);
		for (uint32_t g = 0; g < n_sums; g++) {
			order[g] = g;
		}
		auto key = [&](uint32_t g, uint32_t k) {
			return std::make_tuple(q.aggs[g].col[k], q.aggs[g].add[k], q.aggs[g].mul[k]);
		};
		std::stable_sort(order.begin(), order.end(), [&](uint32_t l, uint32_t r) {
			const uint32_t nl = q.aggs[l].n_factors, nr = q.aggs[r].n_factors;
			for (uint32_t k = 0; k < std::min(nl, nr); k++) {
				if (key(l, k) != key(r, k)) {
					return key(l, k) < key(r, k);
				}
			}
			return nl < nr;
		});
		q.n_steps = 0;
		int prev = -1;  // aggregate whose full product the running term holds
		for (uint32_t g : order) {
			const QueryAgg &qa = q.aggs[g];
			uint32_t shared = 0;
			if (prev >= 0 && q.aggs[prev].n_factors > 0 && q.aggs[prev].n_factors <= qa.n_factors) {
				shared = q.aggs[prev].n_factors;
				for (uint32_t k = 0; k < shared; k++) {
					if (key((uint32_t)prev, k) != key(g, k)) {
						shared = 0;
						break;
					}
				}
			}
			if (qa.n_factors == shared || qa.n_factors == 0) {
				// COUNT(*) (term = 1), or the same product again: one step with x = 1
				QueryStep &st = q.steps[q.n_steps++];
				st.add = 1;
				st.mul = 0;
				st.col = 0;
				st.fresh = qa.n_factors == 0 ? 1 : 0;
				st.emit = (uint8_t)(g + 1);
			} else {
				for (uint32_t k = shared; k < qa.n_factors; k++) {
					QueryStep &st = q.steps[q.n_steps++];
					st.add = qa.add[k];
					st.mul = qa.mul[k];
					st.col = (uint16_t)qa.col[k];
					st.fresh = k == 0 ? 1 : 0;
					st.emit = k + 1 == qa.n_factors ? (uint8_t)(g + 1) : 0;
				}
			}
			prev = qa.n_factors ? (int)g : -1;
		}
	}
	// 64-bit products: bound |value| per term from the analyzed min / max of the scanned segments (type range when a
	// segment was never analyzed or mixes signs), then every prod_k (|add| + |mul| * bound) against 2^62
	long double max_prod = ldexpl(1.0L, 70);
	{
		long double bound[kQueryMaxCols];
		for (uint32_t c = 0; c < n_terms; c++) {
			const adacom_column *col = terms[c].col;
			const long double type_bound = col->t.is_signed ? ldexpl(1.0L, 8 * col->t.size - 1) : ldexpl(1.0L, 8 * col->t.size);
			long double b = 0;
			for (uint64_t g = 0; g < col->n_segs && b < type_bound; g++) {
				const SegDesc &d = col->h_desc[g];
				if (d.count == 0 || d.row_start + d.count <= row_begin || d.row_start >= row_end) {
					continue;
				}
				if (!(d.state & kSegAnalyzed) || d.min_factor == ~0ull) {
					b = type_bound;
				} else if (col->t.is_signed) {
					const int64_t lo = (int64_t)d.min_factor, hi = (int64_t)d.max_factor;
					if (lo >= 0 && hi < 0) {
						b = type_bound;  // unsigned order of sign-extended patterns: the segment mixes signs
					} else {
						b = fmaxl(b, fmaxl(fabsl((long double)lo), fabsl((long double)hi)));
					}
				} else {
					b = fmaxl(b, (long double)d.max_factor);
				}
			}
			bound[c] = fminl(b, type_bound);
		}
		bool narrow = true;
		for (uint32_t g = 0; g < n_sums && narrow; g++) {
			long double prod = 1;
			for (uint32_t k = 0; k < sums[g].n_factors; k++) {
				prod *= fabsl((long double)sums[g].add[k]) + fabsl((long double)sums[g].mul[k]) * bound[sums[g].term[k]];
			}
			narrow = prod < ldexpl(1.0L, 62);
		}
		q.narrow_products = narrow ? 1u : 0u;
		// 32-bit mode of the dense kernel: every value and every factor fits int32
		bool narrow32 = true;
		for (uint32_t c = 0; c < n_terms; c++) {
			narrow32 = narrow32 && bound[c] < ldexpl(1.0L, 31);
		}
		for (uint32_t g = 0; g < n_sums; g++) {
			for (uint32_t k = 0; k < sums[g].n_factors; k++) {
				narrow32 = narrow32 && fabsl((long double)sums[g].add[k]) +
				                               fabsl((long double)sums[g].mul[k]) * bound[sums[g].term[k]] <
				                           ldexpl(1.0L, 31);
			}
		}
		q.narrow_values = narrow32 ? 1u : 0u;
		max_prod = 0;
		for (uint32_t g = 0; g < n_sums; g++) {
			long double prod = 1;
			for (uint32_t k = 0; k < sums[g].n_factors; k++) {
				prod *= fabsl((long double)sums[g].add[k]) + fabsl((long double)sums[g].mul[k]) * bound[sums[g].term[k]];
			}
			max_prod = fmaxl(max_prod, prod);
		}
	}
	// row-block size and pipeline depth: the largest block whose worst-case staged bytes leave room for two
	// CTAs per SM (shrunk while the table offers fewer than ~4 blocks per SM)
	const uint64_t n_rows = row_end - row_begin;
	auto stage_bytes = [&](uint32_t R) -> uint64_t {
		return ((uint64_t)R * width_bits / 8 + (uint64_t)n_terms * kQueryMaxPieces * 32 + 64 + 127) & ~127ull;
	};
	uint32_t R = 0, stages = 0;
	// GROUP BY with per-thread int64 accumulators in shared memory (dense selections skip the survivor lists): allowed
	// when no thread's partial sum can leave int64 -- a thread adds one row per 32-row group of its warp, R / 256 rows
	// per block, over at most ceil(blocks / SMs) blocks -- and the accumulators fit beside two stages (two CTAs per SM:
	// 228 KiB - 2 x (1 KiB reserved + ~26 KiB static) leaves 87 KiB of dynamic shared memory each)
	q.dense_groups = 0;
	if (n_groups && q.narrow_products) {
		const uint32_t dense_smem = scan_columns_group_smem(n_groups, n_sums, 1);
		for (uint32_t r : {2048u, 4096u}) {  // lane = row in the dense path: the small block leaves room for a third CTA
			const uint64_t blocks = (n_rows + r - 1) / r;
			const uint64_t ctas = std::min<uint64_t>(std::max<uint64_t>(blocks, 1), (uint64_t)ctx->sm_count);
			const long double rows_per_thread = (long double)((blocks + ctas - 1) / ctas) * (r / (32u * kQueryConsumerWarps));
			if (!R && 2 * stage_bytes(r) + dense_smem <= 87u * 1024u && max_prod * rows_per_thread < ldexpl(1.0L, 62)) {
				R = r;
				stages = 2;
				q.dense_groups = 1;
			}
		}
	}
	const uint32_t group_smem = n_groups ? scan_columns_group_smem(n_groups, n_sums, 0) : 0u;
	// ADACOM_QUERY_STAGES (developer knob): pipeline depth per CTA; with one stage the overlap of copy and compute comes
	// from the other resident CTAs of the SM (more consumer warps per SM for the same shared memory)
	static const uint32_t forced_stages = getenv("ADACOM_QUERY_STAGES") ? (uint32_t)atoi(getenv("ADACOM_QUERY_STAGES")) : 0u;
	for (uint32_t r : {4096u, 2048u}) {  // <= kQueryMaxBlockRows: one 32-row group per consumer lane
		for (uint32_t st : {3u, 2u, 1u}) {
			if (forced_stages && st != forced_stages) {
				continue;
			}
			// + ~16 KiB static (+ the GROUP BY scratch): three or more CTAs per SM
			if (!R && st * stage_bytes(r) + group_smem <= 48u * 1024u) {
				R = r;
				stages = st;
			}
		}
	}
	for (uint32_t limit : {84u * 1024u, 196u * 1024u}) {  // two CTAs per SM, then one
		if (!R && 2 * stage_bytes(2048) + group_smem <= limit) {
			R = 2048;
			stages = 2;
		}
	}
	if (!R) {
		return fail(ctx, ADACOM_ERR_UNSUPPORTED, "adacom_scan_columns: the referenced columns are too wide to stage together");
	}
	while (R > 2048 && n_rows / R < 4ull * (uint64_t)ctx->sm_count) {
		R /= 2;
	}
	// a row block may touch at most kQueryMaxPieces consecutive segments of a column
	const int ri = R == 8192 ? 2 : (R == 4096 ? 1 : 0);
	for (uint32_t c = 0; c < n_terms; c++) {
		const adacom_column *col = terms[c].col;
		if (row_begin % R == 0 && col->max_span[ri] <= (uint32_t)kQueryMaxPieces) {
			continue;  // no aligned block of the whole column spans more (precomputed with the tile table)
		}
		for (uint64_t r0 = row_begin; r0 < row_end; r0 += R) {
			const uint64_t last = std::min<uint64_t>(r0 + R, row_end) - 1;
			uint64_t g = col->vec_seg[last / kVectorSize];
			while (g + 1 < col->n_segs && last >= col->h_desc[g].row_start + col->h_desc[g].count) {
				g++;
			}
			if (g - col->vec_seg[r0 / kVectorSize] + 1 > (uint64_t)kQueryMaxPieces) {
				return fail(ctx, ADACOM_ERR_UNSUPPORTED,
				            "adacom_scan_columns: a row block spans more than 3 segments of one column");
			}
		}
	}
	// grouped scans over field-comparable columns run warp-per-block on "uniform ranges" (scan_columns_dense_kernel): the
	// rows are cut at every segment boundary of every column, so a block sees exactly one segment per column
	q.dense_ranges = nullptr;
	if (n_groups && q.narrow_products && !getenv("ADACOM_QUERY_PIPELINED") &&
	    max_prod * ((long double)n_rows / 32.0L + 1.0L) < ldexpl(1.0L, 62)) {
		auto host_extend = [](uint64_t x, int size, bool sg) -> uint64_t {
			const unsigned shift = 64u - 8u * (unsigned)size;
			return sg ? (uint64_t)(((int64_t)(x << shift)) >> shift) : ((x << shift) >> shift);
		};
		bool ok = true;
		uint64_t win = 0;
		std::vector<uint64_t> cuts;
		cuts.push_back(row_begin);
		cuts.push_back(row_end);
		for (uint32_t c = 0; c < n_terms && ok; c++) {
			const adacom_column *col = terms[c].col;
			uint32_t wmax = 1;
			for (uint64_t g = 0; g < col->n_segs && ok; g++) {
				const SegDesc &d = col->h_desc[g];
				if (d.count == 0 || d.row_start + d.count <= row_begin || d.row_start >= row_end) {
					continue;
				}
				const uint64_t mn = d.min_factor != ~0ull ? d.min_factor : 0ull;
				const uint64_t range = d.width < 64 ? (1ull << d.width) - 1 : ~0ull, top = mn + range;
				const uint64_t e_lo = host_extend(mn, col->t.size, col->t.is_signed), e_hi = host_extend(top, col->t.size, col->t.is_signed);
				ok = (d.state & kSegRepacked) && d.width <= 32 && top >= mn && e_hi - e_lo == range &&
				     (col->t.is_signed ? (int64_t)e_lo <= (int64_t)e_hi : e_lo <= e_hi);
				wmax = std::max<uint32_t>(wmax, d.width);
				if (d.row_start > row_begin) {
					cuts.push_back(d.row_start);
				}
			}
			win += (uint64_t)kDenseBlockRows * wmax / 8 + 48;
		}
		win = (win + 127) & ~127ull;
		const uint64_t smem = 4 * win + 4ull * kQueryMaxSteps * 32 + (uint64_t)n_groups * n_sums * kDenseThreads * 8 +
		                      (uint64_t)n_groups * kDenseThreads * 4;
		if (ok && smem <= 100u * 1024u) {
			std::vector<uint64_t> key = {row_begin, row_end, n_terms};
			for (uint32_t c = 0; c < n_terms; c++) {
				key.push_back((uint64_t)(uintptr_t)terms[c].col);
				key.push_back(terms[c].col->epoch);
			}
			if (key != ctx->dense_key) {
				std::sort(cuts.begin(), cuts.end());
				cuts.erase(std::unique(cuts.begin(), cuts.end()), cuts.end());
				std::vector<DenseRange> ranges;
				std::vector<uint64_t> cursor(n_terms, 0);
				uint32_t blocks = 0;
				for (size_t i = 0; i + 1 < cuts.size(); i++) {
					DenseRange r;
					memset(&r, 0, sizeof(r));
					r.row_begin = cuts[i];
					r.rows = (uint32_t)(cuts[i + 1] - cuts[i]);
					r.first_block = blocks;
					for (uint32_t c = 0; c < n_terms; c++) {
						const adacom_column *col = terms[c].col;
						uint64_t &g = cursor[c];
						if (g == 0) {
							g = col->vec_seg[cuts[i] / kVectorSize];
						}
						while (g + 1 < col->n_segs && cuts[i] >= col->h_desc[g].row_start + col->h_desc[g].count) {
							g++;
						}
						r.seg[c] = (uint32_t)g;
					}
					blocks += (r.rows + kDenseBlockRows - 1) / kDenseBlockRows;
					ranges.push_back(r);
				}
				if (ranges.size() > ctx->dense_capacity) {
					if (ctx->d_dense_ranges) {
						CK(ctx, cudaFreeAsync(ctx->d_dense_ranges, ctx->stream));
					}
					ctx->dense_capacity = ranges.size() * 2;
					CK(ctx, cudaMallocAsync((void **)&ctx->d_dense_ranges, ctx->dense_capacity * sizeof(DenseRange), ctx->stream));
				}
				CK(ctx, cudaMemcpyAsync(ctx->d_dense_ranges, ranges.data(), ranges.size() * sizeof(DenseRange), cudaMemcpyHostToDevice,
				                        ctx->stream));
				CK(ctx, cudaStreamSynchronize(ctx->stream));  // `ranges` is a stack vector
				ctx->dense_n_ranges = (uint32_t)ranges.size();
				ctx->dense_blocks = blocks;
				ctx->dense_key = key;
			}
			q.dense_ranges = ctx->d_dense_ranges;
			q.dense_n_ranges = ctx->dense_n_ranges;
			q.dense_blocks = ctx->dense_blocks;
			q.dense_window = (uint32_t)win;
		}
	}
	// ungrouped scans run warp-per-block (scan_columns_sparse_kernel): a warp stages the predicate columns of 2048 rows
	// (<= 3 pieces per column, each rounded out to 16 bytes on both sides) in its own shared-memory window
	q.sparse_window = 0;
	if (n_groups == 0 && !getenv("ADACOM_QUERY_PIPELINED")) {
		uint64_t win = 0;
		for (uint32_t c = 0; c < n_terms; c++) {
			if (terms[c].has_predicate) {
				win += 2048ull * std::min<uint32_t>(terms[c].col->wmax ? terms[c].col->wmax : 8u * terms[c].col->t.size, 32u) / 8 +
				       (uint64_t)kQueryMaxPieces * 48;
			}
		}
		win = (win + 127) & ~127ull;
		if (std::max<uint64_t>(win, 4096) * 8 <= 200u * 1024u) {
			q.sparse_window = (uint32_t)std::max<uint64_t>(win, 4096);  // >= the survivor list of a block (2048 x u16)
		}
	}
	q.n_cols = n_terms;
	q.n_aggs = n_sums;
	q.row_begin = row_begin;
	q.row_end = row_end;
	q.block_rows = R;
	q.stages = stages;
	q.stage_bytes = (uint32_t)stage_bytes(R);
	q.literal = (flags & ADACOM_LITERAL) ? 1 : 0;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_columns_async(adacom_ctx *ctx, const adacom_scan_term *terms, uint32_t n_terms,
                                                   const adacom_sum_expr *sums, uint32_t n_sums, uint64_t row_begin,
                                                   uint64_t row_end, int flags, const uint64_t *device_bitmap_in,
                                                   uint64_t *device_bitmap_out, adacom_agg *device_out) {
	if (!ctx || !device_out) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_columns_async: NULL argument");
	}
	CtxLock lock(ctx);
	QueryArgs q;
	adacom_status st = plan_scan_columns(ctx, terms, n_terms, sums, n_sums, row_begin, row_end, flags, q);
	if (st != ADACOM_OK) {
		return st;
	}
	q.in_bitmap = (const uint32_t *)device_bitmap_in;
	q.out_bitmap = (uint32_t *)device_bitmap_out;
	q.out = (Agg *)device_out;
	CK(ctx, cudaMemsetAsync(device_out, 0, n_sums * sizeof(Agg), ctx->stream));
	LAUNCH(ctx, launch_scan_columns(q, ctx->sm_count, ctx->stream));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_columns(adacom_ctx *ctx, const adacom_scan_term *terms, uint32_t n_terms,
                                             const adacom_sum_expr *sums, uint32_t n_sums, uint64_t row_begin,
                                             uint64_t row_end, int flags, const uint64_t *device_bitmap_in,
                                             uint64_t *device_bitmap_out, adacom_agg *out) {
	if (!ctx || !out || n_sums > (uint32_t)kQueryMaxAggs) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_columns: NULL argument");
	}
	if (!ctx->d_aggs) {
		CtxLock lock(ctx);
		CK(ctx, cudaMalloc((void **)&ctx->d_aggs, kQueryMaxAggs * kQueryMaxGroups * sizeof(Agg)));
		CK(ctx, cudaHostAlloc((void **)&ctx->h_aggs, kQueryMaxAggs * kQueryMaxGroups * sizeof(Agg), cudaHostAllocDefault));
	}
	adacom_status st = adacom_scan_columns_async(ctx, terms, n_terms, sums, n_sums, row_begin, row_end, flags,
	                                             device_bitmap_in, device_bitmap_out, (adacom_agg *)ctx->d_aggs);
	if (st != ADACOM_OK) {
		return st;
	}
	CtxLock lock(ctx);
	CK(ctx, cudaMemcpyAsync(ctx->h_aggs, ctx->d_aggs, n_sums * sizeof(Agg), cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	memcpy(out, ctx->h_aggs, n_sums * sizeof(Agg));
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_columns_grouped(adacom_ctx *ctx, const adacom_scan_term *terms, uint32_t n_terms,
                                                     uint32_t group_term, uint32_t n_groups,
                                                     const adacom_sum_expr *sums, uint32_t n_sums, uint64_t row_begin,
                                                     uint64_t row_end, int flags, const uint64_t *device_bitmap_in,
                                                     uint64_t *device_bitmap_out, adacom_agg *out, int out_on_device) {
	if (!ctx || !out || n_groups == 0 || n_sums > (uint32_t)kQueryMaxAggs) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_scan_columns_grouped: NULL argument, no groups or more than 8 sums");
	}
	CtxLock lock(ctx);
	if (!ctx->d_aggs) {
		CK(ctx, cudaMalloc((void **)&ctx->d_aggs, kQueryMaxAggs * kQueryMaxGroups * sizeof(Agg)));
		CK(ctx, cudaHostAlloc((void **)&ctx->h_aggs, kQueryMaxAggs * kQueryMaxGroups * sizeof(Agg), cudaHostAllocDefault));
	}
	QueryArgs q;
	adacom_status st = plan_scan_columns(ctx, terms, n_terms, sums, n_sums, row_begin, row_end, flags, q, group_term, n_groups);
	if (st != ADACOM_OK) {
		return st;
	}
	const size_t bytes = (size_t)n_groups * n_sums * sizeof(Agg);
	q.in_bitmap = (const uint32_t *)device_bitmap_in;
	q.out_bitmap = (uint32_t *)device_bitmap_out;
	q.out = out_on_device ? (Agg *)out : ctx->d_aggs;
	CK(ctx, cudaMemsetAsync(q.out, 0, bytes, ctx->stream));
	LAUNCH(ctx, launch_scan_columns(q, ctx->sm_count, ctx->stream));
	if (!out_on_device) {
		CK(ctx, cudaMemcpyAsync(ctx->h_aggs, ctx->d_aggs, bytes, cudaMemcpyDeviceToHost, ctx->stream));
		CK(ctx, cudaStreamSynchronize(ctx->stream));
		memcpy(out, ctx->h_aggs, bytes);
	}
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// fetch_row
// ---------------------------------------------------------------------------------------------------

extern "C" adacom_status adacom_fetch_rows(adacom_column *col, uint64_t n, const uint32_t *seg_idx,
                                           const uint32_t *row_in_seg, int flags, void *out, int on_device) {
	if (!col || (n && (!seg_idx || !row_in_seg || !out))) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_fetch_rows: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (n == 0) {
		return ADACOM_OK;
	}
	const int literal = (flags & ADACOM_LITERAL) ? 1 : 0;
	if (on_device) {
		LAUNCH(ctx, launch_fetch_rows(col->t, col->d_desc, col->raw, n, seg_idx, row_in_seg, literal, out, ctx->stream));
		return ADACOM_OK;
	}
	if (n <= kSmallFetch) {
		// [segments | rows | answers] in the host-mapped page: one launch, one synchronisation
		uint32_t *h_seg = (uint32_t *)ctx->h_small, *h_row = h_seg + kSmallFetch;
		for (uint64_t q = 0; q < n; q++) {
			if (seg_idx[q] >= col->n_segs || row_in_seg[q] >= col->h_desc[seg_idx[q]].count) {
				return fail(ctx, ADACOM_ERR_INVALID, "adacom_fetch_rows: (segment, row) out of range");
			}
			h_seg[q] = seg_idx[q];
			h_row[q] = row_in_seg[q];
		}
		const uint32_t *d_seg = (const uint32_t *)ctx->d_small;
		if (n == 1) {
			// the single-row slot: the pair travels with the launch, the device reads nothing back from the host
			LAUNCH(ctx, launch_fetch_one(col->t, col->d_desc, col->raw, seg_idx[0], row_in_seg[0], literal,
			                             ctx->d_small + kSmallFetch * 8, ctx->stream));
		} else {
			LAUNCH(ctx, launch_fetch_rows(col->t, col->d_desc, col->raw, n, d_seg, d_seg + kSmallFetch, literal,
			                              ctx->d_small + kSmallFetch * 8, ctx->stream));
		}
		CK(ctx, cudaStreamSynchronize(ctx->stream));
		memcpy(out, ctx->h_small + kSmallFetch * 8, n * col->t.size);
		return ADACOM_OK;
	}
	// copied into the pinned staging buffer [segments | rows | answers] with two block copies; the (segment, row) pairs are
	// validated by the kernel (the lowest offending query index comes back through a host-mapped word): walking the
	// descriptors on the host cost more than the transfer and the kernel together
	const size_t in_bytes = (size_t)n * 8;
	adacom_status st = ensure_stage(ctx, in_bytes + (size_t)n * col->t.size);
	if (st != ADACOM_OK) {
		return st;
	}
	uint32_t *h_seg = (uint32_t *)ctx->h_stage, *h_row = h_seg + n;
	memcpy(h_seg, seg_idx, (size_t)n * 4);
	memcpy(h_row, row_in_seg, (size_t)n * 4);
	unsigned long long *h_bad = ctx->h_totals + 3, *d_bad = ctx->d_totals + 3;
	*h_bad = ~0ull;
	uint32_t *d_seg = (uint32_t *)ctx->d_stage;
	void *d_out = ctx->d_stage + in_bytes;
	CK(ctx, cudaMemcpyAsync(d_seg, h_seg, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
	LAUNCH(ctx, launch_fetch_rows(col->t, col->d_desc, col->raw, n, d_seg, d_seg + n, literal, d_out, ctx->stream,
	                              (uint32_t)col->n_segs, d_bad));
	CK(ctx, cudaMemcpyAsync(ctx->h_stage + in_bytes, d_out, n * col->t.size, cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	if (*h_bad != ~0ull) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_fetch_rows: (segment, row) out of range at query " + std::to_string(*h_bad));
	}
	memcpy(out, ctx->h_stage + in_bytes, n * col->t.size);
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// host-buffer convenience path
// ---------------------------------------------------------------------------------------------------

extern "C" adacom_status adacom_compress_host(adacom_ctx *ctx, adacom_ptype type, const void *host_values,
                                              uint64_t n_segments, const uint64_t *seg_counts, uint32_t seg_capacity,
                                              int flags, adacom_segment_meta *metas_out, uint64_t *host_words_out,
                                              uint64_t host_words_capacity, uint64_t *total_words_out,
                                              adacom_column **col_out) {
	if (!ctx || !col_out || (!host_values && n_segments)) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_compress_host: NULL argument");
	}
	adacom_column *col = nullptr;
	adacom_status st = adacom_column_create(ctx, type, n_segments, seg_counts, seg_capacity, &col);
	if (st != ADACOM_OK) {
		return st;
	}
	auto bail = [&](adacom_status s) {
		adacom_column_destroy(col);
		return s;
	};
	st = adacom_column_upload_raw(col, host_values, 0, col->n_values);
	if (st != ADACOM_OK) {
		return bail(st);
	}
	st = adacom_analyze(col, 0, n_segments, flags, nullptr, 0);
	if (st != ADACOM_OK) {
		return bail(st);
	}
	st = adacom_encode(col, 0, n_segments, flags);
	if (st != ADACOM_OK) {
		return bail(st);
	}
	const uint64_t words = adacom_column_packed_words(col);
	if (total_words_out) {
		*total_words_out = words;
	}
	if (host_words_out) {
		if (host_words_capacity < words) {
			return bail(fail(ctx, ADACOM_ERR_INVALID, "adacom_compress_host: host_words_capacity too small"));
		}
		st = adacom_column_download_packed(col, 0, words, host_words_out);
		if (st != ADACOM_OK) {
			return bail(st);
		}
	}
	if (metas_out) {
		st = adacom_column_get_meta(col, 0, n_segments, metas_out);
		if (st != ADACOM_OK) {
			return bail(st);
		}
	}
	*col_out = col;
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// checkpoint image (persisted form of a column): layout planning on the host (metadata-sized, what
// ColumnCheckpointState / PartialBlockManager do in the reference), block assembly + checksums on the device
// ---------------------------------------------------------------------------------------------------

static_assert(sizeof(adacom_image_header) == 64 && sizeof(adacom_data_pointer) == 64, "image records are 64 bytes");
static_assert(ADACOM_BLOCK_ALLOC_SIZE == kBlockAllocSize && ADACOM_BLOCK_SIZE == kBlockSize, "block geometry");

namespace {

// duckdb::Checksum over whole words (src/common/checksum.cpp:6-25); only used for the few KiB of image metadata
uint64_t meta_checksum(const adacom_image_header &h, const std::vector<adacom_data_pointer> &ptrs) {
	adacom_image_header hz = h;
	hz.meta_checksum = 0;
	uint64_t r = 5381;
	const uint64_t *p = (const uint64_t *)&hz;
	for (size_t i = 0; i < sizeof(hz) / 8; i++) {
		r ^= p[i] * 0xbf58476d1ce4e5b9ull;
	}
	p = (const uint64_t *)ptrs.data();
	for (size_t i = 0; i < ptrs.size() * sizeof(adacom_data_pointer) / 8; i++) {
		r ^= p[i] * 0xbf58476d1ce4e5b9ull;
	}
	return r;
}

struct ImagePlan {
	adacom_image_header header;
	std::vector<adacom_data_pointer> ptrs;
	std::vector<CkPiece> pieces;        // grouped by block, increasing offset inside a block
	std::vector<uint32_t> piece_begin;  // n_blocks + 1
};

// Block placement of FlushSegment + PartialBlockManager::{GetBlockAllocation, RegisterPartialBlock}
// (column_checkpoint_state.cpp:107-139, partial_block_manager.cpp:14-77) with the default limits.
struct BlockPlacer {
	static constexpr uint32_t kMaxPartial = kBlockSize / 5 * 4;  // DEFAULT_MAX_PARTIAL_BLOCK_SIZE
	std::multimap<uint32_t, std::pair<int64_t, uint32_t>> partial;  // space left -> (block id, offset in block)
	int64_t n_blocks = 0;
	void place(uint32_t segment_size, int64_t &block_id, uint32_t &offset) {
		bool found = false;
		if (segment_size <= kMaxPartial) {
			auto it = partial.lower_bound(segment_size);
			if (it != partial.end()) {
				block_id = it->second.first;
				offset = it->second.second;
				partial.erase(it);
				found = true;
			}
		}
		if (!found) {
			block_id = n_blocks++;
			offset = 0;
		}
		const uint32_t new_size = (segment_size + offset + 7u) & ~7u;  // AlignValue
		const uint32_t left = kBlockSize - new_size;
		if (left >= kBlockSize - kMaxPartial) {
			partial.insert({left, {block_id, new_size}});
		}
	}
};

adacom_status plan_image(adacom_column *col, ImagePlan &plan) {
	adacom_status st = sync_mirror(col);
	if (st != ADACOM_OK) {
		return st;
	}
	const uint32_t S = (uint32_t)col->t.size;
	plan.ptrs.assign(col->n_segs, adacom_data_pointer());
	std::vector<std::vector<CkPiece>> per_block;
	BlockPlacer placer;
	for (uint64_t g = 0; g < col->n_segs; g++) {
		const SegDesc &d = col->h_desc[g];
		adacom_data_pointer &p = plan.ptrs[g];
		memset(&p, 0, sizeof(p));
		const bool repacked = d.state & kSegRepacked;
		p.row_start = d.row_start;
		p.tuple_count = d.count;
		p.block_id = ADACOM_INVALID_BLOCK;
		p.min_factor = d.min_factor;
		p.max_factor = d.max_factor;
		p.capacity = d.capacity;
		p.compression_type = (d.state & kSegUncompressedFn) ? 1 : 10;
		p.width = d.width;
		p.state = d.state;
		p.type_size = (uint8_t)S;
		if (d.count == 0) {
			continue;  // FlushSegment returns before allocating anything (column_checkpoint_state.cpp:94-96)
		}
		p.payload_bytes = repacked ? (uint32_t)(8 * (((uint64_t)d.count * d.width + 63) >> 6)) : d.count * S;
		if (!repacked && !col->raw) {
			return fail(col->ctx, ADACOM_ERR_STATE, "adacom_column_checkpoint: a slot-width segment has no raw arena");
		}
		placer.place(p.payload_bytes, p.block_id, p.offset);
		CkPiece pc;
		pc.addr = repacked ? d.packed_ptr : (uint64_t)(col->raw + d.row_start * S);
		pc.block_off = p.offset;
		pc.bytes = p.payload_bytes;
		pc.pad_words = 0;
		pc.elem_size = repacked ? 8 : S;
		if ((size_t)p.block_id >= per_block.size()) {
			per_block.resize(p.block_id + 1);
		}
		per_block[p.block_id].push_back(pc);
	}
	plan.piece_begin.assign(per_block.size() + 1, 0);
	plan.pieces.clear();
	for (size_t b = 0; b < per_block.size(); b++) {
		plan.piece_begin[b] = (uint32_t)plan.pieces.size();
		plan.pieces.insert(plan.pieces.end(), per_block[b].begin(), per_block[b].end());
	}
	plan.piece_begin[per_block.size()] = (uint32_t)plan.pieces.size();
	adacom_image_header &h = plan.header;
	memset(&h, 0, sizeof(h));
	memcpy(h.magic, ADACOM_IMAGE_MAGIC, 8);
	h.version = ADACOM_IMAGE_VERSION;
	h.ptype = (uint32_t)col->ptype;
	h.n_segments = col->n_segs;
	h.n_blocks = per_block.size();
	h.n_values = col->n_values;
	h.blocks_offset = (sizeof(h) + col->n_segs * sizeof(adacom_data_pointer) + 4095) & ~4095ull;
	h.total_bytes = h.blocks_offset + h.n_blocks * (uint64_t)kBlockAllocSize;
	h.meta_checksum = meta_checksum(h, plan.ptrs);
	return ADACOM_OK;
}

// pieces + piece_begin in one stream-ordered allocation
adacom_status upload_pieces(adacom_ctx *ctx, const std::vector<CkPiece> &pieces, const std::vector<uint32_t> &begin,
                            CkPiece **d_pieces, uint32_t **d_begin) {
	const size_t pb = (pieces.size() * sizeof(CkPiece) + 15) & ~(size_t)15;
	uint8_t *buf = nullptr;
	CK(ctx, cudaMallocAsync((void **)&buf, pb + begin.size() * sizeof(uint32_t) + 16, ctx->stream));
	if (!pieces.empty()) {
		CK(ctx, cudaMemcpyAsync(buf, pieces.data(), pieces.size() * sizeof(CkPiece), cudaMemcpyHostToDevice, ctx->stream));
	}
	CK(ctx, cudaMemcpyAsync(buf + pb, begin.data(), begin.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
	*d_pieces = (CkPiece *)buf;
	*d_begin = (uint32_t *)(buf + pb);
	return ADACOM_OK;
}

}  // namespace

extern "C" adacom_status adacom_column_checkpoint_size(adacom_column *col, uint64_t *bytes_out) {
	if (!col || !bytes_out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_column_checkpoint_size: NULL argument");
	}
	CtxLock lock(col->ctx);
	ImagePlan plan;
	adacom_status st = plan_image(col, plan);
	if (st != ADACOM_OK) {
		return st;
	}
	*bytes_out = plan.header.total_bytes;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_column_checkpoint(adacom_column *col, void *out, uint64_t capacity, int out_on_device,
                                                  uint64_t *bytes_out) {
	if (!col || !out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_column_checkpoint: NULL argument");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	CK(ctx, cudaSetDevice(ctx->device));
	ImagePlan plan;
	adacom_status st = plan_image(col, plan);
	if (st != ADACOM_OK) {
		return st;
	}
	const adacom_image_header &h = plan.header;
	if (capacity < h.total_bytes) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_column_checkpoint: capacity below adacom_column_checkpoint_size");
	}
	if (h.n_blocks >= (1ull << 32)) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_column_checkpoint: too many blocks");
	}
	// metadata area: header, pointer table, zero padding
	std::vector<uint8_t> meta(h.blocks_offset, 0);
	memcpy(meta.data(), &h, sizeof(h));
	if (!plan.ptrs.empty()) {
		memcpy(meta.data() + sizeof(h), plan.ptrs.data(), plan.ptrs.size() * sizeof(adacom_data_pointer));
	}
	CkPiece *d_pieces = nullptr;
	uint32_t *d_begin = nullptr;
	st = upload_pieces(ctx, plan.pieces, plan.piece_begin, &d_pieces, &d_begin);
	if (st != ADACOM_OK) {
		return st;
	}
	const uint64_t block_bytes = h.n_blocks * (uint64_t)kBlockAllocSize;
	uint8_t *d_blocks = nullptr;
	if (out_on_device) {
		if ((uintptr_t)out % 8 != 0) {
			return fail(ctx, ADACOM_ERR_INVALID, "adacom_column_checkpoint: device image must be 8-byte aligned");
		}
		d_blocks = (uint8_t *)out + h.blocks_offset;
		CK(ctx, cudaMemcpyAsync(out, meta.data(), meta.size(), cudaMemcpyHostToDevice, ctx->stream));
	} else {
		memcpy(out, meta.data(), meta.size());
		CK(ctx, cudaMallocAsync((void **)&d_blocks, block_bytes ? block_bytes : 8, ctx->stream));
	}
	LAUNCH(ctx, launch_checkpoint_blocks(d_pieces, d_begin, d_blocks, (uint32_t)h.n_blocks, ctx->sm_count, ctx->stream));
	if (!out_on_device) {
		if (block_bytes) {
			CK(ctx, cudaMemcpyAsync((uint8_t *)out + h.blocks_offset, d_blocks, block_bytes, cudaMemcpyDeviceToHost,
			                        ctx->stream));
		}
		CK(ctx, cudaFreeAsync(d_blocks, ctx->stream));
	}
	CK(ctx, cudaFreeAsync(d_pieces, ctx->stream));
	// the staging vectors above are pageable: the copies were staged before the calls returned; the image itself
	// must be complete when a host caller reads it
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	if (bytes_out) {
		*bytes_out = h.total_bytes;
	}
	return ADACOM_OK;
}

static adacom_status restore_locked(adacom_ctx *ctx, const void *image, uint64_t bytes, int image_on_device,
                                    adacom_column **col_io) {
	CK(ctx, cudaSetDevice(ctx->device));
	adacom_image_header h;
	if (bytes < sizeof(h)) {
		return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_column_restore: image shorter than its header");
	}
	if (image_on_device) {
		CK(ctx, cudaMemcpyAsync(&h, image, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
		CK(ctx, cudaStreamSynchronize(ctx->stream));
	} else {
		memcpy(&h, image, sizeof(h));
	}
	TypeInfo t;
	if (memcmp(h.magic, ADACOM_IMAGE_MAGIC, 8) != 0 || h.version != ADACOM_IMAGE_VERSION) {
		return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_column_restore: bad magic / version");
	}
	if (!type_info((adacom_ptype)h.ptype, t)) {
		return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_column_restore: unknown physical type");
	}
	if (h.n_segments >= (1ull << 32) || h.n_blocks >= (1ull << 32) ||
	    h.blocks_offset != ((sizeof(h) + h.n_segments * sizeof(adacom_data_pointer) + 4095) & ~4095ull) ||
	    h.total_bytes != h.blocks_offset + h.n_blocks * (uint64_t)kBlockAllocSize || h.total_bytes > bytes) {
		return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_column_restore: inconsistent header / truncated image");
	}
	std::vector<adacom_data_pointer> ptrs(h.n_segments);
	if (h.n_segments) {
		const uint8_t *src = (const uint8_t *)image + sizeof(h);
		if (image_on_device) {
			CK(ctx, cudaMemcpyAsync(ptrs.data(), src, ptrs.size() * sizeof(adacom_data_pointer), cudaMemcpyDeviceToHost,
			                        ctx->stream));
			CK(ctx, cudaStreamSynchronize(ctx->stream));
		} else {
			memcpy(ptrs.data(), src, ptrs.size() * sizeof(adacom_data_pointer));
		}
	}
	if (meta_checksum(h, ptrs) != h.meta_checksum) {
		return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_column_restore: metadata checksum mismatch");
	}
	// validate the pointers, lay out the packed extent
	const uint32_t S = (uint32_t)t.size;
	uint64_t row = 0, words = 0;
	bool any_raw = false;
	for (uint64_t g = 0; g < h.n_segments; g++) {
		const adacom_data_pointer &p = ptrs[g];
		const bool repacked = p.state & kSegRepacked;
		const uint64_t expect = repacked ? 8 * ((p.tuple_count * p.width + 63) >> 6) : p.tuple_count * S;
		bool ok = p.row_start == row && p.type_size == S && p.width >= 1 && p.width <= 64 &&
		          (uint64_t)p.capacity * S <= kBlockSize && p.tuple_count <= p.capacity && p.payload_bytes == expect &&
		          (repacked ? p.width < 8 * S : p.width == 8 * S);
		if (p.tuple_count == 0) {
			ok = ok && p.block_id == ADACOM_INVALID_BLOCK;
		} else {
			ok = ok && p.block_id >= 0 && (uint64_t)p.block_id < h.n_blocks && p.offset % 8 == 0 &&
			     (uint64_t)p.offset + p.payload_bytes <= kBlockSize;
		}
		if (!ok) {
			return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_column_restore: bad data pointer " + std::to_string(g));
		}
		row += p.tuple_count;
		if (repacked) {
			words += ((p.payload_bytes / 8 + kSegAlignWords - 1) / kSegAlignWords) * kSegAlignWords;
		} else if (p.tuple_count) {
			any_raw = true;
		}
	}
	if (row != h.n_values) {
		return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_column_restore: row count mismatch");
	}
	adacom_column *col = nullptr;
	adacom_status st = column_alloc(ctx, (adacom_ptype)h.ptype, h.n_segments, &col);
	if (st != ADACOM_OK) {
		return st;
	}
	*col_io = col;
	col->n_values = row;
	Extent ext;
	ext.words = words;
	ext.capacity = words;
	ext.seg_end = h.n_segments;
	CK(ctx, cudaMallocAsync((void **)&ext.base, words * 8 + kArenaPad, ctx->stream));
	col->extents.push_back(ext);
	col->latest = col->extents.size() - 1;
	CK(ctx, cudaMallocAsync((void **)&col->d_desc, std::max<uint64_t>(h.n_segments, 1) * sizeof(SegDesc), ctx->stream));
	if (any_raw) {
		st = ensure_raw(col);
		if (st != ADACOM_OK) {
			return st;
		}
	}
	std::vector<std::vector<CkPiece>> per_block(h.n_blocks);
	uint64_t woff = 0;
	for (uint64_t g = 0; g < h.n_segments; g++) {
		const adacom_data_pointer &p = ptrs[g];
		SegDesc &d = col->h_desc[g];
		memset(&d, 0, sizeof(d));
		const bool repacked = p.state & kSegRepacked;
		d.row_start = p.row_start;
		d.min_factor = p.min_factor;
		d.max_factor = p.max_factor;
		d.count = (uint32_t)p.tuple_count;
		d.capacity = p.capacity;
		d.width = p.width;
		d.state = p.state;
		CkPiece pc;
		pc.block_off = p.offset;
		pc.bytes = p.payload_bytes;
		if (repacked) {
			const uint64_t padded = ((p.payload_bytes / 8 + kSegAlignWords - 1) / kSegAlignWords) * kSegAlignWords;
			d.word_offset = woff;
			d.packed_ptr = (uint64_t)(ext.base + woff);
			pc.addr = d.packed_ptr;
			pc.pad_words = (uint32_t)padded;
			pc.elem_size = 8;
			woff += padded;
			col->extents[0].live++;
		} else {
			pc.addr = (uint64_t)(col->raw + p.row_start * S);
			pc.pad_words = 0;
			pc.elem_size = S;
		}
		if (p.tuple_count) {
			per_block[p.block_id].push_back(pc);
		}
	}
	std::vector<CkPiece> pieces;
	std::vector<uint32_t> begin(h.n_blocks + 1, 0);
	for (uint64_t b = 0; b < h.n_blocks; b++) {
		begin[b] = (uint32_t)pieces.size();
		std::sort(per_block[b].begin(), per_block[b].end(),
		          [](const CkPiece &a, const CkPiece &c) { return a.block_off < c.block_off; });
		uint32_t end = 0;
		for (const CkPiece &pc : per_block[b]) {
			if (pc.block_off < end) {
				return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_column_restore: overlapping payloads in block " + std::to_string(b));
			}
			end = pc.block_off + ((pc.bytes + 7u) & ~7u);
		}
		pieces.insert(pieces.end(), per_block[b].begin(), per_block[b].end());
	}
	begin[h.n_blocks] = (uint32_t)pieces.size();
	if (h.n_segments) {
		CK(ctx, cudaMemcpyAsync(col->d_desc, col->h_desc.data(), h.n_segments * sizeof(SegDesc), cudaMemcpyHostToDevice,
		                        ctx->stream));
	}
	st = build_tiles(col);
	if (st != ADACOM_OK) {
		return st;
	}
	CkPiece *d_pieces = nullptr;
	uint32_t *d_begin = nullptr;
	st = upload_pieces(ctx, pieces, begin, &d_pieces, &d_begin);
	if (st != ADACOM_OK) {
		return st;
	}
	const uint64_t block_bytes = h.n_blocks * (uint64_t)kBlockAllocSize;
	const uint8_t *d_blocks = (const uint8_t *)image + h.blocks_offset;
	uint8_t *temp = nullptr;
	if (!image_on_device) {
		CK(ctx, cudaMallocAsync((void **)&temp, block_bytes ? block_bytes : 8, ctx->stream));
		if (block_bytes) {
			CK(ctx, cudaMemcpyAsync(temp, d_blocks, block_bytes, cudaMemcpyHostToDevice, ctx->stream));
		}
		d_blocks = temp;
	} else if ((uintptr_t)image % 8 != 0) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_column_restore: device image must be 8-byte aligned");
	}
	// verdict word: host-mapped totals slot 2 would race with an encode's ticket spin on other columns; use the
	// context's aggregate scratch instead (device) and copy it back
	unsigned long long *d_bad = (unsigned long long *)ctx->d_agg;
	CK(ctx, cudaMemsetAsync(d_bad, 0xff, sizeof(unsigned long long), ctx->stream));
	LAUNCH(ctx, launch_restore_blocks(d_pieces, d_begin, d_blocks, (uint32_t)h.n_blocks, d_bad, ctx->sm_count, ctx->stream));
	unsigned long long bad = 0;
	CK(ctx, cudaMemcpyAsync(&bad, d_bad, sizeof(bad), cudaMemcpyDeviceToHost, ctx->stream));
	if (temp) {
		CK(ctx, cudaFreeAsync(temp, ctx->stream));
	}
	CK(ctx, cudaFreeAsync(d_pieces, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	if (bad != ~0ull) {
		return fail(ctx, ADACOM_ERR_CORRUPT,
		            "adacom_column_restore: computed checksum does not match stored checksum in block " +
		                std::to_string(bad));
	}
	return ADACOM_OK;
}

extern "C" adacom_status adacom_column_restore(adacom_ctx *ctx, const void *image, uint64_t bytes, int image_on_device,
                                               adacom_column **out) {
	if (!ctx || !image || !out) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_column_restore: NULL argument");
	}
	*out = nullptr;
	adacom_column *col = nullptr;
	adacom_status st;
	std::string err;
	{
		CtxLock lock(ctx);
		st = restore_locked(ctx, image, bytes, image_on_device, &col);
		err = ctx->err;
	}
	if (st != ADACOM_OK) {
		if (col) {
			adacom_column_destroy(col);
			ctx->err = err;
		}
		return st;
	}
	*out = col;
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------
// neighbour codecs: DuckDB bitpacking / RLE segment images (kernels_codecs.cu)
// ---------------------------------------------------------------------------------------------------

struct adacom_codec_column {
	adacom_ctx *ctx = nullptr;
	int codec = 0;
	TypeInfo t {4, false};
	uint64_t n_segs = 0, n_values = 0, image_bytes = 0;
	uint8_t *d_images = nullptr;
	CodecSeg *d_segs = nullptr;
	std::vector<CodecSeg> h_segs;
	Job *d_jobs = nullptr;        // (segment, 2048-value group) in segment order
	std::vector<uint32_t> job_first;
	CodecGroup *d_groups = nullptr;  // one record per entry of d_jobs, resolved at creation (codec_groups_kernel)
	Job *d_sum_jobs = nullptr;    // RLE SUM: (segment, tile of 2048 runs)
	std::vector<uint32_t> sum_job_first;
	uint32_t *d_run_start = nullptr;
};

extern "C" adacom_status adacom_codec_column_create(adacom_ctx *ctx, adacom_codec codec, adacom_ptype type,
                                                    uint64_t n_segments, const uint8_t *const *segment_images,
                                                    const uint64_t *segment_bytes, const uint64_t *tuple_counts,
                                                    adacom_codec_column **out) {
	if (!ctx || !out || (n_segments && (!segment_images || !segment_bytes || !tuple_counts)) ||
	    ((int)codec != ADACOM_CODEC_BITPACKING && (int)codec != ADACOM_CODEC_RLE) || n_segments >= (1ull << 31)) {
		return fail(ctx, ADACOM_ERR_INVALID, "adacom_codec_column_create: bad arguments");
	}
	TypeInfo t;
	if (!type_info(type, t)) {
		return fail(ctx, ADACOM_ERR_UNSUPPORTED, "adacom_codec_column_create: unsupported physical type");
	}
	CtxLock lock(ctx);
	CK(ctx, cudaSetDevice(ctx->device));
	std::unique_ptr<adacom_codec_column> col(new adacom_codec_column());
	col->ctx = ctx;
	col->codec = (int)codec;
	col->t = t;
	col->n_segs = n_segments;
	col->h_segs.resize(n_segments);
	std::vector<Job> jobs, sum_jobs;
	col->job_first.assign(n_segments + 1, 0);
	col->sum_job_first.assign(n_segments + 1, 0);
	uint64_t offset = 0, row = 0, runs = 0;
	for (uint64_t g = 0; g < n_segments; g++) {
		const uint64_t bytes = segment_bytes[g], count = tuple_counts[g];
		if (bytes < 8 || bytes > kBlockSize || count >= (1ull << 32) || !segment_images[g]) {
			return fail(ctx, ADACOM_ERR_INVALID, "adacom_codec_column_create: malformed segment image");
		}
		uint64_t head;
		memcpy(&head, segment_images[g], 8);
		CodecSeg &sd = col->h_segs[g];
		sd.image_offset = offset;
		sd.row_start = row;
		sd.bytes = (uint32_t)bytes;
		sd.count = (uint32_t)count;
		sd.first_run = (uint32_t)runs;
		sd.reserved = 0;
		const uint64_t n_groups = (count + kVectorSize - 1) / kVectorSize;
		if (codec == ADACOM_CODEC_BITPACKING) {
			// the metadata entries sit right below `head`: one per group, all inside the image
			if (head > bytes || head < 8 + 4 * n_groups) {
				return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_codec_column_create: bitpacking metadata offset out of range");
			}
			for (uint64_t k = 0; k < n_groups; k++) {
				uint32_t entry;
				memcpy(&entry, segment_images[g] + head - 4 * (k + 1), 4);
				const uint32_t mode = entry >> 24, off = entry & 0x00FFFFFFu;
				if (mode < 1 || mode > 4 || off < 8 || off + (uint64_t)t.size > head) {
					return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_codec_column_create: bitpacking group metadata malformed");
				}
			}
		} else {
			if (head > bytes || head < 8 || ((bytes - head) & 1)) {
				return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_codec_column_create: RLE count offset out of range");
			}
			const uint64_t n_runs = (bytes - head) / 2;
			if (8 + n_runs * (uint64_t)t.size > head) {
				return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_codec_column_create: RLE values overlap the run lengths");
			}
			uint64_t total = 0;
			for (uint64_t k = 0; k < n_runs; k++) {
				uint16_t c;
				memcpy(&c, segment_images[g] + head + 2 * k, 2);
				total += c;
			}
			if (total < count) {
				return fail(ctx, ADACOM_ERR_CORRUPT, "adacom_codec_column_create: RLE runs shorter than the tuple count");
			}
			for (uint64_t k = 0; k < (n_runs + kVectorSize - 1) / kVectorSize; k++) {
				sum_jobs.push_back(Job {(uint32_t)g, (uint32_t)k});
			}
			runs += n_runs;
		}
		for (uint64_t k = 0; k < n_groups; k++) {
			jobs.push_back(Job {(uint32_t)g, (uint32_t)k});
		}
		col->job_first[g + 1] = (uint32_t)jobs.size();
		col->sum_job_first[g + 1] = (uint32_t)sum_jobs.size();
		offset += (bytes + 15) & ~15ull;
		row += count;
	}
	col->n_values = row;
	col->image_bytes = offset;
	CK(ctx, cudaMallocAsync((void **)&col->d_images, offset + kArenaPad, ctx->stream));
	CK(ctx, cudaMemsetAsync(col->d_images + offset, 0, kArenaPad, ctx->stream));
	for (uint64_t g = 0; g < n_segments; g++) {
		CK(ctx, cudaMemcpyAsync(col->d_images + col->h_segs[g].image_offset, segment_images[g], segment_bytes[g],
		                        cudaMemcpyHostToDevice, ctx->stream));
	}
	CK(ctx, cudaMallocAsync((void **)&col->d_segs, std::max<size_t>(n_segments, 1) * sizeof(CodecSeg), ctx->stream));
	CK(ctx, cudaMemcpyAsync(col->d_segs, col->h_segs.data(), n_segments * sizeof(CodecSeg), cudaMemcpyHostToDevice, ctx->stream));
	CK(ctx, cudaMallocAsync((void **)&col->d_jobs, std::max<size_t>(jobs.size(), 1) * sizeof(Job), ctx->stream));
	CK(ctx, cudaMemcpyAsync(col->d_jobs, jobs.data(), jobs.size() * sizeof(Job), cudaMemcpyHostToDevice, ctx->stream));
	if (codec == ADACOM_CODEC_RLE) {
		CK(ctx, cudaMallocAsync((void **)&col->d_sum_jobs, std::max<size_t>(sum_jobs.size(), 1) * sizeof(Job), ctx->stream));
		CK(ctx, cudaMemcpyAsync(col->d_sum_jobs, sum_jobs.data(), sum_jobs.size() * sizeof(Job), cudaMemcpyHostToDevice, ctx->stream));
		CK(ctx, cudaMallocAsync((void **)&col->d_run_start, std::max<uint64_t>(runs, 1) * sizeof(uint32_t), ctx->stream));
		CodecArgs a;
		a.images = col->d_images;
		a.segs = col->d_segs;
		a.run_start = col->d_run_start;
		LAUNCH(ctx, launch_rle_run_starts(a, 0, (uint32_t)n_segments, ctx->sm_count, ctx->stream));
	}
	{
		CK(ctx, cudaMallocAsync((void **)&col->d_groups, std::max<size_t>(jobs.size(), 1) * sizeof(CodecGroup), ctx->stream));
		CodecArgs a;
		a.images = col->d_images;
		a.segs = col->d_segs;
		a.run_start = col->d_run_start;
		a.jobs = col->d_jobs;
		a.n_jobs = (uint32_t)jobs.size();
		LAUNCH(ctx, launch_codec_groups(col->t, col->codec, a, col->d_groups, ctx->stream));
	}
	CK(ctx, cudaStreamSynchronize(ctx->stream));  // the host vectors and the caller's images may go away now
	*out = col.release();
	return ADACOM_OK;
}

extern "C" void adacom_codec_column_destroy(adacom_codec_column *col) {
	if (!col) {
		return;
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	cudaSetDevice(ctx->device);
	for (void *p : {(void *)col->d_images, (void *)col->d_segs, (void *)col->d_jobs, (void *)col->d_sum_jobs,
	                (void *)col->d_run_start, (void *)col->d_groups}) {
		if (p) {
			cudaFreeAsync(p, ctx->stream);
		}
	}
	cudaStreamSynchronize(ctx->stream);
	delete col;
}

extern "C" uint64_t adacom_codec_column_num_values(const adacom_codec_column *col) {
	return col ? col->n_values : 0;
}

extern "C" uint64_t adacom_codec_column_image_bytes(const adacom_codec_column *col) {
	return col ? col->image_bytes : 0;
}

static bool codec_range_ok(const adacom_codec_column *col, uint64_t b, uint64_t e) {
	return col && b <= e && e <= col->n_segs;
}

static CodecArgs codec_args(adacom_codec_column *col, uint64_t b, uint64_t e, bool sum_jobs) {
	CodecArgs a;
	a.images = col->d_images;
	a.segs = col->d_segs;
	a.run_start = col->d_run_start;
	if (sum_jobs) {
		a.jobs = col->d_sum_jobs + col->sum_job_first[b];
		a.n_jobs = col->sum_job_first[e] - col->sum_job_first[b];
	} else {
		a.jobs = col->d_jobs + col->job_first[b];
		a.groups = col->d_groups + col->job_first[b];
		a.n_jobs = col->job_first[e] - col->job_first[b];
	}
	a.out_row0 = b < col->n_segs ? col->h_segs[b].row_start : 0;
	return a;
}

extern "C" adacom_status adacom_codec_decode(adacom_codec_column *col, uint64_t seg_begin, uint64_t seg_end, void *out,
                                             int out_on_device) {
	if (!codec_range_ok(col, seg_begin, seg_end) || !out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_codec_decode: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (seg_begin == seg_end) {
		return ADACOM_OK;
	}
	const CodecSeg &last = col->h_segs[seg_end - 1];
	const uint64_t rows = last.row_start + last.count - col->h_segs[seg_begin].row_start;
	if (rows == 0) {
		return ADACOM_OK;
	}
	void *d_out = out;
	if (!out_on_device) {
		CK(ctx, cudaMallocAsync(&d_out, rows * col->t.size, ctx->stream));
	}
	CodecArgs a = codec_args(col, seg_begin, seg_end, false);
	a.out = d_out;
	LAUNCH(ctx, launch_codec_scan(col->t, col->codec, 0, a, ctx->sm_count, ctx->stream));
	if (!out_on_device) {
		CK(ctx, cudaMemcpyAsync(out, d_out, rows * col->t.size, cudaMemcpyDeviceToHost, ctx->stream));
		CK(ctx, cudaFreeAsync(d_out, ctx->stream));
		CK(ctx, cudaStreamSynchronize(ctx->stream));
	}
	return ADACOM_OK;
}

extern "C" adacom_status adacom_codec_scan_sum(adacom_codec_column *col, uint64_t seg_begin, uint64_t seg_end,
                                               adacom_agg *out) {
	if (!codec_range_ok(col, seg_begin, seg_end) || !out) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_codec_scan_sum: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	CK(ctx, cudaMemsetAsync(ctx->d_agg, 0, sizeof(Agg), ctx->stream));
	if (seg_begin != seg_end) {
		CodecArgs a = codec_args(col, seg_begin, seg_end, col->codec == ADACOM_CODEC_RLE);
		a.agg = ctx->d_agg;
		LAUNCH(ctx, launch_codec_scan(col->t, col->codec, 1, a, ctx->sm_count, ctx->stream));
	}
	CK(ctx, cudaMemcpyAsync(ctx->h_agg, ctx->d_agg, sizeof(Agg), cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	memcpy(out, ctx->h_agg, sizeof(Agg));
	if (col->codec == ADACOM_CODEC_RLE && seg_begin != seg_end) {
		// the last run of a segment may be longer than the tuples it still covers; COUNT is the tuple count
		const CodecSeg &last = col->h_segs[seg_end - 1];
		out->count = last.row_start + last.count - col->h_segs[seg_begin].row_start;
	}
	return ADACOM_OK;
}

extern "C" adacom_status adacom_codec_fetch_rows(adacom_codec_column *col, uint64_t n, const uint32_t *seg_idx,
                                                 const uint32_t *row_in_seg, void *out, int on_device) {
	if (!col || (n && (!seg_idx || !row_in_seg || !out)) || n >= (1ull << 26)) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_codec_fetch_rows: bad arguments");
	}
	adacom_ctx *ctx = col->ctx;
	CtxLock lock(ctx);
	if (n == 0) {
		return ADACOM_OK;
	}
	CodecArgs a = codec_args(col, 0, 0, false);
	if (on_device) {
		a.out = out;
		LAUNCH(ctx, launch_codec_fetch(col->t, col->codec, a, seg_idx, row_in_seg, n, ctx->stream));
		return ADACOM_OK;
	}
	for (uint64_t q = 0; q < n; q++) {
		if (seg_idx[q] >= col->n_segs || row_in_seg[q] >= col->h_segs[seg_idx[q]].count) {
			return fail(ctx, ADACOM_ERR_INVALID, "adacom_codec_fetch_rows: (segment, row) out of range");
		}
	}
	uint32_t *d_ids = nullptr;
	void *d_out = nullptr;
	CK(ctx, cudaMallocAsync((void **)&d_ids, 2 * n * sizeof(uint32_t), ctx->stream));
	CK(ctx, cudaMallocAsync(&d_out, n * col->t.size, ctx->stream));
	CK(ctx, cudaMemcpyAsync(d_ids, seg_idx, n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
	CK(ctx, cudaMemcpyAsync(d_ids + n, row_in_seg, n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
	a.out = d_out;
	LAUNCH(ctx, launch_codec_fetch(col->t, col->codec, a, d_ids, d_ids + n, n, ctx->stream));
	CK(ctx, cudaMemcpyAsync(out, d_out, n * col->t.size, cudaMemcpyDeviceToHost, ctx->stream));
	CK(ctx, cudaFreeAsync(d_ids, ctx->stream));
	CK(ctx, cudaFreeAsync(d_out, ctx->stream));
	CK(ctx, cudaStreamSynchronize(ctx->stream));
	return ADACOM_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Pull-model scan stream (include/adacom_gpu.h: adacom_scan_stream_*).  Decode kernels go onto the context's stream
// (ordered with the encodes / uncompacts of the column, so a window never reads a descriptor in mid-update); the
// device->host copies of the windows run on the scan's own stream behind an event, and the consumer waits on events
// only -- never on the context mutex, never on a stream synchronisation.
// ---------------------------------------------------------------------------------------------------------------
struct adacom_scan_stream {
	adacom_column *col = nullptr;
	uint64_t seg_begin = 0, seg_end = 0;
	uint32_t win_segs = 32, depth = 3;
	int flags = 0;
	cudaStream_t copy_stream = nullptr;
	uint8_t *d_ring = nullptr, *h_ring = nullptr;
	size_t win_bytes = 0, ring_bytes = 0;
	struct Win {
		uint64_t index = ~0ull;  // window number held (or in flight) in this slot
		cudaEvent_t decoded = nullptr, arrived = nullptr;
		bool waited = false;     // the consumer has already synchronised on `arrived`
		bool used = false;       // `arrived` has been recorded at least once
	};
	std::vector<Win> wins;
	uint64_t n_windows = 0;
	// first segment of every window (+ seg_end).  Default: a read-ahead ramp of 1, 2, 4, ... win_segs segments, so the
	// first vector of a short scan (a point filter that touches one segment) waits for 256 KiB, not for 8 MiB
	std::vector<uint64_t> win_start;
};

namespace {

uint64_t window_first(const adacom_scan_stream *s, uint64_t w) {
	return s->win_start[w];
}
uint64_t window_end(const adacom_scan_stream *s, uint64_t w) {
	return s->win_start[w + 1];
}

// enqueue window w into its slot: decode on the context stream, D2H on the copy stream
adacom_status issue_window(adacom_scan_stream *s, uint64_t w) {
	adacom_column *col = s->col;
	adacom_ctx *ctx = col->ctx;
	adacom_scan_stream::Win &win = s->wins[w % s->depth];
	const uint64_t b = window_first(s, w), e = window_end(s, w);
	const uint64_t rows = range_rows(col, b, e);
	uint8_t *d_buf = s->d_ring + (w % s->depth) * s->win_bytes;
	uint8_t *h_buf = s->h_ring + (w % s->depth) * s->win_bytes;
	CtxLock lock(ctx);
	CK(ctx, cudaSetDevice(ctx->device));
	if (win.used) {
		// the slot's previous copy must have left the device buffer before the kernel overwrites it
		CK(ctx, cudaStreamWaitEvent(ctx->stream, win.arrived, 0));
	}
	if (rows) {
		DecodeArgs a;
		a.desc = col->d_desc;
		a.jobs = col->d_tiles + col->tile_first[b];
		a.n_jobs = col->tile_first[e] - col->tile_first[b];
		a.raw = col->raw;
		a.literal = (s->flags & ADACOM_LITERAL) ? 1 : 0;
		a.out = d_buf;
		a.out_row0 = col->h_desc[b].row_start;
		LAUNCH(ctx, launch_decode(col->t, kModeFlat, a, ctx->sm_count, ctx->stream));
	}
	CK(ctx, cudaEventRecord(win.decoded, ctx->stream));
	CK(ctx, cudaStreamWaitEvent(s->copy_stream, win.decoded, 0));
	if (rows) {
		CK(ctx, cudaMemcpyAsync(h_buf, d_buf, rows * col->t.size, cudaMemcpyDeviceToHost, s->copy_stream));
	}
	CK(ctx, cudaEventRecord(win.arrived, s->copy_stream));
	win.index = w;
	win.waited = false;
	win.used = true;
	return ADACOM_OK;
}

}  // namespace

extern "C" adacom_status adacom_scan_stream_open(adacom_column *col, uint64_t seg_begin, uint64_t seg_end,
                                                 uint32_t window_segments, uint32_t depth, int flags,
                                                 adacom_scan_stream **out) {
	if (!range_ok(col, seg_begin, seg_end) || !out || depth == 1) {
		return fail(col ? col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_scan_stream_open: bad arguments");
	}
	*out = nullptr;
	adacom_ctx *ctx = col->ctx;
	adacom_scan_stream *s = new (std::nothrow) adacom_scan_stream();
	if (!s) {
		return fail(ctx, ADACOM_ERR_NOMEM, "out of host memory");
	}
	s->col = col;
	s->seg_begin = seg_begin;
	s->seg_end = seg_end;
	s->win_segs = window_segments ? window_segments : 32;
	s->depth = depth ? depth : 3;
	s->flags = flags;
	{
		uint64_t at = seg_begin, step = window_segments ? s->win_segs : 1;
		while (at < seg_end) {
			s->win_start.push_back(at);
			at += step;
			step = step * 2 < s->win_segs ? step * 2 : s->win_segs;
		}
		if (s->win_start.empty()) {
			s->win_start.push_back(seg_begin);
		}
		s->n_windows = seg_end > seg_begin ? s->win_start.size() : 0;
		s->win_start.push_back(seg_end);
	}
	if (s->n_windows < s->depth) {
		s->depth = s->n_windows < 2 ? 2 : (uint32_t)s->n_windows;
	}
	uint64_t max_rows = 0;
	for (uint64_t w = 0; w < s->n_windows; w++) {
		const uint64_t r = range_rows(col, window_first(s, w), window_end(s, w));
		max_rows = r > max_rows ? r : max_rows;
	}
	s->win_bytes = ((size_t)max_rows * col->t.size + kArenaPad + 255) & ~(size_t)255;
	s->wins.resize(s->depth);
	auto bail = [&](cudaError_t e, const char *what) {
		std::string msg = std::string(what) + ": " + cudaGetErrorString(e);
		adacom_scan_stream_close(s);
		return fail(ctx, e == cudaErrorMemoryAllocation ? ADACOM_ERR_NOMEM : ADACOM_ERR_CUDA, msg);
	};
	cudaError_t e;
	const size_t need = s->win_bytes * s->depth;
	std::vector<cudaEvent_t> events;
	{
		CtxLock lock(ctx);
		if ((e = cudaSetDevice(ctx->device)) != cudaSuccess) {
			return bail(e, "cudaSetDevice");
		}
		// a ring of a closed stream that is large enough (smallest such)
		int best = -1;
		for (size_t i = 0; i < ctx->ring_cache.size(); i++) {
			if (ctx->ring_cache[i].bytes >= need && (best < 0 || ctx->ring_cache[i].bytes < ctx->ring_cache[best].bytes)) {
				best = (int)i;
			}
		}
		if (best >= 0) {
			adacom_ctx::ScanRing r = std::move(ctx->ring_cache[best]);
			ctx->ring_cache.erase(ctx->ring_cache.begin() + best);
			s->d_ring = r.d_ring;
			s->h_ring = r.h_ring;
			s->ring_bytes = r.bytes;
			s->copy_stream = r.copy_stream;
			events = std::move(r.events);
		}
	}
	if (!s->copy_stream) {
		if ((e = cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) {
			return bail(e, "cudaStreamCreateWithFlags");
		}
		if ((e = cudaMalloc((void **)&s->d_ring, need)) != cudaSuccess) {
			return bail(e, "cudaMalloc(scan ring)");
		}
		if ((e = cudaHostAlloc((void **)&s->h_ring, need, cudaHostAllocDefault)) != cudaSuccess) {
			return bail(e, "cudaHostAlloc(scan ring)");
		}
		s->ring_bytes = need;
	}
	for (auto &w : s->wins) {
		for (cudaEvent_t *ev : {&w.decoded, &w.arrived}) {
			if (!events.empty()) {
				*ev = events.back();
				events.pop_back();
			} else if ((e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming)) != cudaSuccess) {
				return bail(e, "cudaEventCreateWithFlags");
			}
		}
	}
	for (auto ev : events) {
		cudaEventDestroy(ev);
	}
	// run ahead right away: the first depth-1 windows are in flight before the first pull
	for (uint64_t w = 0; w + 1 < s->depth && w < s->n_windows; w++) {
		adacom_status st = issue_window(s, w);
		if (st != ADACOM_OK) {
			adacom_scan_stream_close(s);
			return st;
		}
	}
	*out = s;
	return ADACOM_OK;
}

extern "C" adacom_status adacom_scan_stream_segment(adacom_scan_stream *s, uint64_t seg, const void **values,
                                                    uint64_t *count) {
	if (!s || !values || seg < s->seg_begin || seg >= s->seg_end) {
		return fail(s ? s->col->ctx : nullptr, ADACOM_ERR_INVALID, "adacom_scan_stream_segment: segment outside the stream");
	}
	adacom_column *col = s->col;
	adacom_ctx *ctx = col->ctx;
	const uint64_t w = (uint64_t)(std::upper_bound(s->win_start.begin(), s->win_start.end(), seg) - s->win_start.begin()) - 1;
	adacom_scan_stream::Win *win = &s->wins[w % s->depth];
	if (win->index != w) {
		// not resident and not in flight: (re)start the pipeline here
		adacom_status st = issue_window(s, w);
		if (st != ADACOM_OK) {
			return st;
		}
	}
	if (!win->waited) {
		// first touch of this window: keep the ring full -- windows w+1 .. w+depth-2 (the slot of w-1 may still be
		// referenced by the consumer's last vector and is left alone)
		for (uint64_t k = w + 1; k + 1 < w + s->depth && k < s->n_windows; k++) {
			if (s->wins[k % s->depth].index != k) {
				adacom_status st = issue_window(s, k);
				if (st != ADACOM_OK) {
					return st;
				}
			}
		}
		CK(ctx, cudaEventSynchronize(win->arrived));
		win->waited = true;
	}
	const uint64_t row0 = col->h_desc[window_first(s, w)].row_start;
	*values = s->h_ring + (w % s->depth) * s->win_bytes + (col->h_desc[seg].row_start - row0) * col->t.size;
	if (count) {
		*count = col->h_desc[seg].count;
	}
	return ADACOM_OK;
}

extern "C" void adacom_scan_stream_close(adacom_scan_stream *s) {
	if (!s) {
		return;
	}
	adacom_ctx *ctx = s->col->ctx;
	{
		CtxLock lock(ctx);
		cudaSetDevice(ctx->device);
	}
	if (s->copy_stream) {
		cudaStreamSynchronize(s->copy_stream);
	}
	if (s->copy_stream && s->d_ring && s->h_ring) {
		// the decode kernels of this stream's windows sit on the context stream: by the time a later stream reuses
		// the ring its first kernel is ordered behind them there, and the copies have drained above
		adacom_ctx::ScanRing r;
		r.d_ring = s->d_ring;
		r.h_ring = s->h_ring;
		r.bytes = s->ring_bytes;
		r.copy_stream = s->copy_stream;
		for (auto &w : s->wins) {
			for (cudaEvent_t ev : {w.decoded, w.arrived}) {
				if (ev) {
					r.events.push_back(ev);
				}
			}
		}
		CtxLock lock(ctx);
		if (ctx->ring_cache.size() < 16) {
			ctx->ring_cache.push_back(std::move(r));
			delete s;
			return;
		}
	}
	for (auto &w : s->wins) {
		if (w.arrived) {
			cudaEventDestroy(w.arrived);
		}
		if (w.decoded) {
			cudaEventDestroy(w.decoded);
		}
	}
	if (s->h_ring) {
		cudaFreeHost(s->h_ring);
	}
	if (s->d_ring) {
		cudaFree(s->d_ring);
	}
	if (s->copy_stream) {
		cudaStreamDestroy(s->copy_stream);
	}
	delete s;
}
