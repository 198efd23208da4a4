// succinct_gpu.cpp -- see succinct_gpu.hpp.  Compiled against the reference's headers; forwards every
// CompressionFunction slot of the succinct codec to the C ABI in include/adacom_gpu.h.
#include "succinct_gpu.hpp"

#include "adacom_gpu.h"
#include "duckdb/common/types/null_value.hpp"
#include "duckdb/common/types/vector.hpp"
#include "duckdb/main/config.hpp"
#include "duckdb/storage/buffer_manager.hpp"
#include "duckdb/storage/segment/uncompressed.hpp"
#include "duckdb/storage/statistics/numeric_statistics.hpp"
#include "duckdb/storage/table/append_state.hpp"
#include "duckdb/storage/table/column_data_checkpointer.hpp"
#include "duckdb/storage/table/column_segment.hpp"
#include "duckdb/storage/table/scan_state.hpp"
#include "duckdb/storage/table/row_group.hpp"

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

namespace duckdb {

namespace {

//===--------------------------------------------------------------------===//
// Runtime: one codec context per process (one GPU), errors -> the reference's two exception types
//===--------------------------------------------------------------------===//
//! A pool of codec contexts: ADACOM_DEVICES GPUs (default 1) x ADACOM_STREAMS contexts per GPU (default 4).  Every
//! context owns its stream and serialises only its own calls, so DuckDB's scan workers (one row group per task,
//! row_group_collection.cpp:119-153) run concurrently.  Row group g lives on GPU g mod G: all segments of a
//! row group are on the same device, nothing moves between GPUs.
struct GpuRuntime {
	std::vector<adacom_ctx *> contexts; // [device * streams + k]
	idx_t devices = 1, streams = 4;
	adacom_status status = ADACOM_OK;
	std::string error;

	GpuRuntime() {
		if (const char *env = getenv("ADACOM_DEVICES")) {
			devices = MaxValue<idx_t>(1, (idx_t)atoi(env));
		}
		if (const char *env = getenv("ADACOM_STREAMS")) {
			streams = MaxValue<idx_t>(1, (idx_t)atoi(env));
		}
		idx_t first = 0;
		if (const char *env = getenv("ADACOM_FIRST_DEVICE")) {
			first = (idx_t)atoi(env);
		}
		for (idx_t d = 0; d < devices && status == ADACOM_OK; d++) {
			for (idx_t k = 0; k < streams && status == ADACOM_OK; k++) {
				adacom_ctx *ctx = nullptr;
				status = adacom_ctx_create((int)(first + d), &ctx);
				if (status == ADACOM_OK) {
					contexts.push_back(ctx);
				} else {
					error = adacom_last_error(nullptr);
				}
			}
		}
	}
};

//! ADACOM_BINDING_STATS=1: wall time spent inside the C-ABI calls, printed when the process exits
struct BindingStats {
	std::atomic<uint64_t> decode_ns {0}, decode_calls {0}, compact_ns {0}, compact_calls {0}, upload_ns {0},
	    upload_calls {0}, slice_ns {0}, slice_calls {0};
	bool enabled = getenv("ADACOM_BINDING_STATS") != nullptr;
	~BindingStats() {
		if (enabled) {
			fprintf(stderr,
			        "[adacom binding] decode %llu calls %.3f ms | compact %llu calls %.3f ms | upload %llu calls %.3f ms | "
			        "slices %llu calls %.3f ms\n",
			        (unsigned long long)decode_calls, decode_ns / 1e6, (unsigned long long)compact_calls, compact_ns / 1e6,
			        (unsigned long long)upload_calls, upload_ns / 1e6, (unsigned long long)slice_calls, slice_ns / 1e6);
		}
	}
};
BindingStats &Stats() {
	static BindingStats stats;
	return stats;
}
struct StatTimer {
	std::atomic<uint64_t> &ns, &calls;
	std::chrono::steady_clock::time_point t0;
	bool on;
	StatTimer(std::atomic<uint64_t> &ns_p, std::atomic<uint64_t> &calls_p) : ns(ns_p), calls(calls_p), on(Stats().enabled) {
		if (on) {
			t0 = std::chrono::steady_clock::now();
		}
	}
	~StatTimer() {
		if (on) {
			ns += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
			calls++;
		}
	}
};

GpuRuntime &Runtime() {
	static GpuRuntime runtime;
	if (runtime.status != ADACOM_OK) {
		// no CPU fallback: without the GPU library / device the codec is unavailable, loudly
		throw InternalException("succinct GPU codec unavailable: %s", runtime.error);
	}
	return runtime;
}

//! context of a segment: device = row group mod G, stream slot spread over the segments of the device
adacom_ctx *GpuContextFor(idx_t row_start) {
	auto &rt = Runtime();
	idx_t row_group = row_start / RowGroup::ROW_GROUP_SIZE;
	idx_t device = row_group % rt.devices;
	idx_t slot = (row_start / 32768) % rt.streams;
	return rt.contexts[device * rt.streams + slot];
}

void Check(adacom_status st, const char *what) {
	if (st == ADACOM_OK) {
		return;
	}
	if (st == ADACOM_ERR_NOMEM) {
		throw std::bad_alloc(); // sdsl memory_management.hpp:326-329
	}
	throw InternalException("succinct GPU codec, %s failed with status %d", what, (int)st);
}

adacom_ptype GpuTypeOf(PhysicalType type) {
	// the enumerators of adacom_ptype ARE duckdb::PhysicalType values (types.hpp), checked here
	static_assert((int)PhysicalType::UINT8 == ADACOM_UINT8 && (int)PhysicalType::INT8 == ADACOM_INT8 &&
	                  (int)PhysicalType::UINT16 == ADACOM_UINT16 && (int)PhysicalType::INT16 == ADACOM_INT16 &&
	                  (int)PhysicalType::UINT32 == ADACOM_UINT32 && (int)PhysicalType::INT32 == ADACOM_INT32 &&
	                  (int)PhysicalType::UINT64 == ADACOM_UINT64 && (int)PhysicalType::INT64 == ADACOM_INT64,
	              "adacom_ptype mirrors duckdb::PhysicalType");
	return (adacom_ptype)(int)type;
}

//===--------------------------------------------------------------------===//
// Per-segment state (ColumnSegment::segment_state): host slots while appending, device column afterwards
//===--------------------------------------------------------------------===//
struct SuccinctGpuSegmentState : public CompressedSegmentState {
	~SuccinctGpuSegmentState() override {
		if (column) {
			adacom_column_destroy(column);
		}
	}

	//! the context (GPU + stream) this segment lives on
	adacom_ctx *ctx = nullptr;
	//! one-segment device column (raw slots and, once compacted, the packed payload)
	adacom_column *column = nullptr;
	//! the slot-width array of an open segment (what succinct_vec holds in the reference, column_segment.cpp:101-105);
	//! released once the segment is compacted on the device
	unique_ptr<data_t[]> slots;
	//! validity words of the slots (bit i = row i valid), allocated when the first NULL is appended: the reference
	//! keeps NULL rows out of min/max (succinct.cpp:284-292), so adacom_analyze needs the mask
	std::vector<uint64_t> validity;
	//! values of `slots` that are already on the device
	idx_t resident = 0;
	//! ColumnSegment::GetDataSize() of the device state (0 = not known yet)
	idx_t data_size = 0;
	//! segment.count at which the device state was last compacted (Compact() is idempotent: the reference's lazy
	//! `if (!compacted) Compact()` on the scan path is unlocked, column_segment.cpp:157-160, so two scan threads can
	//! both decide to compact the same segment -- the second call must not find the raw slots gone)
	idx_t compacted_for = (idx_t)-1;
	//! serialises upload / re-encode against concurrent scans (the reference's bit_compression_lock)
	std::mutex lock;
};

SuccinctGpuSegmentState &StateOf(ColumnSegment &segment) {
	auto state = segment.GetSegmentState();
	if (!state) {
		throw InternalException("succinct GPU codec: segment without device state");
	}
	return (SuccinctGpuSegmentState &)*state;
}

//! make the first segment.count slots resident on the device (no-op once uploaded)
void EnsureResident(ColumnSegment &segment, SuccinctGpuSegmentState &state) {
	if (state.column && state.resident == segment.count) {
		return;
	}
	StatTimer timer(Stats().upload_ns, Stats().upload_calls);
	if (state.column) { // the segment grew since the last upload: rebuild (appends and scans rarely interleave)
		adacom_column_destroy(state.column);
		state.column = nullptr;
	}
	uint64_t count = segment.count;
	auto capacity = (uint32_t)(segment.SegmentSize() / segment.type_size);
	if (!state.ctx) {
		state.ctx = GpuContextFor(segment.start);
	}
	Check(adacom_column_create(state.ctx, GpuTypeOf(segment.type.InternalType()), 1, &count, capacity, &state.column),
	      "column_create");
	if (count) {
		if (!state.slots) {
			throw InternalException("succinct GPU codec: segment has neither host slots nor a device column");
		}
		Check(adacom_column_upload_raw(state.column, state.slots.get(), 0, count), "upload");
	}
	state.resident = count;
	state.data_size = 0;
	state.compacted_for = (idx_t)-1;
}

unique_ptr<CompressedSegmentState> SuccinctGpuInitSegment(ColumnSegment &segment, block_id_t block_id) {
	// the slot array is allocated by the first append (scan-only segments never need one)
	return make_unique<SuccinctGpuSegmentState>();
}

//===--------------------------------------------------------------------===//
// Analyze (checkpoint entry points; the reference only counts rows, succinct.cpp:19-42)
//===--------------------------------------------------------------------===//
struct SuccinctGpuAnalyzeState : public AnalyzeState {
	idx_t count = 0;
	uint64_t min_factor = NumericLimits<uint64_t>::Maximum();
	uint64_t max_factor = 0;
};

unique_ptr<AnalyzeState> SuccinctGpuInitAnalyze(ColumnData &col_data, PhysicalType type) {
	return make_unique<SuccinctGpuAnalyzeState>();
}

template <class T>
bool SuccinctGpuAnalyze(AnalyzeState &state_p, Vector &input, idx_t count) {
	auto &state = (SuccinctGpuAnalyzeState &)state_p;
	UnifiedVectorFormat vdata;
	input.ToUnifiedFormat(count, vdata);
	auto values = (T *)vdata.data;
	for (idx_t i = 0; i < count; i++) {
		auto idx = vdata.sel->get_index(i);
		if (vdata.validity.RowIsValid(idx)) { // NULLs do not take part in min/max (succinct.cpp:284-292)
			state.min_factor = MinValue<uint64_t>(state.min_factor, uint64_t(values[idx]));
			state.max_factor = MaxValue<uint64_t>(state.max_factor, uint64_t(values[idx]));
		}
	}
	state.count += count;
	return true;
}

//! score = the bytes the column will occupy: 9 + 8 * ceil(n * w / 64) per segment (sdsl size_in_bytes)
template <class T>
idx_t SuccinctGpuFinalAnalyze(AnalyzeState &state_p) {
	auto &state = (SuccinctGpuAnalyzeState &)state_p;
	uint64_t range = state.max_factor >= state.min_factor ? state.max_factor - state.min_factor : 0;
	idx_t width = 1;
	while (width < 64 && (range >> width) != 0) {
		width++;
	}
	if (width >= 8 * sizeof(T)) {
		return sizeof(T) * state.count; // no re-pack (column_segment.cpp:363)
	}
	idx_t per_segment = Storage::BLOCK_SIZE / sizeof(T);
	idx_t segments = (state.count + per_segment - 1) / per_segment;
	return 9 * segments + 8 * ((state.count * width + 63) / 64);
}

//===--------------------------------------------------------------------===//
// Compress (checkpoint entry points): append into transient succinct segments, flush each full one
//===--------------------------------------------------------------------===//
struct SuccinctGpuCompressState : public CompressionState {
	explicit SuccinctGpuCompressState(ColumnDataCheckpointer &checkpointer_p) : checkpointer(checkpointer_p) {
		NewSegment(checkpointer.GetRowGroup().start);
	}

	void NewSegment(idx_t row_start) {
		segment = ColumnSegment::CreateTransientSegment(checkpointer.GetDatabase(), checkpointer.GetType(), row_start);
		segment->InitializeAppend(append_state);
	}
	void Flush() {
		idx_t bytes = segment->FinalizeAppend(append_state);
		checkpointer.GetCheckpointState().FlushSegment(move(segment), bytes);
	}

	ColumnDataCheckpointer &checkpointer;
	unique_ptr<ColumnSegment> segment;
	ColumnAppendState append_state;
};

unique_ptr<CompressionState> SuccinctGpuInitCompression(ColumnDataCheckpointer &checkpointer,
                                                        unique_ptr<AnalyzeState> state) {
	return make_unique<SuccinctGpuCompressState>(checkpointer);
}

void SuccinctGpuCompress(CompressionState &state_p, Vector &data, idx_t count) {
	auto &state = (SuccinctGpuCompressState &)state_p;
	UnifiedVectorFormat vdata;
	data.ToUnifiedFormat(count, vdata);
	idx_t offset = 0;
	while (count > 0) {
		idx_t appended = state.segment->Append(state.append_state, vdata, offset, count);
		offset += appended;
		count -= appended;
		if (count > 0) { // segment full: hand it to the checkpoint state and open the next one
			idx_t next_start = state.segment->start + state.segment->count;
			state.Flush();
			state.NewSegment(next_start);
		}
	}
}

void SuccinctGpuFinalizeCompress(CompressionState &state_p) {
	auto &state = (SuccinctGpuCompressState &)state_p;
	state.Flush();
}

//===--------------------------------------------------------------------===//
// Scan: one batched decode of the segment per scan state, then 2048-value slices
//===--------------------------------------------------------------------===//
//! pinned host buffers for the per-segment decode (cudaHostAlloc is expensive: keep them)
struct PinnedPool {
	std::mutex lock;
	std::vector<std::pair<data_ptr_t, idx_t>> free_list;

	data_ptr_t Get(adacom_ctx *ctx, idx_t bytes, idx_t &capacity) {
		{
			std::lock_guard<std::mutex> guard(lock);
			for (idx_t i = 0; i < free_list.size(); i++) {
				if (free_list[i].second >= bytes) {
					auto entry = free_list[i];
					free_list[i] = free_list.back();
					free_list.pop_back();
					capacity = entry.second;
					return entry.first;
				}
			}
		}
		void *ptr = nullptr;
		capacity = MaxValue<idx_t>(bytes, Storage::BLOCK_ALLOC_SIZE);
		Check(adacom_host_alloc(ctx, capacity, &ptr), "host_alloc");
		return (data_ptr_t)ptr;
	}
	void Put(data_ptr_t ptr, idx_t capacity) {
		std::lock_guard<std::mutex> guard(lock);
		free_list.emplace_back(ptr, capacity);
	}
};

PinnedPool &Pool() {
	static PinnedPool pool;
	return pool;
}

struct SuccinctGpuScanState : public SegmentScanState {
	~SuccinctGpuScanState() override {
		if (decoded) {
			Pool().Put(decoded, capacity);
		}
	}
	//! the decoded segment (count * sizeof(T) bytes, pinned), filled by the first scan call
	data_ptr_t decoded = nullptr;
	idx_t capacity = 0;
	idx_t decoded_count = 0;
	//! the segment the buffer belongs to (a scan state moves from segment to segment)
	ColumnSegment *owner = nullptr;
};

unique_ptr<SegmentScanState> SuccinctGpuInitScan(ColumnSegment &segment) {
	return make_unique<SuccinctGpuScanState>();
}

template <class T>
void SuccinctGpuScanPartial(ColumnSegment &segment, ColumnScanState &state, idx_t scan_count, Vector &result,
                            idx_t result_offset) {
	auto &seg_state = StateOf(segment);
	auto start = segment.GetRelativeIndex(state.row_index);
	result.SetVectorType(VectorType::FLAT_VECTOR);
	auto target = FlatVector::GetData(result);
	auto scan = (SuccinctGpuScanState *)state.scan_state.get();
	if (scan && scan->owner == &segment && scan->decoded_count == segment.count) {
		// every pull after the first: a slice of the decoded segment
		StatTimer timer(Stats().slice_ns, Stats().slice_calls);
		memcpy(target + result_offset * sizeof(T), scan->decoded + start * sizeof(T), scan_count * sizeof(T));
		return;
	}
	std::lock_guard<std::mutex> guard(seg_state.lock);
	EnsureResident(segment, seg_state);
	if (!scan) { // no scan state (index fetch paths): decode just the requested slice
		Check(adacom_scan_partial(seg_state.column, 0, start, scan_count, 0, target, result_offset, 0), "scan_partial");
		return;
	}
	// first pull of this segment: ONE launch decodes all of it into a pinned buffer
	StatTimer timer(Stats().decode_ns, Stats().decode_calls);
	idx_t bytes = segment.count * sizeof(T) + 8;
	if (scan->decoded && scan->capacity < bytes) {
		Pool().Put(scan->decoded, scan->capacity);
		scan->decoded = nullptr;
	}
	if (!scan->decoded) {
		scan->decoded = Pool().Get(seg_state.ctx, bytes, scan->capacity);
	}
	Check(adacom_decode(seg_state.column, 0, 1, 0, scan->decoded, 0), "decode");
	scan->decoded_count = segment.count;
	scan->owner = &segment;
	memcpy(target + result_offset * sizeof(T), scan->decoded + start * sizeof(T), scan_count * sizeof(T));
}

template <class T>
void SuccinctGpuScan(ColumnSegment &segment, ColumnScanState &state, idx_t scan_count, Vector &result) {
	SuccinctGpuScanPartial<T>(segment, state, scan_count, result, 0);
}

//===--------------------------------------------------------------------===//
// Fetch (the intended semantics of succinct.cpp:244-260: ONE value, the one at row_id)
//===--------------------------------------------------------------------===//
template <class T>
void SuccinctGpuFetchRow(ColumnSegment &segment, ColumnFetchState &state, row_t row_id, Vector &result,
                         idx_t result_idx) {
	auto &seg_state = StateOf(segment);
	std::lock_guard<std::mutex> guard(seg_state.lock);
	EnsureResident(segment, seg_state);
	uint32_t seg = 0, row = (uint32_t)row_id;
	Check(adacom_fetch_rows(seg_state.column, 1, &seg, &row, 0, FlatVector::GetData(result) + result_idx * sizeof(T), 0),
	      "fetch_rows");
}

//===--------------------------------------------------------------------===//
// Append (succinct.cpp:262-330): slots on the host while the segment is open, min/max kept on the segment
//===--------------------------------------------------------------------===//
unique_ptr<CompressionAppendState> SuccinctGpuInitAppend(ColumnSegment &segment) {
	auto &buffer_manager = BufferManager::GetBufferManager(segment.db);
	return make_unique<CompressionAppendState>(buffer_manager.Pin(segment.block));
}

template <class T>
idx_t SuccinctGpuAppend(CompressionAppendState &append_state, ColumnSegment &segment, SegmentStatistics &stats,
                        UnifiedVectorFormat &data, idx_t offset, idx_t count) {
	auto &seg_state = StateOf(segment);
	idx_t capacity = segment.SegmentSize() / sizeof(T);
	idx_t copied = MinValue<idx_t>(count, capacity - segment.count);
	auto source = (T *)data.data;
	if (!seg_state.slots) {
		seg_state.slots = unique_ptr<data_t[]>(new data_t[segment.SegmentSize()]);
	}
	auto target = (T *)seg_state.slots.get() + segment.count;
	uint64_t min_factor = NumericLimits<uint64_t>::Maximum(), max_factor = 0;
	if (data.validity.AllValid()) {
		for (idx_t i = 0; i < copied; i++) {
			auto idx = data.sel->get_index(offset + i);
			NumericStatistics::Update<T>(stats, source[idx]);
			target[i] = source[idx];
			min_factor = MinValue<uint64_t>(min_factor, uint64_t(source[idx]));
			max_factor = MaxValue<uint64_t>(max_factor, uint64_t(source[idx]));
		}
	} else {
		for (idx_t i = 0; i < copied; i++) {
			auto idx = data.sel->get_index(offset + i);
			if (data.validity.RowIsValid(idx)) {
				NumericStatistics::Update<T>(stats, source[idx]);
				target[i] = source[idx];
				min_factor = MinValue<uint64_t>(min_factor, uint64_t(source[idx]));
				max_factor = MaxValue<uint64_t>(max_factor, uint64_t(source[idx]));
			} else {
				target[i] = NullValue<T>(); // the NULL gap (succinct.cpp:288-292)
				if (seg_state.validity.empty()) {
					seg_state.validity.assign((capacity + 63) / 64, ~uint64_t(0));
				}
				idx_t row = segment.count + i;
				seg_state.validity[row >> 6] &= ~(uint64_t(1) << (row & 63));
			}
		}
	}
	segment.UpdateMinFactor(min_factor);
	segment.UpdateMaxFactor(max_factor);
	segment.count += copied; // Q6: rows copied, not rows requested
	return copied;
}

//! succinct.cpp:324-330: bytes of the segment.  Compaction is triggered by ColumnSegment::Append (segment full,
//! column_segment.cpp:266-268), by the first scan (:157-160) or by the policy -- never here.
template <class T>
idx_t SuccinctGpuFinalizeAppend(ColumnSegment &segment, SegmentStatistics &stats) {
	return segment.count * sizeof(T);
}

template <class T>
CompressionFunction SuccinctGpuGetFunction(PhysicalType type) {
	return CompressionFunction(CompressionType::COMPRESSION_SUCCINCT, type, SuccinctGpuInitAnalyze,
	                           SuccinctGpuAnalyze<T>, SuccinctGpuFinalAnalyze<T>, SuccinctGpuInitCompression,
	                           SuccinctGpuCompress, SuccinctGpuFinalizeCompress, SuccinctGpuInitScan, SuccinctGpuScan<T>,
	                           SuccinctGpuScanPartial<T>, SuccinctGpuFetchRow<T>, UncompressedFunctions::EmptySkip,
	                           SuccinctGpuInitSegment, SuccinctGpuInitAppend, SuccinctGpuAppend<T>,
	                           SuccinctGpuFinalizeAppend<T>, nullptr);
}

} // namespace

//===--------------------------------------------------------------------===//
// Public: function table + the policy-side bodies
//===--------------------------------------------------------------------===//
CompressionFunction SuccinctGpuFun::GetFunction(PhysicalType type) {
	switch (type) {
	case PhysicalType::INT8:
		return SuccinctGpuGetFunction<int8_t>(type);
	case PhysicalType::UINT8:
		return SuccinctGpuGetFunction<uint8_t>(type);
	case PhysicalType::INT16:
		return SuccinctGpuGetFunction<int16_t>(type);
	case PhysicalType::UINT16:
		return SuccinctGpuGetFunction<uint16_t>(type);
	case PhysicalType::INT32:
		return SuccinctGpuGetFunction<int32_t>(type);
	case PhysicalType::UINT32:
		return SuccinctGpuGetFunction<uint32_t>(type);
	case PhysicalType::INT64:
		return SuccinctGpuGetFunction<int64_t>(type);
	case PhysicalType::UINT64:
		return SuccinctGpuGetFunction<uint64_t>(type);
	default: // INT128 included (Q6): stays uncompressed
		throw InternalException("Unsupported type for SuccinctGpuFun::GetFunction");
	}
}

bool SuccinctGpuFun::TypeIsSupported(PhysicalType type) {
	switch (type) {
	case PhysicalType::INT8:
	case PhysicalType::UINT8:
	case PhysicalType::INT16:
	case PhysicalType::UINT16:
	case PhysicalType::INT32:
	case PhysicalType::UINT32:
	case PhysicalType::INT64:
	case PhysicalType::UINT64:
		return true;
	default:
		return false;
	}
}

idx_t SuccinctGpuFun::SlotWidthSize(idx_t segment_size, idx_t type_size) {
	// sdsl::size_in_bytes of the slot-width int_vector the reference allocates (column_segment.cpp:101-105)
	idx_t bits = (segment_size / type_size) * type_size * 8;
	return 9 + 8 * ((bits + 63) / 64);
}

unique_ptr<CompressedSegmentState> SuccinctGpuFun::InitSegmentFromBlock(ColumnSegment &segment, data_ptr_t block_ptr) {
	auto state = make_unique<SuccinctGpuSegmentState>();
	state->ctx = GpuContextFor(segment.start);
	uint64_t count = segment.count;
	auto capacity = (uint32_t)(segment.SegmentSize() / segment.type_size);
	Check(adacom_column_create(state->ctx, GpuTypeOf(segment.type.InternalType()), 1, &count, capacity, &state->column),
	      "column_create");
	if (count) {
		Check(adacom_column_upload_raw(state->column, block_ptr, 0, count), "upload");
		Check(adacom_ctx_synchronize(state->ctx), "synchronize"); // the block may be evicted once we return
	}
	state->resident = count;
	return move(state);
}

idx_t SuccinctGpuFun::Compact(ColumnSegment &segment, bool zero_extended_minmax) {
	auto &seg_state = StateOf(segment);
	std::lock_guard<std::mutex> guard(seg_state.lock);
	if (seg_state.column && seg_state.resident == segment.count && seg_state.compacted_for == segment.count &&
	    seg_state.data_size) {
		return seg_state.data_size; // already compacted at this size (a concurrent or repeated Compact())
	}
	StatTimer timer(Stats().compact_ns, Stats().compact_calls);
	EnsureResident(segment, seg_state);
	int flags = 0;
	if (DBConfig::GetConfig(segment.db).succinct_padded_to_next_byte_enabled) {
		flags |= ADACOM_PADDED; // column_segment.cpp:356-359
	}
	if (zero_extended_minmax) {
		flags |= ADACOM_ZERO_EXT_MINMAX; // BitCompressFromUncompressed, column_segment.cpp:390-400
	}
	// NULL rows stay out of min/max on the append path (succinct.cpp:284-292); the adaptive encoder looks at every
	// slot of the block (column_segment.cpp:392-399), so it gets no mask
	const uint64_t *validity = (!zero_extended_minmax && !seg_state.validity.empty()) ? seg_state.validity.data() : nullptr;
	Check(adacom_analyze(seg_state.column, 0, 1, flags, validity, 0), "analyze");
	Check(adacom_encode(seg_state.column, 0, 1, flags), "encode");
	uint64_t bytes = 0;
	Check(adacom_column_data_size(seg_state.column, 0, 1, &bytes), "data_size");
	seg_state.data_size = bytes;
	seg_state.compacted_for = segment.count;
	adacom_segment_meta meta;
	Check(adacom_column_get_meta(seg_state.column, 0, 1, &meta), "get_meta");
	if (meta.state & ADACOM_SEG_REPACKED) {
		// the payload lives in the packed arena now: drop the slot copies on both sides
		Check(adacom_column_release_raw(seg_state.column), "release_raw");
		seg_state.slots.reset();
	}
	return bytes;
}

void SuccinctGpuFun::UncompactToBlock(ColumnSegment &segment, data_ptr_t block_ptr) {
	auto &seg_state = StateOf(segment);
	std::lock_guard<std::mutex> guard(seg_state.lock);
	EnsureResident(segment, seg_state);
	if (segment.count) {
		Check(adacom_decode(seg_state.column, 0, 1, 0, block_ptr, 0), "decode");
	}
}

idx_t SuccinctGpuFun::DataSize(ColumnSegment &segment) {
	auto state = (SuccinctGpuSegmentState *)segment.GetSegmentState();
	if (!state || !state->column || !state->data_size) {
		// open segment: the slot-width vector of the reference
		return SlotWidthSize(segment.SegmentSize(), segment.type_size);
	}
	return state->data_size;
}

} // namespace duckdb
