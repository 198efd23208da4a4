// succinct_gpu.cpp -- SuccinctGpuFun (the CompressionFunction vtable) and ColumnSegment re-hosted on the GPU
// codec.  Control flow follows the reference call sites one to one; every codec loop is a C-ABI call.
//
//   SuccinctFun::GetFunction / TypeIsSupported ... src/storage/compression/succinct.cpp:335-382
//   SuccinctScan / SuccinctScanPartial ............ succinct.cpp:123-144, 232-240
//   SuccinctFetchRow (intended semantics, Q3) ..... succinct.cpp:244-260
//   SuccinctInitAppend / Append / FinalizeAppend .. succinct.cpp:264-330
//   InitCompression / Compress / FinalizeCompress . succinct.cpp:86-119
//   ColumnSegment::{CreateTransientSegment, Scan, ScanPartial, FetchRow, Append, Compact, Uncompact,
//   GetDataSize} .................................. src/storage/table/column_segment.cpp:45-346
#include "adacom_host.hpp"

#include <algorithm>
#include <limits>

namespace adacom_host {

// ---------------------------------------------------------------------------------------------------
// types
// ---------------------------------------------------------------------------------------------------

idx_t GetTypeIdSize(PhysicalType type) {
	switch (type) {
	case PhysicalType::BOOL:
	case PhysicalType::UINT8:
	case PhysicalType::INT8:
		return 1;
	case PhysicalType::UINT16:
	case PhysicalType::INT16:
		return 2;
	case PhysicalType::UINT32:
	case PhysicalType::INT32:
	case PhysicalType::FLOAT:
		return 4;
	case PhysicalType::UINT64:
	case PhysicalType::INT64:
	case PhysicalType::DOUBLE:
		return 8;
	case PhysicalType::INT128:
	case PhysicalType::VARCHAR:
		return 16;
	default:
		throw InternalException("Invalid PhysicalType for GetTypeIdSize");
	}
}

bool TypeIsInteger(PhysicalType type) {
	return (type >= PhysicalType::UINT8 && type <= PhysicalType::INT64) || type == PhysicalType::INT128;
}

adacom_ptype ToAdacomType(PhysicalType type) {
	if (!SuccinctGpuFun::TypeIsSupported(type)) {
		throw InternalException("Unsupported type for FixedSizeSuccinct::GetFunction");
	}
	return (adacom_ptype)(uint8_t)type;  // the C ABI uses duckdb's PhysicalType values
}

// ---------------------------------------------------------------------------------------------------
// database instance
// ---------------------------------------------------------------------------------------------------

DatabaseInstance::DatabaseInstance(const DBConfig *cfg)
    : column_segment_catalog(cfg ? cfg->compression_rate : 0.90, cfg ? cfg->background_interval_s : 10.0) {
	if (cfg) {
		config.succinct_enabled = cfg->succinct_enabled;
		config.succinct_extract_prefix_enabled = cfg->succinct_extract_prefix_enabled;
		config.succinct_padded_to_next_byte_enabled = cfg->succinct_padded_to_next_byte_enabled;
		config.adaptive_succinct_compression_enabled = cfg->adaptive_succinct_compression_enabled;
		config.gpu_device = cfg->gpu_device;
		config.gpu_devices = cfg->gpu_devices;
		config.compression_rate = cfg->compression_rate;
		config.background_interval_s = cfg->background_interval_s;
	}
}

DatabaseInstance::~DatabaseInstance() {
	column_segment_catalog.StopBackgroundThread();
	for (auto &kv : gpus) {
		adacom_ctx_destroy(kv.second);
	}
}

adacom_ctx *DatabaseInstance::Gpu() {
	return Gpu(config.gpu_device);
}

adacom_ctx *DatabaseInstance::Gpu(int device) {
	if (device < 0) {
		device = config.gpu_device;
	}
	std::lock_guard<std::mutex> guard(gpu_lock);
	auto it = gpus.find(device);
	if (it != gpus.end()) {
		return it->second;
	}
	adacom_ctx *ctx = nullptr;
	adacom_status st = adacom_ctx_create(device, &ctx);
	if (st != ADACOM_OK) {
		throw InternalException(std::string("succinct GPU codec unavailable: ") + adacom_last_error(nullptr));
	}
	gpus[device] = ctx;
	if (device == config.gpu_device) {
		gpu = ctx;
	}
	return ctx;
}

void DatabaseInstance::Throw(const char *what) {
	std::string msg = std::string(what) + ": " + (gpu ? adacom_last_error(gpu) : adacom_last_error(nullptr));
	throw InternalException(msg);
}

#define GPU_CHECK(db, call)              \
	do {                                 \
		adacom_status st__ = (call);     \
		if (st__ == ADACOM_ERR_NOMEM) {  \
			throw std::bad_alloc();      \
		}                                \
		if (st__ != ADACOM_OK) {         \
			(db).Throw(#call);           \
		}                                \
	} while (0)

CompressionFunction *DBConfig::GetCompressionFunction(CompressionType type, PhysicalType data_type) {
	std::lock_guard<std::mutex> guard(lock);  // compression_config.cpp:85-94
	uint32_t key = ((uint32_t)type << 8) | (uint32_t)data_type;
	auto it = functions.find(key);
	if (it != functions.end()) {
		return it->second.get();
	}
	std::unique_ptr<CompressionFunction> fn(new CompressionFunction(SuccinctGpuFun::GetFunction(data_type)));
	fn->type = type;  // the UNCOMPRESSED entry shares the slots; segments in that state read their flat block
	auto *raw = fn.get();
	functions[key] = std::move(fn);
	return raw;
}

// ---------------------------------------------------------------------------------------------------
// ColumnSegment
// ---------------------------------------------------------------------------------------------------

static idx_t SlotVectorBytes(idx_t segment_size) {
	// sdsl::size_in_bytes of the slot-width int_vector allocated by the ctor (column_segment.cpp:101-105)
	return 9 + 8 * ((segment_size * 8 + 63) / 64);
}

unique_ptr<ColumnSegment> ColumnSegment::CreateTransientSegment(DatabaseInstance &db, PhysicalType type, idx_t start,
                                                                idx_t segment_size) {
	auto &config = db.config;
	bool integer = TypeIsInteger(type);
	if (integer && !SuccinctGpuFun::TypeIsSupported(type)) {
		integer = false;  // INT128 passes TypeIsInteger but has no succinct function (Q6): stays uncompressed
	}
	CompressionFunction *function;
	bool succinct_possible;
	if (integer && config.succinct_enabled && !config.adaptive_succinct_compression_enabled) {
		succinct_possible = true;
		function = config.GetCompressionFunction(CompressionType::COMPRESSION_SUCCINCT, type);
	} else {
		succinct_possible = integer && config.succinct_enabled;
		if (!integer) {
			throw InternalException("adacom_host only hosts integer column segments");
		}
		function = config.GetCompressionFunction(CompressionType::COMPRESSION_UNCOMPRESSED, type);
		db.buffer_manager.AddToDataSize((int64_t)segment_size);
	}
	return unique_ptr<ColumnSegment>(new ColumnSegment(db, type, start, segment_size, function, succinct_possible,
	                                                   config.adaptive_succinct_compression_enabled, integer));
}

ColumnSegment::ColumnSegment(DatabaseInstance &db_p, PhysicalType type_p, idx_t start_p, idx_t segment_size_p,
                             CompressionFunction *function_p, bool succinct_possible_p, bool background,
                             bool is_data_segment_p)
    : db(db_p), type(type_p), type_size(GetTypeIdSize(type_p)), start(start_p), function(function_p),
      segment_size(segment_size_p), succinct_possible(succinct_possible_p), background_compaction_enabled(background),
      is_data_segment(is_data_segment_p) {
	host_block->resize(segment_size);
	staging_validity.assign((segment_size / type_size + 63) / 64, ~0ull);
	if (function->type == CompressionType::COMPRESSION_SUCCINCT) {
		db.buffer_manager.AddToDataSize((int64_t)SlotVectorBytes(segment_size));
	}
	db.column_segment_catalog.AddColumnSegment(this);
}

ColumnSegment::~ColumnSegment() {
	db.column_segment_catalog.RemoveColumnSegment(this);
	if (gpu_column && owns_gpu_column) {
		adacom_column_destroy(gpu_column);
	}
}

SegmentBatchOwner *ColumnSegment::BatchOwner() const {
	return column_data;
}

int ColumnSegment::Device() const {
	return device < 0 ? db.config.gpu_device : device;
}

void ColumnSegment::EnsureDeviceColumn() {
	if (!gpu_column) {
		uint64_t cnt = count;
		GPU_CHECK(db, adacom_column_create(db.Gpu(device), ToAdacomType(type), 1, &cnt,
		                                   (uint32_t)(segment_size / type_size), &gpu_column));
		gpu_index = 0;
		owns_gpu_column = true;
		uploaded = 0;
	}
}

void ColumnSegment::EnsureStaging() {
	// a member of a shared device column gives its host block back after the upload; appends (and host-side
	// reads of an uncompacted segment) bring it back from the device
	if (HostBlock()->size() == segment_size) {
		return;
	}
	auto block = std::make_shared<std::vector<uint8_t>>(segment_size);
	if (gpu_column && count) {
		GPU_CHECK(db, adacom_decode(gpu_column, gpu_index, gpu_index + 1, 0, block->data(), 0));
	}
	std::lock_guard<std::mutex> guard(bit_compression_lock);
	host_block.swap(block);
}

void ColumnSegment::DetachFromBatch() {
	// the segment is about to change (append after Uncompact): its slot in the shared column goes stale, the
	// segment continues on its own
	if (gpu_column && !owns_gpu_column) {
		EnsureStaging();
		gpu_column = nullptr;
		gpu_index = 0;
		owns_gpu_column = true;
		uploaded = 0;
	}
}

void ColumnSegment::Upload() {
	// the device column is created at Compact() time with the final count (segments are append-only until then)
	if (gpu_column && owns_gpu_column && adacom_column_num_values(gpu_column) != count) {
		adacom_column_destroy(gpu_column);
		gpu_column = nullptr;
	}
	EnsureDeviceColumn();
	if (uploaded < count) {
		GPU_CHECK(db, adacom_column_upload_raw(gpu_column, host_block->data() + uploaded * type_size, uploaded,
		                                       count - uploaded));
		uploaded = count;
	}
}

adacom_segment_meta ColumnSegment::Meta() const {
	adacom_segment_meta m;
	memset(&m, 0, sizeof(m));
	if (gpu_column) {
		adacom_status st = adacom_column_get_meta(gpu_column, gpu_index, gpu_index + 1, &m);
		if (st != ADACOM_OK) {
			const_cast<DatabaseInstance &>(db).Throw("adacom_column_get_meta");
		}
	}
	return m;
}

uint64_t ColumnSegment::GetMinFactor() {
	return gpu_column ? Meta().min_factor : UINT64_MAX;
}
uint64_t ColumnSegment::GetMax() {
	return gpu_column ? Meta().max_factor : 0;
}
uint8_t ColumnSegment::Width() {
	return gpu_column ? Meta().width : (uint8_t)(8 * type_size);
}

void ColumnSegment::InitializeScan(ColumnScanState &state) {
	state.scan_state = function->init_scan(*this);
}

void ColumnSegment::Scan(ColumnScanState &state, idx_t scan_count, Vector &result, idx_t result_offset,
                         bool entire_vector) {
	if (entire_vector) {
		Scan(state, scan_count, result);
	} else {
		ScanPartial(state, scan_count, result, result_offset);
	}
}

void ColumnSegment::Skip(ColumnScanState &state) {
	function->skip(*this, state, state.row_index - state.internal_index);
	state.internal_index = state.row_index;
}

void ColumnSegment::Scan(ColumnScanState &state, idx_t scan_count, Vector &result) {
	db.column_segment_catalog.AddReadAccess(this);
	if (!compacted && !background_compaction_enabled) {
		Compact();  // lazy encode on first read (column_segment.cpp:157-160); a ColumnData flushes its whole batch
	}
	{
		std::lock_guard<std::mutex> guard(bit_compression_lock);
		if (force_reinitializing_scan_state || !state.scan_state) {
			InitializeScan(state);
			force_reinitializing_scan_state = false;
		}
	}
	function->scan_vector(*this, state, scan_count, result);
}

void ColumnSegment::ScanPartial(ColumnScanState &state, idx_t scan_count, Vector &result, idx_t result_offset) {
	db.column_segment_catalog.AddReadAccess(this);
	if (!compacted && !background_compaction_enabled) {
		Compact();
	}
	{
		std::lock_guard<std::mutex> guard(bit_compression_lock);
		if (force_reinitializing_scan_state || !state.scan_state) {
			InitializeScan(state);
			force_reinitializing_scan_state = false;
		}
	}
	function->scan_partial(*this, state, scan_count, result, result_offset);
}

void ColumnSegment::FetchRow(ColumnFetchState &state, row_t row_id, Vector &result, idx_t result_idx) {
	function->fetch_row(*this, state, row_id - (row_t)this->start, result, result_idx);
}

void ColumnSegment::InitializeAppend(ColumnAppendState &state) {
	if (!function->init_append) {
		throw InternalException("Attempting to init append to a segment without init_append method");
	}
	state.append_state = function->init_append(*this);
}

idx_t ColumnSegment::Append(ColumnAppendState &state, UnifiedVectorFormat &append_data, idx_t offset, idx_t count_p) {
	if (!function->append) {
		throw InternalException("Attempting to append to a segment without append method");
	}
	bool uncompacted = false;
	if (IsBitCompressed()) {
		Uncompact();  // decode back to flat before appending (column_segment.cpp:253-259)
		InitializeAppend(state);
		uncompacted = true;
	}
	if (count < segment_size / type_size) {
		DetachFromBatch();  // the slot in the shared device column is fixed-size and about to go stale
		EnsureStaging();
	}
	idx_t copy_count = function->append(*state.append_state, *this, stats, append_data, offset, count_p);
	num_elements += count_p;  // requested, not copied: quirk Q6 kept for the trigger below
	if (!compacted && !background_compaction_enabled && (num_elements >= segment_size / type_size || uncompacted)) {
		if (column_data && !gpu_column && !uncompacted) {
			column_data->SegmentSealed();  // deferred: the column compacts its sealed segments a batch at a time
		} else {
			Compact();
		}
	}
	return copy_count;
}

idx_t ColumnSegment::FinalizeAppend(ColumnAppendState &) {
	if (!function->finalize_append) {
		throw InternalException("Attempting to call FinalizeAppend on a segment without a finalize_append method");
	}
	return function->finalize_append(*this, stats);
}

void ColumnSegment::Compact() {
	if (compacted || !function || num_elements == 0 || !succinct_possible) {
		return;  // column_segment.cpp:278
	}
	if (column_data) {
		// member of a ColumnData: compaction is a batch operation of the column (one launch set for all pending
		// segments, or a one-element list for a member that the policy uncompacted earlier)
		if (!gpu_column) {
			column_data->CompactPending(true);
		} else {
			column_data->CompactMany({this});
		}
		return;
	}
	const int flags = CompactFlags();
	const idx_t size_before = GetDataSize();
	Upload();
	// analyze: the min/max SuccinctAppendLoop tracks while appending, taken by the reduction kernel
	GPU_CHECK(db, adacom_analyze(gpu_column, 0, 1, flags, staging_has_nulls ? staging_validity.data() : nullptr, 0));
	// encode: BitCompressFromSuccinct / BitCompressFromUncompressed
	GPU_CHECK(db, adacom_encode(gpu_column, 0, 1, flags));
	FinishCompact(size_before, Meta());
}

int ColumnSegment::CompactFlags() const {
	int flags = db.config.succinct_padded_to_next_byte_enabled ? ADACOM_PADDED : 0;
	if (function->type != CompressionType::COMPRESSION_SUCCINCT) {
		flags |= ADACOM_ZERO_EXT_MINMAX;  // BitCompressFromUncompressed scans zero-extended raw bytes (Q2b)
	}
	return flags;
}

void ColumnSegment::FinishCompact(idx_t size_before, const adacom_segment_meta &m) {
	{
		std::lock_guard<std::mutex> guard(bit_compression_lock);
		if (function->type != CompressionType::COMPRESSION_SUCCINCT) {
			function = db.config.GetCompressionFunction(CompressionType::COMPRESSION_SUCCINCT, type);
		}
		compacted = true;
		epoch++;
	}
	compacted_size = m.size_in_bytes;
	if (m.min_factor != UINT64_MAX) {
		stats.has_value = true;
		stats.umin = m.min_factor;
		stats.umax = m.max_factor;
		stats.min = (int64_t)m.min_factor;
		stats.max = (int64_t)m.max_factor;
	}
	db.buffer_manager.AddToDataSize((int64_t)m.size_in_bytes - (int64_t)size_before);
}

void ColumnSegment::Uncompact() {
	if (!compacted || !function || function->type != CompressionType::COMPRESSION_SUCCINCT) {
		return;  // column_segment.cpp:327-330
	}
	if (column_data && gpu_column && !owns_gpu_column) {
		column_data->UncompactMany({this});
		return;
	}
	const idx_t compressed_size = GetDataSize();
	GPU_CHECK(db, adacom_uncompact(gpu_column, gpu_index, gpu_index + 1, 0));
	// the flat block lives on the host again (appends continue there)
	{
		auto block = std::make_shared<std::vector<uint8_t>>(segment_size);
		GPU_CHECK(db, adacom_decode(gpu_column, gpu_index, gpu_index + 1, 0, block->data(), 0));
		std::lock_guard<std::mutex> guard(bit_compression_lock);
		host_block.swap(block);
	}
	FinishUncompact(compressed_size);
}

void ColumnSegment::FinishUncompact(idx_t compressed_size) {
	{
		std::lock_guard<std::mutex> guard(bit_compression_lock);
		function = db.config.GetCompressionFunction(CompressionType::COMPRESSION_UNCOMPRESSED, type);
		compacted = false;
		force_reinitializing_scan_state = true;
		epoch++;
	}
	db.buffer_manager.AddToDataSize((int64_t)segment_size - (int64_t)compressed_size);
}

idx_t ColumnSegment::GetDataSize() const {
	if (!is_data_segment) {
		return 0;
	}
	if (function->type == CompressionType::COMPRESSION_SUCCINCT) {
		if (compacted && gpu_column) {
			return compacted_size;
		}
		return SlotVectorBytes(segment_size);
	}
	return segment_size;
}

idx_t ColumnSegment::SuccinctSize() const {
	return function->type == CompressionType::COMPRESSION_SUCCINCT ? GetDataSize() : 0;
}

// ---------------------------------------------------------------------------------------------------
// the plugin entry points
// ---------------------------------------------------------------------------------------------------

namespace {

// ---- analyze: real meaning for the three checkpoint-style slots (the reference's are counting stubs,
// fixed_size_uncompressed.cpp:27-41): stage the vectors, run the analyze + layout kernels once, return the
// exact succinct size in bytes (sum of 9 + 8*ceil(n*w/64) over the segments the rows would fill).
struct SuccinctGpuAnalyzeState : public AnalyzeState {
	DatabaseInstance *db = nullptr;
	PhysicalType type = PhysicalType::INT64;
	std::vector<uint8_t> values;
	idx_t count = 0;
};

DatabaseInstance *g_analyze_db = nullptr;  // ColumnData is opaque here; tests bind the instance explicitly

unique_ptr<AnalyzeState> SuccinctInitAnalyze(ColumnData &, PhysicalType type) {
	auto state = unique_ptr<SuccinctGpuAnalyzeState>(new SuccinctGpuAnalyzeState());
	state->db = g_analyze_db;
	state->type = type;
	return unique_ptr<AnalyzeState>(state.release());
}

bool SuccinctAnalyze(AnalyzeState &state_p, Vector &input, idx_t count) {
	auto &state = (SuccinctGpuAnalyzeState &)state_p;
	idx_t S = GetTypeIdSize(state.type);
	state.values.insert(state.values.end(), input.buffer.data(), input.buffer.data() + count * S);
	state.count += count;
	return true;
}

idx_t SuccinctFinalAnalyze(AnalyzeState &state_p) {
	auto &state = (SuccinctGpuAnalyzeState &)state_p;
	idx_t S = GetTypeIdSize(state.type);
	if (!state.db || state.count == 0) {
		return S * state.count;  // the reference's score when no device is bound
	}
	DatabaseInstance &db = *state.db;
	idx_t cap = BLOCK_SIZE / S;
	std::vector<uint64_t> counts;
	for (idx_t left = state.count; left > 0;) {
		idx_t c = std::min(left, cap);
		counts.push_back(c);
		left -= c;
	}
	int flags = db.config.succinct_padded_to_next_byte_enabled ? ADACOM_PADDED : 0;
	adacom_column *col = nullptr;
	uint64_t total_words = 0;
	GPU_CHECK(db, adacom_compress_host(db.Gpu(), ToAdacomType(state.type), state.values.data(), counts.size(),
	                                   counts.data(), 0, flags, nullptr, nullptr, 0, &total_words, &col));
	uint64_t bytes = 0;
	adacom_status st = adacom_column_data_size(col, 0, counts.size(), &bytes);
	adacom_column_destroy(col);
	if (st != ADACOM_OK) {
		db.Throw("adacom_column_data_size");
	}
	return bytes;
}

// ---- compress: same shape as the reference's SuccinctCompressState (succinct.cpp:47-119) ---------------
struct SuccinctGpuCompressState : public CompressionState {
	ColumnDataCheckpointer &checkpointer;
	unique_ptr<ColumnSegment> current_segment;
	ColumnAppendState append_state;
	explicit SuccinctGpuCompressState(ColumnDataCheckpointer &cp) : checkpointer(cp) {
	}
};

}  // namespace

// the only pieces of ColumnDataCheckpointer the compress slots use (column_data_checkpointer.cpp:157-193)
class ColumnDataCheckpointer {
public:
	ColumnDataCheckpointer(DatabaseInstance &db, PhysicalType type, idx_t row_start)
	    : db(db), type(type), row_start(row_start) {
	}
	DatabaseInstance &db;
	PhysicalType type;
	idx_t row_start;
	std::vector<unique_ptr<ColumnSegment>> flushed;  // ColumnCheckpointState::FlushSegment's sink
};

namespace {

void CreateEmptySegment(SuccinctGpuCompressState &state, idx_t row_start) {
	state.current_segment = ColumnSegment::CreateTransientSegment(state.checkpointer.db, state.checkpointer.type, row_start);
	state.current_segment->InitializeAppend(state.append_state);
}

unique_ptr<CompressionState> SuccinctInitCompression(ColumnDataCheckpointer &checkpointer, unique_ptr<AnalyzeState>) {
	auto state = unique_ptr<SuccinctGpuCompressState>(new SuccinctGpuCompressState(checkpointer));
	CreateEmptySegment(*state, checkpointer.row_start);
	return unique_ptr<CompressionState>(state.release());
}

void SuccinctCompress(CompressionState &state_p, Vector &data, idx_t count) {
	auto &state = (SuccinctGpuCompressState &)state_p;
	UnifiedVectorFormat vdata;
	vdata.data = data.buffer.data();
	idx_t offset = 0;
	while (count > 0) {
		idx_t appended = state.current_segment->Append(state.append_state, vdata, offset, count);
		if (appended == count) {
			return;
		}
		auto next_start = state.current_segment->start + state.current_segment->count;
		state.current_segment->FinalizeAppend(state.append_state);
		state.current_segment->Compact();
		state.checkpointer.flushed.push_back(std::move(state.current_segment));
		CreateEmptySegment(state, next_start);
		offset += appended;
		count -= appended;
	}
}

void SuccinctFinalizeCompress(CompressionState &state_p) {
	auto &state = (SuccinctGpuCompressState &)state_p;
	state.current_segment->FinalizeAppend(state.append_state);
	state.current_segment->Compact();
	state.checkpointer.flushed.push_back(std::move(state.current_segment));
}

// ---- scan ---------------------------------------------------------------------------------------------
// A segment of a ColumnData is served by the column's scan stream (adacom_scan_stream): windows of 32 segments are
// decoded by ONE launch each and copied into a pinned ring that runs ahead of the pulls, across segment borders --
// the state of the next segment adopts the stream of the one the scan has just left (DuckDB keeps that state alive in
// ColumnScanState::previous_states until the next ScanVector).  A whole vector is handed out zero-copy
// (FlatVector::SetData, like FixedSizeScan); a partial one is a memcpy slice.  A lone segment decodes itself once
// (one launch + one D2H of <= 256 KiB) -- the reference re-copies the packed vector on every call (quirk Q4).
struct GpuScanStream {
	adacom_scan_stream *handle = nullptr;
	adacom_column *column = nullptr;
	uint64_t seg_begin = 0, seg_end = 0;
	~GpuScanStream() {
		adacom_scan_stream_close(handle);
	}
};

struct SuccinctGpuScanState : public SegmentScanState {
	std::shared_ptr<GpuScanStream> stream;
	std::shared_ptr<std::vector<uint8_t>> hot;  // host block of a hot segment a zero-copy vector points into
	const uint8_t *values = nullptr;  // this segment inside the stream's current window
	std::vector<uint8_t> decoded;     // lone segments
	uint64_t epoch = ~0ull;
};

unique_ptr<SegmentScanState> SuccinctInitScan(ColumnSegment &) {
	return unique_ptr<SegmentScanState>(new SuccinctGpuScanState());
}

const uint8_t *StreamValues(ColumnSegment &segment, ColumnScanState &state, SuccinctGpuScanState &scan_state) {
	if (scan_state.values) {
		return scan_state.values;
	}
	if (!scan_state.stream && !state.previous_states.empty() && state.previous_states.back()) {
		auto *prev = dynamic_cast<SuccinctGpuScanState *>(state.previous_states.back().get());
		if (prev && prev->stream && prev->stream->column == segment.gpu_column &&
		    segment.gpu_index >= prev->stream->seg_begin && segment.gpu_index < prev->stream->seg_end) {
			scan_state.stream = prev->stream;
		}
	}
	if (!scan_state.stream) {
		auto stream = std::make_shared<GpuScanStream>();
		stream->column = segment.gpu_column;
		stream->seg_begin = segment.gpu_index;
		stream->seg_end = adacom_column_num_segments(segment.gpu_column);
		GPU_CHECK(segment.db, adacom_scan_stream_open(segment.gpu_column, stream->seg_begin, stream->seg_end, 0, 0, 0,
		                                              &stream->handle));
		scan_state.stream = std::move(stream);
	}
	const void *values = nullptr;
	GPU_CHECK(segment.db, adacom_scan_stream_segment(scan_state.stream->handle, segment.gpu_index, &values, nullptr));
	scan_state.values = (const uint8_t *)values;
	return scan_state.values;
}

void SuccinctScanPartial(ColumnSegment &segment, ColumnScanState &state, idx_t scan_count, Vector &result,
                         idx_t result_offset) {
	auto start = segment.GetRelativeIndex(state.row_index);
	if (start + scan_count > segment.count) {
		throw InternalException("SuccinctScanPartial: scan runs past the end of the segment");
	}
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (result_offset == 0) {
		result.external = nullptr;  // a zero-copy view of the previous pull ends here
	}
	idx_t S = segment.type_size;
	data_ptr_t target_ptr = FlatVector::GetData(result) + result_offset * S;
	auto block = segment.HostBlock();  // keeps the block alive for the copy even if the policy drops it meanwhile
	const bool host_resident = block->size() == segment.SegmentSize();
	if (!segment.gpu_column && host_resident) {
		// not compacted yet (adaptive mode): the slot-width image is the flat block (FixedSizeScanPartial)
		memcpy(target_ptr, block->data() + start * S, scan_count * S);
		return;
	}
	if (!segment.gpu_column) {
		throw InternalException("SuccinctScanPartial: segment has neither a host block nor a device column");
	}
	auto &scan_state = (SuccinctGpuScanState &)*state.scan_state;
	if (segment.column_data && !segment.owns_gpu_column) {
		if (!segment.compacted && host_resident) {
			// a hot segment (uncompacted by the policy): its flat block is in host memory
			memcpy(target_ptr, block->data() + start * S, scan_count * S);
			return;
		}
		memcpy(target_ptr, StreamValues(segment, state, scan_state) + start * S, scan_count * S);
		return;
	}
	if (!segment.compacted && host_resident) {
		memcpy(target_ptr, block->data() + start * S, scan_count * S);
		return;
	}
	if (scan_state.epoch != segment.epoch) {
		scan_state.decoded.resize(segment.count * S);
		GPU_CHECK(segment.db, adacom_decode(segment.gpu_column, segment.gpu_index, segment.gpu_index + 1, 0,
		                                    scan_state.decoded.data(), 0));
		scan_state.epoch = segment.epoch;
	}
	memcpy(target_ptr, scan_state.decoded.data() + start * S, scan_count * S);
}

void SuccinctScan(ColumnSegment &segment, ColumnScanState &state, idx_t scan_count, Vector &result) {
	if (segment.gpu_column && segment.column_data && !segment.owns_gpu_column) {
		// entire vector out of one segment: zero-copy view into the pinned window
		auto start = segment.GetRelativeIndex(state.row_index);
		if (start + scan_count > segment.count) {
			throw InternalException("SuccinctScan: scan runs past the end of the segment");
		}
		auto &scan_state = (SuccinctGpuScanState &)*state.scan_state;
		result.SetVectorType(VectorType::FLAT_VECTOR);
		auto block = segment.HostBlock();
		if (!segment.compacted && block->size() == segment.SegmentSize()) {
			scan_state.hot = block;  // the zero-copy vector outlives this call: the scan state holds the block
			FlatVector::SetData(result, block->data() + start * segment.type_size);  // hot: host-resident flat block
			return;
		}
		FlatVector::SetData(result, (data_ptr_t)StreamValues(segment, state, scan_state) + start * segment.type_size);
		return;
	}
	SuccinctScanPartial(segment, state, scan_count, result, 0);
}

void SuccinctFetchRow(ColumnSegment &segment, ColumnFetchState &, row_t row_id, Vector &result, idx_t result_idx) {
	if (row_id < 0 || (idx_t)row_id >= segment.count) {
		throw InternalException("SuccinctFetchRow: row_id outside the segment");
	}
	idx_t S = segment.type_size;
	data_ptr_t target_ptr = FlatVector::GetData(result) + result_idx * S;
	auto block = segment.HostBlock();
	if (block->size() == segment.SegmentSize() && (!segment.gpu_column || !segment.compacted)) {
		memcpy(target_ptr, block->data() + (idx_t)row_id * S, S);  // host-resident flat block (open or hot segment)
		return;
	}
	uint32_t seg = (uint32_t)segment.gpu_index, row = (uint32_t)row_id;
	GPU_CHECK(segment.db, adacom_fetch_rows(segment.gpu_column, 1, &seg, &row, 0, target_ptr, 0));
}

void EmptySkip(ColumnSegment &, ColumnScanState &, idx_t) {
}

// ---- append -------------------------------------------------------------------------------------------
unique_ptr<CompressionAppendState> SuccinctInitAppend(ColumnSegment &) {
	return unique_ptr<CompressionAppendState>(new CompressionAppendState());
}

template <class T>
idx_t SuccinctAppend(CompressionAppendState &, ColumnSegment &segment, SegmentStatistics &, UnifiedVectorFormat &data,
                     idx_t offset, idx_t count) {
	idx_t max_tuple_count = segment.SegmentSize() / sizeof(T);
	idx_t copy_count = std::min<idx_t>(count, max_tuple_count - segment.count);
	auto sdata = (const T *)data.data;
	auto target = (T *)segment.host_block->data();
	if ((!data.sel || !data.sel->sel) && data.validity.AllValid()) {
		// flat, all-valid vector (the common case of a bulk load): one block copy
		memcpy(target + segment.count, sdata + offset, copy_count * sizeof(T));
		segment.count += copy_count;
		return copy_count;
	}
	SelectionVector identity;
	const SelectionVector *sel = data.sel ? data.sel : &identity;
	for (idx_t i = 0; i < copy_count; i++) {
		auto source_idx = sel->get_index(offset + i);
		auto target_idx = segment.count + i;
		if (data.validity.RowIsValid(source_idx)) {
			target[target_idx] = sdata[source_idx];
		} else {
			// NullValue<T>() in the gap (succinct.cpp:288-292); the analyze kernel skips the row
			target[target_idx] = std::numeric_limits<T>::min();
			segment.staging_validity[target_idx >> 6] &= ~(1ull << (target_idx & 63));
			segment.staging_has_nulls = true;
		}
	}
	segment.count += copy_count;
	return copy_count;
}

template <class T>
idx_t SuccinctFinalizeAppend(ColumnSegment &segment, SegmentStatistics &) {
	return segment.count * sizeof(T);
}

template <class T>
CompressionFunction SuccinctGetFunction(PhysicalType type) {
	CompressionFunction f;
	f.type = CompressionType::COMPRESSION_SUCCINCT;
	f.data_type = type;
	f.init_analyze = SuccinctInitAnalyze;
	f.analyze = SuccinctAnalyze;
	f.final_analyze = SuccinctFinalAnalyze;
	f.init_compression = SuccinctInitCompression;
	f.compress = SuccinctCompress;
	f.compress_finalize = SuccinctFinalizeCompress;
	f.init_scan = SuccinctInitScan;
	f.scan_vector = SuccinctScan;
	f.scan_partial = SuccinctScanPartial;
	f.fetch_row = SuccinctFetchRow;
	f.skip = EmptySkip;
	f.init_segment = nullptr;
	f.init_append = SuccinctInitAppend;
	f.append = SuccinctAppend<T>;
	f.finalize_append = SuccinctFinalizeAppend<T>;
	f.revert_append = nullptr;
	return f;
}

}  // namespace

void BindAnalyzeDatabase(DatabaseInstance *db) {
	g_analyze_db = db;
}

ColumnDataCheckpointer *NewCheckpointer(DatabaseInstance &db, PhysicalType type, idx_t row_start) {
	return new ColumnDataCheckpointer(db, type, row_start);
}
void DeleteCheckpointer(ColumnDataCheckpointer *cp) {
	delete cp;
}
std::vector<unique_ptr<ColumnSegment>> &CheckpointerSegments(ColumnDataCheckpointer &cp) {
	return cp.flushed;
}

CompressionFunction SuccinctGpuFun::GetFunction(PhysicalType data_type) {
	switch (data_type) {
	case PhysicalType::INT8:
		return SuccinctGetFunction<int8_t>(data_type);
	case PhysicalType::UINT8:
		return SuccinctGetFunction<uint8_t>(data_type);
	case PhysicalType::INT16:
		return SuccinctGetFunction<int16_t>(data_type);
	case PhysicalType::UINT16:
		return SuccinctGetFunction<uint16_t>(data_type);
	case PhysicalType::INT32:
		return SuccinctGetFunction<int32_t>(data_type);
	case PhysicalType::UINT32:
		return SuccinctGetFunction<uint32_t>(data_type);
	case PhysicalType::INT64:
		return SuccinctGetFunction<int64_t>(data_type);
	case PhysicalType::UINT64:
		return SuccinctGetFunction<uint64_t>(data_type);
	default:
		throw InternalException("Unsupported type for FixedSizeSuccinct::GetFunction");
	}
}

bool SuccinctGpuFun::TypeIsSupported(PhysicalType type) {
	switch (type) {
	case PhysicalType::INT8:
	case PhysicalType::INT16:
	case PhysicalType::INT32:
	case PhysicalType::INT64:
	case PhysicalType::UINT8:
	case PhysicalType::UINT16:
	case PhysicalType::UINT32:
	case PhysicalType::UINT64:
		return true;
	default:
		return false;
	}
}

}  // namespace adacom_host
