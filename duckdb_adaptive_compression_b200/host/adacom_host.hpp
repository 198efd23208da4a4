// adacom_host.hpp -- C++ host side above the C ABI: the reference's own plugin / segment / policy surface
// for the succinct path, re-hosted on the GPU codec.
//
// Everything here mirrors a reference interface by name, argument meaning and error behaviour so that the
// storage layer's call sites read the same (paths relative to the reference repository root):
//
//   CompressionFunction + typedefs .. src/include/duckdb/function/compression_function.hpp:29-177
//   SuccinctFun ..................... src/include/duckdb/function/compression/compression.hpp:56-59,
//                                     src/storage/compression/succinct.cpp:335-382
//   ColumnSegment ................... src/include/duckdb/storage/table/column_segment.hpp:60,129-214,
//                                     src/storage/table/column_segment.cpp:45-110,133-346
//   ColumnSegmentCatalog ............ src/include/duckdb/catalog/catalog_entry/column_segment_catalog.hpp:14-54,
//                                     src/catalog/catalog_entry/column_segment_catalog.cpp:13-135
//   DBConfig succinct flags ......... src/include/duckdb/main/config.hpp:189-197
//   BufferManager data-size counter . src/include/duckdb/storage/buffer_manager.hpp:71-82
//
// The stand-ins for Vector / UnifiedVectorFormat / ColumnScanState carry only what the succinct path reads.
// All arithmetic happens in libadacom_gpu.so; this layer owns no codec loop (and no CPU fallback).
#pragma once

#include "../../include/adacom_gpu.h"

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace adacom_host {

using idx_t = uint64_t;
using row_t = int64_t;
using data_ptr_t = uint8_t *;
using block_id_t = int64_t;
using std::unique_ptr;

constexpr idx_t STANDARD_VECTOR_SIZE = 2048;   // src/include/duckdb/common/vector_size.hpp:17
constexpr idx_t BLOCK_SIZE = 262144 - 8;       // Storage::BLOCK_SIZE, src/include/duckdb/common/constants.hpp:101-106

struct InternalException : public std::runtime_error {
	explicit InternalException(const std::string &msg) : std::runtime_error(msg) {
	}
};

// src/include/duckdb/common/types.hpp (values of the integer physical types)
enum class PhysicalType : uint8_t {
	BOOL = 1, UINT8 = 2, INT8 = 3, UINT16 = 4, INT16 = 5, UINT32 = 6, INT32 = 7, UINT64 = 8, INT64 = 9,
	FLOAT = 11, DOUBLE = 12, VARCHAR = 200, INT128 = 204
};
idx_t GetTypeIdSize(PhysicalType type);
bool TypeIsInteger(PhysicalType type);  // src/common/types.cpp:318-320 (INT128 included, quirk Q6)

// src/include/duckdb/common/enums/compression_type.hpp:16-29
enum class CompressionType : uint8_t { COMPRESSION_AUTO = 0, COMPRESSION_UNCOMPRESSED = 1, COMPRESSION_SUCCINCT = 10 };

enum class VectorType : uint8_t { FLAT_VECTOR = 0 };

// ---- minimal Vector / UnifiedVectorFormat ---------------------------------------------------------------
struct Vector {
	explicit Vector(PhysicalType type) : type(type), buffer(STANDARD_VECTOR_SIZE * GetTypeIdSize(type)) {
	}
	Vector(PhysicalType type, idx_t capacity) : type(type), buffer(capacity * GetTypeIdSize(type)) {
	}
	void SetVectorType(VectorType t) {
		vector_type = t;
	}
	PhysicalType type;
	VectorType vector_type = VectorType::FLAT_VECTOR;
	std::vector<uint8_t> buffer;
	data_ptr_t external = nullptr;  // zero-copy view set by FlatVector::SetData (cleared by the next pull)
};
struct FlatVector {
	static data_ptr_t GetData(Vector &v) {
		return v.external ? v.external : v.buffer.data();
	}
	//! src/include/duckdb/common/types/vector.hpp: FlatVector::SetData -- the vector points at caller-owned memory
	static void SetData(Vector &v, data_ptr_t data) {
		v.external = data;
	}
};
struct SelectionVector {
	const uint32_t *sel = nullptr;  // nullptr = identity
	idx_t get_index(idx_t i) const {
		return sel ? sel[i] : i;
	}
};
struct ValidityMask {
	const uint64_t *mask = nullptr;  // nullptr = all valid
	bool AllValid() const {
		return mask == nullptr;
	}
	bool RowIsValid(idx_t row) const {
		return !mask || ((mask[row >> 6] >> (row & 63)) & 1);
	}
};
struct UnifiedVectorFormat {
	const SelectionVector *sel = nullptr;
	data_ptr_t data = nullptr;
	ValidityMask validity;
};

// ---- state objects of the plugin ABI (compression_function.hpp:29-51) -----------------------------------
struct AnalyzeState { virtual ~AnalyzeState() = default; };
struct CompressionState { virtual ~CompressionState() = default; };
struct CompressedSegmentState { virtual ~CompressedSegmentState() = default; };
struct SegmentScanState { virtual ~SegmentScanState() = default; };
struct CompressionAppendState { virtual ~CompressionAppendState() = default; };

class ColumnSegment;
class ColumnData;
class ColumnDataCheckpointer;
struct SegmentStatistics {
	bool has_value = false;
	int64_t min = 0, max = 0;      // signed view
	uint64_t umin = 0, umax = 0;   // unsigned view
};
struct ColumnScanState {  // src/include/duckdb/storage/table/scan_state.hpp:43-74
	ColumnSegment *current = nullptr;
	idx_t row_index = 0;
	idx_t internal_index = 0;
	unique_ptr<SegmentScanState> scan_state;
	bool initialized = false;
	//! set by InitializeScanWithOffset: the first pull of such a scan is treated as a point read (hot-vector cache)
	bool probe = false;
	//! the previous pull was served as a point read (a scan that pulls again tells the column it was not one)
	bool was_point_read = false;
	idx_t version = 0;
	//! scan states of the segments a vector has just left: kept alive until the next ScanVector (zero-copy vectors)
	std::vector<unique_ptr<SegmentScanState>> previous_states;
};
struct ColumnFetchState {};
struct ColumnAppendState {
	unique_ptr<CompressionAppendState> append_state;
};

// ---- plugin vtable: same slots, same signatures (compression_function.hpp:65-177) -----------------------
typedef unique_ptr<AnalyzeState> (*compression_init_analyze_t)(ColumnData &col_data, PhysicalType type);
typedef bool (*compression_analyze_t)(AnalyzeState &state, Vector &input, idx_t count);
typedef idx_t (*compression_final_analyze_t)(AnalyzeState &state);
typedef unique_ptr<CompressionState> (*compression_init_compression_t)(ColumnDataCheckpointer &checkpointer,
                                                                       unique_ptr<AnalyzeState> state);
typedef void (*compression_compress_data_t)(CompressionState &state, Vector &scan_vector, idx_t count);
typedef void (*compression_compress_finalize_t)(CompressionState &state);
typedef unique_ptr<SegmentScanState> (*compression_init_segment_scan_t)(ColumnSegment &segment);
typedef void (*compression_scan_vector_t)(ColumnSegment &segment, ColumnScanState &state, idx_t scan_count,
                                          Vector &result);
typedef void (*compression_scan_partial_t)(ColumnSegment &segment, ColumnScanState &state, idx_t scan_count,
                                           Vector &result, idx_t result_offset);
typedef void (*compression_fetch_row_t)(ColumnSegment &segment, ColumnFetchState &state, row_t row_id, Vector &result,
                                        idx_t result_idx);
typedef void (*compression_skip_t)(ColumnSegment &segment, ColumnScanState &state, idx_t skip_count);
typedef unique_ptr<CompressedSegmentState> (*compression_init_segment_t)(ColumnSegment &segment, block_id_t block_id);
typedef unique_ptr<CompressionAppendState> (*compression_init_append_t)(ColumnSegment &segment);
typedef idx_t (*compression_append_t)(CompressionAppendState &append_state, ColumnSegment &segment,
                                      SegmentStatistics &stats, UnifiedVectorFormat &data, idx_t offset, idx_t count);
typedef idx_t (*compression_finalize_append_t)(ColumnSegment &segment, SegmentStatistics &stats);
typedef void (*compression_revert_append_t)(ColumnSegment &segment, idx_t start_row);

class CompressionFunction {
public:
	CompressionType type;
	PhysicalType data_type;
	compression_init_analyze_t init_analyze;
	compression_analyze_t analyze;
	compression_final_analyze_t final_analyze;
	compression_init_compression_t init_compression;
	compression_compress_data_t compress;
	compression_compress_finalize_t compress_finalize;
	compression_init_segment_scan_t init_scan;
	compression_scan_vector_t scan_vector;
	compression_scan_partial_t scan_partial;
	compression_fetch_row_t fetch_row;
	compression_skip_t skip;
	compression_init_segment_t init_segment;
	compression_init_append_t init_append;
	compression_append_t append;
	compression_finalize_append_t finalize_append;
	compression_revert_append_t revert_append;
};

// Drop-in for SuccinctFun (src/storage/compression/succinct.cpp:335-382): same two static entry points.
struct SuccinctGpuFun {
	static CompressionFunction GetFunction(PhysicalType data_type);  // throws InternalException if unsupported
	static bool TypeIsSupported(PhysicalType type);
};

// ---- database-level singletons the segment code reaches for ---------------------------------------------
struct DBConfig {  // src/include/duckdb/main/config.hpp:189-197 (same names, same defaults)
	bool succinct_enabled = true;
	bool succinct_extract_prefix_enabled = true;
	bool succinct_padded_to_next_byte_enabled = false;
	bool adaptive_succinct_compression_enabled = false;
	int gpu_device = 0;                      // B200 ordinal this database instance is pinned to (first device)
	int gpu_devices = 1;                     // number of GPUs the row groups are dealt over (gpu_device .. +n-1)
	double compression_rate = 0.90;          // column_segment_catalog.cpp:65
	double background_interval_s = 10.0;     // column_segment_catalog.cpp:70
	CompressionFunction *GetCompressionFunction(CompressionType type, PhysicalType data_type);  // registry, locked

private:
	std::mutex lock;
	std::unordered_map<uint32_t, std::unique_ptr<CompressionFunction>> functions;
};

class BufferManager {  // the two counters the succinct code adjusts (buffer_manager.hpp:71-82)
public:
	void AddToDataSize(int64_t delta) {
		data_size += delta;
	}
	idx_t GetDataSize() const {
		return (idx_t)data_size.load();
	}

private:
	std::atomic<int64_t> data_size {0};
};

class CompactableSegment;
// The segments of one column live in one device column, so the policy's per-epoch subsets become ONE launch set each
// (adacom_encode_list / adacom_uncompact_list) instead of one launch set per segment.  ColumnData implements it.
class SegmentBatchOwner {
public:
	virtual ~SegmentBatchOwner() = default;
	virtual void CompactMany(const std::vector<CompactableSegment *> &segments) = 0;
	virtual void UncompactMany(const std::vector<CompactableSegment *> &segments) = 0;
};

// Anything the access-frequency policy can compact: ColumnSegment implements it, CPU tests use a fake.
class CompactableSegment {
public:
	virtual ~CompactableSegment() = default;
	virtual void Compact() = 0;
	virtual void Uncompact() = 0;
	virtual idx_t GetDataSize() const = 0;
	virtual bool IsDataSegment() const = 0;
	//! non-null: Compact()/Uncompact() decisions of an epoch for this segment are handed to the owner in one batch
	virtual SegmentBatchOwner *BatchOwner() const {
		return nullptr;
	}
	//! GPU this segment's payload lives on (device-sharded catalog: per-device accounting); -1 = host only
	virtual int Device() const {
		return -1;
	}
	//! reads since the last epoch.  The counter lives in the segment so that AddReadAccess on the scan path (one call
	//! per 2048-value pull, from every scan thread) is one relaxed atomic increment instead of a catalog-wide lock
	std::atomic<idx_t> catalog_reads {0};
	std::atomic<bool> catalog_registered {false};
};

struct AccessStatistics {
	idx_t num_reads;
};

// Same public surface as the reference's ColumnSegmentCatalog; the statistics map is guarded by a mutex and
// Compact/Uncompact run under it (the reference's races, quirk Q5, are not reproduced).
class ColumnSegmentCatalog {
public:
	explicit ColumnSegmentCatalog(double compression_rate = 0.90, double interval_s = 10.0);
	~ColumnSegmentCatalog();
	void AddColumnSegment(CompactableSegment *segment);
	void AddReadAccess(CompactableSegment *segment);
	void RemoveColumnSegment(CompactableSegment *segment);
	void EnableBackgroundThreadCompaction();
	bool BackgroundCompactionEnabled() const {
		return background_compaction_enabled;
	}
	void CompactAllSegments();
	size_t GetTotalDataSize();
	//! the same sum split by the GPU the segments live on (key -1: segments without a device payload)
	std::unordered_map<int, size_t> GetDataSizePerDevice();
	//! One iteration of CompressLowestKSegments' loop body (column_segment_catalog.cpp:71-112): sort by
	//! reads, Compact() the coldest `compression_rate` share, Uncompact() the rest, zero the counters.
	//! Returns {compacted, uncompacted}.
	std::pair<size_t, size_t> RunCompactionEpoch();
	idx_t NumReads(CompactableSegment *segment);
	void StopBackgroundThread();

private:
	void CompressLowestKSegments();
	std::mutex lock;
	std::unordered_map<CompactableSegment *, AccessStatistics> statistics;
	std::atomic<idx_t> event_counter {0};
	bool background_compaction_enabled = false;
	double compression_rate;
	double interval_s;
	std::thread worker;
	std::condition_variable cv;
	bool stop = false;
};

class DatabaseInstance {
public:
	explicit DatabaseInstance(const DBConfig *cfg = nullptr);
	~DatabaseInstance();
	DBConfig config;
	BufferManager buffer_manager;
	ColumnSegmentCatalog column_segment_catalog;
	adacom_ctx *gpu = nullptr;  // created lazily by Gpu(); throws InternalException when no B200 is usable
	adacom_ctx *Gpu();
	//! context of GPU `device` (config.gpu_device when < 0); one per device, created on first use.  Columns are dealt
	//! over config.gpu_devices GPUs by ColumnData (row group g -> device g mod G)
	adacom_ctx *Gpu(int device);
	//! owner of row group g: gpu_device + g mod gpu_devices (segments never span row groups, so nothing else moves)
	int DeviceForRowGroup(idx_t row_group) const {
		return config.gpu_device + (int)(row_group % (idx_t)(config.gpu_devices > 0 ? config.gpu_devices : 1));
	}
	[[noreturn]] void Throw(const char *what);

private:
	std::mutex gpu_lock;
	std::unordered_map<int, adacom_ctx *> gpus;
};

// ---- ColumnSegment ----------------------------------------------------------------------------------------
enum class ColumnSegmentType : uint8_t { TRANSIENT, PERSISTENT };

class ColumnSegment : public CompactableSegment {
public:
	//! ColumnSegment::CreateTransientSegment (column_segment.cpp:45-82): SUCCINCT for integer types when
	//! succinct_enabled && !adaptive, else UNCOMPRESSED (adaptive mode compacts in the background).
	static unique_ptr<ColumnSegment> CreateTransientSegment(DatabaseInstance &db, PhysicalType type, idx_t start,
	                                                        idx_t segment_size = BLOCK_SIZE);
	~ColumnSegment() override;

	void InitializeScan(ColumnScanState &state);
	void Scan(ColumnScanState &state, idx_t scan_count, Vector &result, idx_t result_offset, bool entire_vector);
	void FetchRow(ColumnFetchState &state, row_t row_id, Vector &result, idx_t result_idx);
	void Skip(ColumnScanState &state);
	void InitializeAppend(ColumnAppendState &state);
	idx_t Append(ColumnAppendState &state, UnifiedVectorFormat &data, idx_t offset, idx_t count);
	idx_t FinalizeAppend(ColumnAppendState &state);

	void Compact() override;
	void Uncompact() override;
	idx_t GetDataSize() const override;
	idx_t SuccinctSize() const;
	bool IsDataSegment() const override {
		return is_data_segment;
	}
	idx_t SegmentSize() const {
		return segment_size;
	}
	idx_t GetRelativeIndex(idx_t row_index) const {
		return row_index - start;
	}
	uint64_t GetMinFactor();
	uint64_t GetMax();
	uint8_t Width();
	bool IsBitCompressed() const {
		return compacted;
	}
	SegmentBatchOwner *BatchOwner() const override;
	int Device() const override;

	DatabaseInstance &db;
	PhysicalType type;
	idx_t type_size;
	ColumnSegmentType segment_type = ColumnSegmentType::TRANSIENT;
	idx_t start;
	//! values in the segment; atomic like SegmentBase::count of the reference (src/include/duckdb/storage/table/
	//! segment_base.hpp): a reader that sees the new count also sees the values an append wrote before it
	std::atomic<idx_t> count {0};
	idx_t num_elements = 0;
	CompressionFunction *function;
	SegmentStatistics stats;
	idx_t segment_size;
	bool compacted = false;
	bool succinct_possible = false;
	bool background_compaction_enabled = false;
	bool is_data_segment = false;
	bool force_reinitializing_scan_state = false;
	std::mutex bit_compression_lock;

	// device side: segment `gpu_index` of `gpu_column`.  A lone segment owns a one-segment column of the C ABI; a
	// segment of a ColumnData is a member of the column's shared device column (one launch set per batch, one scan
	// stream per scan).  Values are staged on the host until the first Compact().
	adacom_column *gpu_column = nullptr;
	uint64_t gpu_index = 0;
	bool owns_gpu_column = true;
	ColumnData *column_data = nullptr;
	int device = -1;                      // GPU ordinal of gpu_column (-1: default device of the database)
	//! slot-width host image of the values (== the raw arena): the block of an open segment, of a segment that was never
	//! uploaded (adaptive mode) and of a HOT segment the policy uncompacted; empty once the values only live on the device.
	//! Shared ownership: a scan that handed out a zero-copy vector keeps the block alive while the policy drops it
	//! (the pointer is swapped under bit_compression_lock; the reference's own version of this is a data race, Q5)
	std::shared_ptr<std::vector<uint8_t>> host_block = std::make_shared<std::vector<uint8_t>>();
	std::shared_ptr<std::vector<uint8_t>> HostBlock() {
		std::lock_guard<std::mutex> guard(bit_compression_lock);
		return host_block;
	}
	void DropHostBlock() {
		auto empty = std::make_shared<std::vector<uint8_t>>();
		std::lock_guard<std::mutex> guard(bit_compression_lock);
		host_block.swap(empty);
	}
	std::vector<uint64_t> staging_validity;
	bool staging_has_nulls = false;
	idx_t uploaded = 0;                   // values already copied to the device
	uint64_t epoch = 0;                   // bumped by Compact/Uncompact: scan states re-initialise
	idx_t compacted_size = 0;             // size_in_bytes of the compacted image (GetDataSize without a device call)

private:
	ColumnSegment(DatabaseInstance &db, PhysicalType type, idx_t start, idx_t segment_size, CompressionFunction *function,
	              bool succinct_possible, bool background, bool is_data_segment);
	void Scan(ColumnScanState &state, idx_t scan_count, Vector &result);
	void ScanPartial(ColumnScanState &state, idx_t scan_count, Vector &result, idx_t result_offset);
	void EnsureDeviceColumn();
	void EnsureStaging();
	void DetachFromBatch();
	void Upload();
	int CompactFlags() const;
	void FinishCompact(idx_t size_before, const adacom_segment_meta &m);
	void FinishUncompact(idx_t compressed_size);
	adacom_segment_meta Meta() const;
	friend struct SuccinctGpuFun;
	friend class ColumnData;
};

// ---- table filters pushed to one column (src/include/duckdb/planner/table_filter.hpp:21-27, planner/filter/*.hpp) --
// Same class names, members and enum values as the reference; the constant of a comparison is an integer (the path
// covers the integer physical types).
enum class TableFilterType : uint8_t {
	CONSTANT_COMPARISON = 0, IS_NULL = 1, IS_NOT_NULL = 2, CONJUNCTION_OR = 3, CONJUNCTION_AND = 4
};
enum class ExpressionType : uint8_t {  // src/include/duckdb/common/enums/expression_type.hpp:34-46
	COMPARE_EQUAL = 25, COMPARE_NOTEQUAL = 26, COMPARE_LESSTHAN = 27, COMPARE_GREATERTHAN = 28,
	COMPARE_LESSTHANOREQUALTO = 29, COMPARE_GREATERTHANOREQUALTO = 30
};
class TableFilter {
public:
	explicit TableFilter(TableFilterType filter_type_p) : filter_type(filter_type_p) {
	}
	virtual ~TableFilter() = default;
	TableFilterType filter_type;
};
class ConstantFilter : public TableFilter {  // planner/filter/constant_filter.hpp:17-24
public:
	ConstantFilter(ExpressionType comparison_type_p, int64_t constant_p)
	    : TableFilter(TableFilterType::CONSTANT_COMPARISON), comparison_type(comparison_type_p), constant(constant_p) {
	}
	ExpressionType comparison_type;
	int64_t constant;
};
class IsNullFilter : public TableFilter {
public:
	IsNullFilter() : TableFilter(TableFilterType::IS_NULL) {
	}
};
class IsNotNullFilter : public TableFilter {
public:
	IsNotNullFilter() : TableFilter(TableFilterType::IS_NOT_NULL) {
	}
};
class ConjunctionFilter : public TableFilter {  // planner/filter/conjunction_filter.hpp:15-24
public:
	explicit ConjunctionFilter(TableFilterType filter_type_p) : TableFilter(filter_type_p) {
	}
	std::vector<unique_ptr<TableFilter>> child_filters;
};
class ConjunctionOrFilter : public ConjunctionFilter {
public:
	ConjunctionOrFilter() : ConjunctionFilter(TableFilterType::CONJUNCTION_OR) {
	}
};
class ConjunctionAndFilter : public ConjunctionFilter {
public:
	ConjunctionAndFilter() : ConjunctionFilter(TableFilterType::CONJUNCTION_AND) {
	}
};
//! One adacom_scan_select of a pushed filter: the children of one conjunction level (ColumnData::BuildFilterStages)
struct FilterStage {
	std::vector<adacom_pred> preds;
	adacom_conj conj = ADACOM_CONJ_AND;
};
//! What DuckDB computes above a filtered scan of one column: approved_tuple_count of FilterSelection
//! (column_segment.cpp:685-844) and COUNT(col) / SUM(col) over the approved rows (NULL rows do not enter them)
struct FilterAggregateResult {
	idx_t approved_tuple_count = 0;
	idx_t valid_count = 0;
	uint64_t sum_lo = 0;  // 128-bit two's complement (HUGEINT)
	uint64_t sum_hi = 0;
};

// ---- ColumnData: the owner of a column's segments (src/storage/table/column_data.cpp) ---------------------
// Same entry points as the reference for the succinct path: AppendData (:249-270), InitializeScan[WithOffset]
// (:73-89), ScanVector (:92-139), FetchRow (:304-315).  What is new is below the segments: all sealed segments of the
// column share ONE device column, so
//   * compaction of many segments is one analyze + one encode launch set (CompactPending / CompactMany),
//   * a scan is served by one adacom_scan_stream that runs ahead of the 2048-value pulls across segment borders,
//   * the column sits on one GPU (`device`), chosen by the caller: row group g -> device g mod G.
class ColumnData : public SegmentBatchOwner {
public:
	ColumnData(DatabaseInstance &db, PhysicalType type, idx_t start_row = 0, int device = -1);
	~ColumnData() override;

	void InitializeAppend(ColumnAppendState &state);
	void AppendData(ColumnAppendState &state, UnifiedVectorFormat &vdata, idx_t count);
	void InitializeScan(ColumnScanState &state);
	void InitializeScanWithOffset(ColumnScanState &state, idx_t row_idx);
	idx_t ScanVector(ColumnScanState &state, Vector &result, idx_t remaining);
	void FetchRow(ColumnFetchState &state, row_t row_id, Vector &result, idx_t result_idx);
	//! batched form of FetchRow for index joins: one adacom_fetch_rows launch for all row ids
	void FetchRows(const row_t *row_ids, idx_t count, Vector &result);
	idx_t GetMaxEntry() const;
	ColumnSegment *GetSegment(idx_t row_idx) const;
	//! The table filter of this column and the aggregate above it in one pass over the packed payload (the work of
	//! ColumnData::Select -> ColumnSegment::FilterSelection on every vector, column_segment.cpp:685-844, plus the
	//! SUM / COUNT DuckDB runs on the approved rows): one adacom_scan_select per run of segments that share a
	//! device column.  Leaf filters, conjunctions of leaves (same-kind nesting is flattened) and an AND whose children
	//! include OR groups of leaves (IN lists): one call per group, chained through device selection bitmaps like the
	//! recursion of FilterSelection.  An AND group under an OR (a union of selections) throws.
	FilterAggregateResult FilterAggregate(const TableFilter &filter);
	//! the chain of calls FilterAggregate makes for `filter` (host logic only: no GPU is touched)
	static std::vector<FilterStage> BuildFilterStages(const TableFilter &filter);

	//! move every sealed (full or finalized) segment that is still host-staged into the shared device column and
	//! compact them with one analyze + one encode; called by the append path every `pending_limit` segments, by
	//! the first scan that meets a pending segment, and by ColumnSegmentCatalog::CompactAllSegments
	void CompactPending(bool include_last = true);
	//! Append filled a segment: the batch is flushed once `pending_limit` sealed segments wait on the host
	void SegmentSealed();
	void CompactMany(const std::vector<CompactableSegment *> &segments) override;
	void UncompactMany(const std::vector<CompactableSegment *> &segments) override;

	//! Hot vectors of point reads.  A scan that starts WithOffset and pulls one vector (how a point query reads a
	//! column) or a FetchRow costs a launch and a PCIe round trip (13-37 us) where the reference reads RAM.  Skewed
	//! point reads come back to the same vectors, so a 2048-row vector that is touched a SECOND time is decoded once
	//! into host memory and served from there (direct-mapped, `vector_cache_bytes` per column; rows are append-only,
	//! a partially filled tail vector is refreshed when a read goes past what it held).  Uniform random reads never
	//! touch a vector twice and pay nothing but the look-up.  0 disables.
	idx_t vector_cache_bytes = 32ull << 20;
	struct VectorCacheStats {
		idx_t hits = 0, fills = 0, misses = 0;
		idx_t lean_reads = 0;  // first touches served by one adacom_scan_partial
	};
	VectorCacheStats GetVectorCacheStats();

	DatabaseInstance &db;
	PhysicalType type;
	idx_t start;
	int device;
	std::atomic<idx_t> version {0};
	idx_t pending_limit = 64;  // sealed segments staged on the host before a batch is uploaded (16 MiB)
	bool hot_segments_on_host = true;  // Uncompact() brings the flat block of a (hot) segment back to host memory
	std::vector<unique_ptr<ColumnSegment>> segments;

	//! shared device columns of this ColumnData: the members of batch k are segments whose gpu_column == col
	struct Batch {
		adacom_column *col = nullptr;
		idx_t n = 0;
		uint64_t *validity = nullptr;  // pinned: validity bits of the batch's rows when a member has NULLs
	};
	std::vector<Batch> batches;
	std::mutex lock;
	//! guards the STRUCTURE of `segments` (the reference's SegmentTree::node_lock): an append that opens a new segment
	//! may move the vector's array under a reader.  Leaf lock, held only for the look-up or the push_back (never while
	//! `lock` is waited for; `lock` may be held when it is taken).  Segments are never removed, so the pointers a
	//! look-up returns stay valid.
	mutable std::mutex tree_lock;
	//! the segments as they are now (pointers stay valid; the vector itself may grow behind the snapshot)
	std::vector<ColumnSegment *> SegmentSnapshot() const;

private:
	struct CachedVector {
		idx_t vec = ~(idx_t)0;  // absolute row / 2048
		idx_t rows = 0;        // rows the entry holds (a tail vector may hold fewer than 2048)
		std::vector<uint8_t> data;
	};
	std::vector<CachedVector> vector_cache;
	std::vector<idx_t> vector_seen;  // vectors touched once (direct-mapped filter)
	std::mutex vector_cache_lock;
	VectorCacheStats vector_cache_stats;
	//! scans that began WithOffset and were served as a point read / of those, the ones that pulled again: a parallel
	//! table scan starts every row group WithOffset and always pulls again, so the one-launch first touch (which a
	//! continuing scan pays on top of its stream) is only used while most such scans stop after one vector
	std::atomic<uint32_t> point_reads_started {0}, point_reads_continued {0};
	//! copy `count` values from row `offset` of vector `vec` if the cache holds them; a second touch fills the entry
	bool ReadHotVector(idx_t vec, idx_t offset, idx_t count, data_ptr_t dst);
	//! rows [row, row + count) by ONE adacom_scan_partial when they lie inside one packed member of a shared column
	bool ReadPackedRows(ColumnSegment *seg, idx_t row, idx_t count, data_ptr_t dst);

public:

private:
	void AppendTransientSegment(idx_t start_row);
	//! one shared device column for the given host-staged segments: upload, one analyze, one encode
	void CompactHostSegments(const std::vector<ColumnSegment *> &candidates);
};

adacom_ptype ToAdacomType(PhysicalType type);

// Hooks for the checkpoint-style slots: ColumnData / ColumnDataCheckpointer are opaque to the plugin ABI, so
// the driver (DuckDB's ColumnDataCheckpointer::WriteToDisk, or a test) binds the database instance and reads
// the flushed segments through these.
void BindAnalyzeDatabase(DatabaseInstance *db);
ColumnDataCheckpointer *NewCheckpointer(DatabaseInstance &db, PhysicalType type, idx_t row_start);
void DeleteCheckpointer(ColumnDataCheckpointer *cp);
std::vector<unique_ptr<ColumnSegment>> &CheckpointerSegments(ColumnDataCheckpointer &cp);

}  // namespace adacom_host
