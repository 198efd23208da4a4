// column_data.cpp -- ColumnData re-hosted on the GPU codec: the owner of a column's segments.
//
// Entry points and control flow follow the reference (paths relative to its repository root):
//   ColumnData::InitializeScan / InitializeScanWithOffset .. src/storage/table/column_data.cpp:73-89
//   ColumnData::ScanVector .................................. column_data.cpp:92-139
//   ColumnData::InitializeAppend / AppendData ............... column_data.cpp:222-270
//   ColumnData::FetchRow ..................................... column_data.cpp:304-315
// What the reference does per segment on the CPU (Compact() of each full segment, a deep copy + decode per 2048-value
// pull, one FetchRow per row id) is batched here below the same interface: the sealed segments of the column share one
// device column, so compaction is one analyze + one encode launch set per batch, a scan is one decode launch per
// 32 segments running ahead of the pulls in a pinned ring, and the policy's per-epoch subsets are one launch set each.
#include "adacom_host.hpp"

#include <algorithm>
#include <map>
#include <vector>

namespace adacom_host {

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

ColumnData::ColumnData(DatabaseInstance &db_p, PhysicalType type_p, idx_t start_row, int device_p)
    : db(db_p), type(type_p), start(start_row), device(device_p < 0 ? db_p.config.gpu_device : device_p) {
}

ColumnData::~ColumnData() {
	segments.clear();  // members only drop their reference; the shared device columns go last
	for (auto &b : batches) {
		if (b.col) {
			adacom_column_destroy(b.col);
		}
		if (b.validity) {
			adacom_host_free(db.Gpu(device), b.validity);
		}
	}
}

void ColumnData::AppendTransientSegment(idx_t start_row) {
	auto segment = ColumnSegment::CreateTransientSegment(db, type, start_row);
	segment->column_data = this;
	segment->device = device;
	{
		std::lock_guard<std::mutex> tree_guard(tree_lock);
		segments.push_back(std::move(segment));
	}
	version++;
}

void ColumnData::InitializeAppend(ColumnAppendState &state) {
	std::lock_guard<std::mutex> guard(lock);
	if (segments.empty()) {
		AppendTransientSegment(start);
	}
	segments.back()->InitializeAppend(state);
}

void ColumnData::AppendData(ColumnAppendState &state, UnifiedVectorFormat &vdata, idx_t count) {
	idx_t offset = 0;
	while (true) {
		ColumnSegment *current = segments.back().get();
		idx_t copied_elements = current->Append(state, vdata, offset, count);
		if (copied_elements == count) {
			break;
		}
		{
			std::lock_guard<std::mutex> guard(lock);
			AppendTransientSegment(current->start + current->count);
			segments.back()->InitializeAppend(state);
		}
		offset += copied_elements;
		count -= copied_elements;
	}
}

std::vector<ColumnSegment *> ColumnData::SegmentSnapshot() const {
	std::lock_guard<std::mutex> tree_guard(tree_lock);
	std::vector<ColumnSegment *> snapshot(segments.size());
	for (idx_t i = 0; i < segments.size(); i++) {
		snapshot[i] = segments[i].get();
	}
	return snapshot;
}

idx_t ColumnData::GetMaxEntry() const {
	std::lock_guard<std::mutex> tree_guard(tree_lock);
	return segments.empty() ? 0 : segments.back()->start + segments.back()->count - segments.front()->start;
}

ColumnSegment *ColumnData::GetSegment(idx_t row_idx) const {
	// SegmentTree::GetSegment: binary search over the segment starts
	std::lock_guard<std::mutex> tree_guard(tree_lock);
	auto it = std::upper_bound(segments.begin(), segments.end(), row_idx,
	                           [](idx_t row, const unique_ptr<ColumnSegment> &s) { return row < s->start; });
	if (it == segments.begin()) {
		throw InternalException("ColumnData::GetSegment: row before the first segment");
	}
	--it;
	if (row_idx >= (*it)->start + (*it)->count && it + 1 != segments.end()) {
		throw InternalException("ColumnData::GetSegment: row not covered by a segment");
	}
	return it->get();
}

static ColumnSegment *NextSegment(const ColumnData &col, ColumnSegment *current) {
	std::lock_guard<std::mutex> tree_guard(col.tree_lock);
	auto it = std::lower_bound(col.segments.begin(), col.segments.end(), current->start,
	                           [](const unique_ptr<ColumnSegment> &s, idx_t row) { return s->start < row; });
	while (it != col.segments.end() && it->get() != current) {
		++it;
	}
	if (it == col.segments.end() || it + 1 == col.segments.end()) {
		return nullptr;
	}
	return (it + 1)->get();
}

void ColumnData::InitializeScan(ColumnScanState &state) {
	{
		std::lock_guard<std::mutex> tree_guard(tree_lock);
		state.current = segments.empty() ? nullptr : segments.front().get();
	}
	state.row_index = state.current ? state.current->start : 0;
	state.internal_index = state.row_index;
	state.initialized = false;
	state.probe = false;
	state.was_point_read = false;
	state.version = version;
	state.scan_state.reset();
}

void ColumnData::InitializeScanWithOffset(ColumnScanState &state, idx_t row_idx) {
	state.current = GetSegment(row_idx);
	state.row_index = row_idx;
	state.internal_index = state.current->start;
	state.initialized = false;
	state.probe = true;
	state.was_point_read = false;
	state.version = version;
	state.scan_state.reset();
}

ColumnData::VectorCacheStats ColumnData::GetVectorCacheStats() {
	std::lock_guard<std::mutex> guard(vector_cache_lock);
	return vector_cache_stats;
}

bool ColumnData::ReadPackedRows(ColumnSegment *seg, idx_t row, idx_t count, data_ptr_t dst) {
	if (!seg || !seg->gpu_column || seg->owns_gpu_column || !seg->compacted || row < seg->start ||
	    row + count > seg->start + seg->count) {
		return false;
	}
	GPU_CHECK(db, adacom_scan_partial(seg->gpu_column, seg->gpu_index, row - seg->start, count, 0, dst, 0, 0));
	return true;
}

bool ColumnData::ReadHotVector(idx_t vec, idx_t offset, idx_t count, data_ptr_t dst) {
	const idx_t S = GetTypeIdSize(type);
	const idx_t n_slots = vector_cache_bytes / (STANDARD_VECTOR_SIZE * S);
	if (n_slots == 0) {
		return false;
	}
	{
		std::lock_guard<std::mutex> guard(vector_cache_lock);
		if (vector_cache.size() != n_slots) {
			vector_cache.assign(n_slots, CachedVector());
			vector_seen.assign(4 * n_slots, ~(idx_t)0);
		}
		CachedVector &slot = vector_cache[vec % n_slots];
		if (slot.vec == vec && offset + count <= slot.rows) {
			memcpy(dst, slot.data.data() + offset * S, count * S);
			vector_cache_stats.hits++;
			return true;
		}
		idx_t &seen = vector_seen[vec % vector_seen.size()];
		if (seen != vec) {
			seen = vec;  // first touch: the caller's ordinary path serves it
			vector_cache_stats.misses++;
			return false;
		}
	}
	// second touch (or a tail vector that has grown): decode the whole vector once through the ordinary scan path
	const idx_t first_row = vec * STANDARD_VECTOR_SIZE;
	const idx_t end = start + GetMaxEntry();
	if (first_row < start || first_row >= end) {
		return false;
	}
	const idx_t rows = std::min<idx_t>(STANDARD_VECTOR_SIZE, end - first_row);
	if (offset + count > rows) {
		return false;
	}
	Vector decoded(type);
	if (!ReadPackedRows(GetSegment(first_row), first_row, rows, decoded.buffer.data())) {
		// the vector crosses a segment border or lives (partly) on the host: the ordinary scan path
		ColumnScanState fill_state;
		InitializeScanWithOffset(fill_state, first_row);
		fill_state.probe = false;
		if (ScanVector(fill_state, decoded, rows) != rows) {
			return false;
		}
	}
	const uint8_t *src = FlatVector::GetData(decoded);
	memcpy(dst, src + offset * S, count * S);
	std::lock_guard<std::mutex> guard(vector_cache_lock);
	if (vector_cache.size() == n_slots) {
		CachedVector &slot = vector_cache[vec % n_slots];
		slot.vec = vec;
		slot.rows = rows;
		slot.data.assign(src, src + rows * S);
		vector_cache_stats.fills++;
	}
	return true;
}

idx_t ColumnData::ScanVector(ColumnScanState &state, Vector &result, idx_t remaining) {
	state.previous_states.clear();
	result.external = nullptr;
	if (state.was_point_read) {
		state.was_point_read = false;
		point_reads_continued.fetch_add(1, std::memory_order_relaxed);  // it was the start of a longer scan
	}
	if (state.probe) {
		// the first pull of a scan WithOffset: a point read when it stays inside one aligned vector
		state.probe = false;
		const idx_t offset = state.row_index % STANDARD_VECTOR_SIZE;
		const bool point_shaped = !state.initialized && state.version == version && vector_cache_bytes && remaining > 0 &&
		                          offset + remaining <= STANDARD_VECTOR_SIZE &&
		                          state.row_index + remaining <= start + GetMaxEntry() &&
		                          result.buffer.size() >= remaining * GetTypeIdSize(type);
		bool served = point_shaped &&
		              ReadHotVector(state.row_index / STANDARD_VECTOR_SIZE, offset, remaining, result.buffer.data());
		// first touch of a vector inside one packed segment: one decode of just these rows (one launch, one copy)
		// instead of a scan stream that would decode the segment ahead of a reader who is not coming back
		if (point_shaped && !served) {
			const uint32_t started = point_reads_started.load(std::memory_order_relaxed);
			const uint32_t continued = point_reads_continued.load(std::memory_order_relaxed);
			if (2 * continued <= started && ReadPackedRows(state.current, state.row_index, remaining, result.buffer.data())) {
				served = true;
				std::lock_guard<std::mutex> guard(vector_cache_lock);
				vector_cache_stats.lean_reads++;
			}
		}
		if (point_shaped) {
			// history of the scans that start WithOffset (halved now and then so that the column follows its workload)
			if (point_reads_started.fetch_add(1, std::memory_order_relaxed) >= 256) {
				point_reads_started.store(128, std::memory_order_relaxed);
				point_reads_continued.store(point_reads_continued.load(std::memory_order_relaxed) / 2, std::memory_order_relaxed);
			}
			state.was_point_read = true;
		}
		if (served) {
			db.column_segment_catalog.AddReadAccess(state.current);  // the policy still sees the read
			state.row_index += remaining;
			while (state.row_index >= state.current->start + state.current->count) {
				ColumnSegment *next = NextSegment(*this, state.current);
				if (!next) {
					break;
				}
				state.current = next;
			}
			state.internal_index = state.current->start;
			return remaining;
		}
	}
	if (state.version != version) {
		InitializeScanWithOffset(state, state.row_index);
		state.current->InitializeScan(state);
		state.initialized = true;
	} else if (!state.initialized) {
		state.current->InitializeScan(state);
		state.internal_index = state.current->start;
		state.initialized = true;
	}
	if (state.internal_index < state.row_index) {
		state.current->Skip(state);
	}
	idx_t initial_remaining = remaining;
	while (remaining > 0) {
		idx_t scan_count = std::min<idx_t>(remaining, state.current->start + state.current->count - state.row_index);
		idx_t result_offset = initial_remaining - remaining;
		if (scan_count > 0) {
			state.current->Scan(state, scan_count, result, result_offset, scan_count == initial_remaining);
			state.row_index += scan_count;
			remaining -= scan_count;
		}
		if (remaining > 0) {
			ColumnSegment *next = NextSegment(*this, state.current);
			if (!next) {
				break;
			}
			state.previous_states.emplace_back(std::move(state.scan_state));
			state.current = next;
			state.current->InitializeScan(state);
		}
	}
	state.internal_index = state.row_index;
	return initial_remaining - remaining;
}

void ColumnData::FetchRow(ColumnFetchState &state, row_t row_id, Vector &result, idx_t result_idx) {
	if (vector_cache_bytes && row_id >= (row_t)start && (idx_t)row_id < start + GetMaxEntry() &&
	    ReadHotVector((idx_t)row_id / STANDARD_VECTOR_SIZE, (idx_t)row_id % STANDARD_VECTOR_SIZE, 1,
	                  FlatVector::GetData(result) + result_idx * GetTypeIdSize(type))) {
		return;
	}
	auto segment = GetSegment((idx_t)row_id);
	segment->FetchRow(state, row_id, result, result_idx);
}

void ColumnData::FetchRows(const row_t *row_ids, idx_t count, Vector &result) {
	// the per-row slot is one launch + one PCIe round trip per row id; an index join hands over its whole batch:
	// one adacom_fetch_rows launch per device column the row ids touch
	const idx_t S = GetTypeIdSize(type);
	result.external = nullptr;
	if (result.buffer.size() < count * S) {
		result.buffer.resize(count * S);
	}
	const std::vector<ColumnSegment *> segs = SegmentSnapshot();
	if (segs.empty()) {
		throw InternalException("ColumnData::FetchRows: empty column");
	}
	struct Group {
		adacom_column *column;
		std::vector<uint32_t> seg_idx, row_in_seg;
		std::vector<idx_t> pos;
	};
	std::vector<Group> groups;
	ColumnFetchState fetch_state;
	const idx_t cap = segs.front()->SegmentSize() / S;
	idx_t last_group = 0;
	for (idx_t i = 0; i < count; i++) {
		const idx_t row = (idx_t)row_ids[i];
		// segments are full except the open one: guess by division, fall back to the tree search
		idx_t guess = std::min<idx_t>((row - start) / cap, segs.size() - 1);
		ColumnSegment *seg = segs[guess];
		if (row < seg->start || row >= seg->start + seg->count) {
			seg = GetSegment(row);
			if (row >= seg->start + seg->count) {
				throw InternalException("ColumnData::FetchRows: row id outside the column");
			}
		}
		if (!seg->gpu_column || seg->owns_gpu_column) {
			seg->FetchRow(fetch_state, row_ids[i], result, i);
			continue;
		}
		if (groups.empty() || groups[last_group].column != seg->gpu_column) {
			for (last_group = 0; last_group < groups.size() && groups[last_group].column != seg->gpu_column; last_group++) {
			}
			if (last_group == groups.size()) {
				groups.emplace_back();
				groups.back().column = seg->gpu_column;
			}
		}
		Group &g = groups[last_group];
		g.seg_idx.push_back((uint32_t)seg->gpu_index);
		g.row_in_seg.push_back((uint32_t)(row - seg->start));
		g.pos.push_back(i);
	}
	std::vector<uint8_t> out;
	for (auto &g : groups) {
		const idx_t n = g.pos.size();
		const bool in_place = groups.size() == 1 && n == count;  // every row from one device column: no scatter
		uint8_t *dst = result.buffer.data();
		if (!in_place) {
			out.resize(n * S);
			dst = out.data();
		}
		GPU_CHECK(db, adacom_fetch_rows(g.column, n, g.seg_idx.data(), g.row_in_seg.data(), 0, dst, 0));
		if (!in_place) {
			for (idx_t k = 0; k < n; k++) {
				memcpy(result.buffer.data() + g.pos[k] * S, out.data() + k * S, S);
			}
		}
	}
}

void ColumnData::SegmentSealed() {
	idx_t pending = 0;
	{
		std::lock_guard<std::mutex> guard(lock);
		for (idx_t i = segments.size(); i-- > 0;) {
			auto &s = segments[i];
			if (s->gpu_column || s->compacted) {
				break;
			}
			pending++;
		}
	}
	if (pending >= pending_limit) {
		CompactPending(false);
	}
}

static void AppendBits(std::vector<uint64_t> &dst, idx_t dst_bit, const uint64_t *src, idx_t n_bits) {
	for (idx_t i = 0; i < n_bits; i += 64) {
		uint64_t w = src[i >> 6];
		idx_t m = std::min<idx_t>(64, n_bits - i);
		if (m < 64) {
			w |= ~0ull << m;  // bits past the segment stay "valid"
		}
		idx_t pos = dst_bit + i;
		idx_t sh = pos & 63;
		// dst is initialised to all ones: clear where w is zero
		dst[pos >> 6] &= (w << sh) | ((1ull << sh) - 1);
		if (sh && (pos >> 6) + 1 < dst.size()) {
			dst[(pos >> 6) + 1] &= (w >> (64 - sh)) | (~0ull << sh);
		}
	}
}

void ColumnData::CompactPending(bool include_last) {
	std::vector<ColumnSegment *> pending;
	{
		std::lock_guard<std::mutex> guard(lock);
		for (idx_t i = 0; i < segments.size(); i++) {
			ColumnSegment *s = segments[i].get();
			if (!include_last && i + 1 == segments.size()) {
				continue;  // still open for appends
			}
			pending.push_back(s);
		}
	}
	CompactHostSegments(pending);
}

void ColumnData::CompactHostSegments(const std::vector<ColumnSegment *> &candidates) {
	std::lock_guard<std::mutex> guard(lock);
	std::vector<ColumnSegment *> pending;
	for (auto *s : candidates) {
		if (s->gpu_column || s->compacted || s->count == 0 || !s->succinct_possible) {
			continue;
		}
		pending.push_back(s);
	}
	if (pending.empty()) {
		return;
	}
	const idx_t S = GetTypeIdSize(type);
	const int flags = pending.front()->CompactFlags();
	std::vector<uint64_t> counts(pending.size());
	idx_t rows = 0;
	bool nulls = false;
	for (idx_t i = 0; i < pending.size(); i++) {
		counts[i] = pending[i]->count;
		rows += pending[i]->count;
		nulls |= pending[i]->staging_has_nulls;
	}
	Batch batch;
	batch.n = pending.size();
	adacom_ctx *ctx = db.Gpu(device);
	GPU_CHECK(db, adacom_column_create(ctx, ToAdacomType(type), pending.size(), counts.data(),
	                                   (uint32_t)(pending.front()->SegmentSize() / S), &batch.col));
	batches.push_back(batch);  // owned from here on (freed by the destructor even if a later step throws)
	Batch &b = batches.back();
	idx_t row = 0;
	for (auto *s : pending) {
		GPU_CHECK(db, adacom_column_upload_raw(b.col, s->host_block->data(), row, s->count));
		row += s->count;
	}
	if (nulls) {
		// validity of the whole batch (bit i = row i of the device column), kept in pinned host memory so that later
		// re-compactions of members (the policy) can hand it to the analyze kernel again
		std::vector<uint64_t> bits((rows + 63) / 64 + 1, ~0ull);
		row = 0;
		for (auto *s : pending) {
			AppendBits(bits, row, s->staging_validity.data(), s->count);
			row += s->count;
		}
		void *pinned = nullptr;
		GPU_CHECK(db, adacom_host_alloc(ctx, bits.size() * 8, &pinned));
		memcpy(pinned, bits.data(), bits.size() * 8);
		b.validity = (uint64_t *)pinned;
	}
	GPU_CHECK(db, adacom_analyze(b.col, 0, b.n, flags, b.validity, 0));
	GPU_CHECK(db, adacom_encode(b.col, 0, b.n, flags));
	std::vector<adacom_segment_meta> metas(b.n);
	GPU_CHECK(db, adacom_column_get_meta(b.col, 0, b.n, metas.data()));
	for (idx_t i = 0; i < pending.size(); i++) {
		ColumnSegment *s = pending[i];
		const idx_t size_before = s->GetDataSize();
		s->gpu_column = b.col;
		s->gpu_index = i;
		s->owns_gpu_column = false;
		s->uploaded = s->count;
		s->FinishCompact(size_before, metas[i]);
		s->DropHostBlock();  // the values live on the device now (EnsureStaging brings them back)
	}
}

// the policy's subsets: group the members by device column, ids ascending, one launch set per column
static std::map<adacom_column *, std::vector<ColumnSegment *>> GroupByColumn(const std::vector<CompactableSegment *> &in) {
	std::map<adacom_column *, std::vector<ColumnSegment *>> groups;
	for (auto *c : in) {
		auto *s = static_cast<ColumnSegment *>(c);
		if (s->gpu_column && !s->owns_gpu_column) {
			groups[s->gpu_column].push_back(s);
		}
	}
	for (auto &kv : groups) {
		std::sort(kv.second.begin(), kv.second.end(),
		          [](ColumnSegment *l, ColumnSegment *r) { return l->gpu_index < r->gpu_index; });
	}
	return groups;
}

void ColumnData::CompactMany(const std::vector<CompactableSegment *> &list) {
	std::vector<ColumnSegment *> host_only;
	std::vector<CompactableSegment *> members;
	for (auto *c : list) {
		auto *s = static_cast<ColumnSegment *>(c);
		if (s->compacted || s->num_elements == 0 || !s->succinct_possible) {
			continue;  // column_segment.cpp:278
		}
		if (!s->gpu_column) {
			host_only.push_back(s);  // never uploaded (adaptive mode): only the listed ones move to the device
		} else if (!s->owns_gpu_column) {
			members.push_back(s);
		} else {
			s->column_data = nullptr;  // detached earlier: compacts on its own
			s->Compact();
			s->column_data = this;
		}
	}
	if (!host_only.empty()) {
		std::sort(host_only.begin(), host_only.end(), [](ColumnSegment *l, ColumnSegment *r) { return l->start < r->start; });
		CompactHostSegments(host_only);
	}
	std::lock_guard<std::mutex> guard(lock);
	for (auto &kv : GroupByColumn(members)) {
		std::vector<uint32_t> ids;
		for (auto *s : kv.second) {
			ids.push_back((uint32_t)s->gpu_index);
		}
		const uint64_t *validity = nullptr;
		for (auto &b : batches) {
			if (b.col == kv.first) {
				validity = b.validity;
			}
		}
		const int flags = kv.second.front()->CompactFlags();
		GPU_CHECK(db, adacom_analyze_list(kv.first, ids.size(), ids.data(), flags, validity));
		GPU_CHECK(db, adacom_encode_list(kv.first, ids.size(), ids.data(), flags));
		const uint32_t lo = ids.front(), hi = ids.back() + 1;
		std::vector<adacom_segment_meta> metas(hi - lo);
		GPU_CHECK(db, adacom_column_get_meta(kv.first, lo, hi, metas.data()));
		for (auto *s : kv.second) {
			s->FinishCompact(s->GetDataSize(), metas[s->gpu_index - lo]);
			s->DropHostBlock();  // a re-compacted (cold again) segment gives its host block back
		}
	}
}

void ColumnData::UncompactMany(const std::vector<CompactableSegment *> &list) {
	std::vector<CompactableSegment *> members;
	for (auto *c : list) {
		auto *s = static_cast<ColumnSegment *>(c);
		if (!s->compacted || !s->function || s->function->type != CompressionType::COMPRESSION_SUCCINCT) {
			continue;  // column_segment.cpp:327-330
		}
		if (s->gpu_column && !s->owns_gpu_column) {
			members.push_back(s);
		} else {
			s->column_data = nullptr;
			s->Uncompact();
			s->column_data = this;
		}
	}
	std::lock_guard<std::mutex> guard(lock);
	for (auto &kv : GroupByColumn(members)) {
		std::vector<uint32_t> ids;
		std::vector<idx_t> sizes;
		for (auto *s : kv.second) {
			ids.push_back((uint32_t)s->gpu_index);
			sizes.push_back(s->GetDataSize());
		}
		GPU_CHECK(db, adacom_uncompact_list(kv.first, ids.size(), ids.data(), 0));
		// Uncompact() is what the policy does to the HOT segments: their flat block comes back to the host (as in the
		// reference, where a hot segment is an uncompressed block in RAM), so that scans and row fetches of hot
		// segments are host memory reads while the cold ones stay packed in HBM
		for (idx_t i = 0; i < kv.second.size(); i++) {
			if (hot_segments_on_host) {
				kv.second[i]->EnsureStaging();
			}
			kv.second[i]->FinishUncompact(sizes[i]);
		}
	}
}

// ---- filter pushdown ---------------------------------------------------------------------------------------

static adacom_cmp ToAdacomCmp(ExpressionType t) {
	switch (t) {
	case ExpressionType::COMPARE_EQUAL: return ADACOM_CMP_EQ;
	case ExpressionType::COMPARE_NOTEQUAL: return ADACOM_CMP_NE;
	case ExpressionType::COMPARE_LESSTHAN: return ADACOM_CMP_LT;
	case ExpressionType::COMPARE_GREATERTHAN: return ADACOM_CMP_GT;
	case ExpressionType::COMPARE_LESSTHANOREQUALTO: return ADACOM_CMP_LE;
	case ExpressionType::COMPARE_GREATERTHANOREQUALTO: return ADACOM_CMP_GE;
	default: throw InternalException("Unknown comparison type for filter pushed down to table!");  // column_segment.cpp:771
	}
}

static void FlattenFilter(const TableFilter &filter, TableFilterType level, std::vector<adacom_pred> &preds) {
	switch (filter.filter_type) {
	case TableFilterType::CONSTANT_COMPARISON: {
		auto &f = static_cast<const ConstantFilter &>(filter);
		preds.push_back(adacom_pred {ADACOM_PRED_COMPARE, (int32_t)ToAdacomCmp(f.comparison_type), f.constant});
		break;
	}
	case TableFilterType::IS_NULL:
		preds.push_back(adacom_pred {ADACOM_PRED_IS_NULL, 0, 0});
		break;
	case TableFilterType::IS_NOT_NULL:
		preds.push_back(adacom_pred {ADACOM_PRED_IS_NOT_NULL, 0, 0});
		break;
	case TableFilterType::CONJUNCTION_AND:
	case TableFilterType::CONJUNCTION_OR: {
		if (filter.filter_type != level) {
			throw InternalException("ColumnData::FilterAggregate: AND and OR levels mixed in one filter (chain "
			                        "adacom_scan_select through selection bitmaps)");
		}
		for (auto &child : static_cast<const ConjunctionFilter &>(filter).child_filters) {
			FlattenFilter(*child, level, preds);
		}
		break;
	}
	default:
		throw InternalException("FIXME: unsupported type for filter selection");  // column_segment.cpp:841
	}
}

std::vector<FilterStage> ColumnData::BuildFilterStages(const TableFilter &filter) {
	// The filter becomes a chain of stages, each one call that refines the selection of the one before (the recursion
	// of FilterSelection, column_segment.cpp:688-723): the leaves of the top level, then -- under an AND -- every OR
	// group of leaves (what an IN list or a pushed OR looks like).  An AND group under an OR would need the union of two
	// selections and is refused.
	std::vector<FilterStage> stages;
	const bool top_or = filter.filter_type == TableFilterType::CONJUNCTION_OR;
	FilterStage top;
	top.conj = top_or ? ADACOM_CONJ_OR : ADACOM_CONJ_AND;
	if (filter.filter_type == TableFilterType::CONJUNCTION_AND) {
		std::vector<const TableFilter *> todo {&filter};
		while (!todo.empty()) {
			const TableFilter *f = todo.back();
			todo.pop_back();
			if (f->filter_type == TableFilterType::CONJUNCTION_AND) {
				for (auto &child : static_cast<const ConjunctionFilter *>(f)->child_filters) {
					todo.push_back(child.get());
				}
			} else if (f->filter_type == TableFilterType::CONJUNCTION_OR) {
				FilterStage group;
				group.conj = ADACOM_CONJ_OR;
				FlattenFilter(*f, TableFilterType::CONJUNCTION_OR, group.preds);
				stages.push_back(std::move(group));
			} else {
				FlattenFilter(*f, TableFilterType::CONJUNCTION_AND, top.preds);
			}
		}
	} else {
		FlattenFilter(filter, top_or ? TableFilterType::CONJUNCTION_OR : TableFilterType::CONJUNCTION_AND, top.preds);
	}
	if (!top.preds.empty() || stages.empty()) {
		stages.insert(stages.begin(), std::move(top));
	}
	for (auto &st : stages) {
		if (st.preds.size() > ADACOM_MAX_PREDS) {
			throw InternalException("ColumnData::FilterAggregate: more than ADACOM_MAX_PREDS children in one conjunction");
		}
	}
	return stages;
}

FilterAggregateResult ColumnData::FilterAggregate(const TableFilter &filter) {
	const std::vector<FilterStage> stages = BuildFilterStages(filter);
	// sealed segments that still wait on the host join a shared device column first (what the first scan does)
	CompactPending(true);
	FilterAggregateResult result;
	auto add = [&](const adacom_agg &a) {
		result.approved_tuple_count += a.count;
		result.valid_count += a.reserved;
		const uint64_t lo = result.sum_lo + a.sum_lo;
		result.sum_hi += a.sum_hi + (lo < a.sum_lo ? 1 : 0);
		result.sum_lo = lo;
	};
	std::lock_guard<std::mutex> guard(lock);
	// selection bitmaps between the stages (device memory, 1 bit per row of the largest run)
	adacom_ctx *ctx = db.Gpu(device);
	void *bitmaps[2] = {nullptr, nullptr};
	struct Release {
		adacom_ctx *ctx;
		void **b;
		~Release() {
			adacom_device_free(ctx, b[0]);
			adacom_device_free(ctx, b[1]);
		}
	} release {ctx, bitmaps};
	idx_t bitmap_rows = 0;
	auto run_stages = [&](adacom_column *column, idx_t seg_b, idx_t seg_e, idx_t rows, const uint64_t *validity) {
		if (stages.size() > 1 && rows > bitmap_rows) {
			for (auto &bm : bitmaps) {
				adacom_device_free(ctx, bm);
				bm = nullptr;
				GPU_CHECK(db, adacom_device_alloc(ctx, ((rows + 63) / 64 + 2) * 8, &bm));
			}
			bitmap_rows = rows;
		}
		adacom_agg agg;
		for (idx_t k = 0; k < stages.size(); k++) {
			const bool last = k + 1 == stages.size();
			GPU_CHECK(db, adacom_scan_select(column, seg_b, seg_e, 0, stages[k].preds.data(), (uint32_t)stages[k].preds.size(),
			                                 stages[k].conj, validity, 0, k ? (const uint64_t *)bitmaps[(k - 1) & 1] : nullptr,
			                                 last ? nullptr : (uint64_t *)bitmaps[k & 1], &agg));
		}
		add(agg);
	};
	// runs of consecutive members of one shared device column -> one chain of calls; a segment on its own (detached by
	// an append after Uncompact, or never compacted) is scanned through its own single-segment column
	idx_t i = 0;
	while (i < segments.size()) {
		ColumnSegment *s = segments[i].get();
		if (s->count == 0) {
			i++;
			continue;
		}
		if (s->gpu_column && !s->owns_gpu_column) {
			idx_t j = i + 1, rows = s->count;
			while (j < segments.size() && segments[j]->gpu_column == s->gpu_column && !segments[j]->owns_gpu_column &&
			       segments[j]->gpu_index == s->gpu_index + (j - i)) {
				rows += segments[j]->count;
				j++;
			}
			const uint64_t *validity = nullptr;
			for (auto &b : batches) {
				if (b.col == s->gpu_column) {
					validity = b.validity;
				}
			}
			run_stages(s->gpu_column, s->gpu_index, s->gpu_index + (j - i), rows, validity);
			i = j;
		} else {
			s->Upload();
			run_stages(s->gpu_column, 0, 1, s->count, s->staging_has_nulls ? s->staging_validity.data() : nullptr);
			i++;
		}
	}
	return result;
}

}  // namespace adacom_host
