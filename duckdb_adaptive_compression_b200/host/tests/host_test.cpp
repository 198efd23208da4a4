// host_test.cpp -- drives the C++ shim the way DuckDB drives the reference (needs a B200):
//   ColumnData::AppendData      -> ColumnSegment::Append in 2048-value vectors, new segment when full
//   ColumnData::ScanVector      -> ColumnSegment::Scan with row-group aligned vectors (start = 2048k + c),
//                                  vectors straddling two segments arrive as two scan_partial calls
//   ColumnData::FetchRow        -> ColumnSegment::FetchRow
//   ColumnSegmentCatalog        -> CompactAllSegments / RunCompactionEpoch / GetTotalDataSize
//   ColumnDataCheckpointer      -> init_analyze/analyze/final_analyze, init_compression/compress/compress_finalize
#include "../adacom_host.hpp"

#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <functional>
#include <random>
#include <thread>

using namespace adacom_host;

#define CHECK(cond)                                                           \
	do {                                                                      \
		if (!(cond)) {                                                        \
			fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
			exit(1);                                                          \
		}                                                                     \
	} while (0)

template <class T>
struct Column {  // the slice of ColumnData the succinct path needs
	DatabaseInstance &db;
	PhysicalType type;
	std::vector<unique_ptr<ColumnSegment>> segments;
	ColumnAppendState append_state;
	idx_t rows = 0;
	Column(DatabaseInstance &db, PhysicalType type) : db(db), type(type) {
	}
	void AppendTransientSegment(idx_t start_row) {
		segments.push_back(ColumnSegment::CreateTransientSegment(db, type, start_row));
		segments.back()->InitializeAppend(append_state);
	}
	void AppendData(const T *values, const uint64_t *validity, idx_t count) {  // column_data.cpp:249-270
		UnifiedVectorFormat vdata;
		vdata.data = (data_ptr_t)values;
		vdata.validity.mask = validity;
		idx_t offset = 0;
		while (count > 0) {
			if (segments.empty()) {
				AppendTransientSegment(rows);
			}
			idx_t copied = segments.back()->Append(append_state, vdata, offset, count);
			rows += copied;
			if (copied == count) {
				break;
			}
			AppendTransientSegment(rows);
			offset += copied;
			count -= copied;
		}
	}
	ColumnSegment *GetSegment(idx_t row) {
		for (auto &s : segments) {
			if (row >= s->start && row < s->start + s->count) {
				return s.get();
			}
		}
		return nullptr;
	}
	// ColumnData::ScanVector (column_data.cpp:92-139): one <=2048 vector starting at `row`
	idx_t ScanVector(ColumnScanState &state, idx_t row, Vector &result) {
		idx_t remaining = std::min<idx_t>(STANDARD_VECTOR_SIZE, rows - row);
		idx_t scanned = 0;
		while (remaining > 0) {
			ColumnSegment *seg = GetSegment(row + scanned);
			if (state.current != seg) {
				state.current = seg;
				seg->InitializeScan(state);
			}
			state.row_index = row + scanned;
			idx_t in_seg = std::min<idx_t>(remaining, seg->start + seg->count - state.row_index);
			bool entire = scanned == 0 && in_seg == remaining;
			seg->Scan(state, in_seg, result, scanned, entire);
			scanned += in_seg;
			remaining -= in_seg;
		}
		return scanned;
	}
};

template <class T>
void RunColumn(DatabaseInstance &db, PhysicalType type, idx_t n, uint64_t span, int64_t base, bool nulls) {
	std::mt19937_64 rng(42 + n + span);
	std::vector<T> values(n);
	for (auto &v : values) {
		v = (T)(base + (int64_t)(rng() % span));
	}
	std::vector<uint64_t> validity((n + 63) / 64, ~0ull);
	if (nulls) {
		for (idx_t i = 0; i < n; i += 7) {
			validity[i >> 6] &= ~(1ull << (i & 63));
		}
	}
	Column<T> col(db, type);
	for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
		idx_t m = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - off);
		// each call sees vector-local validity bits, like UnifiedVectorFormat
		std::vector<uint64_t> local((m + 63) / 64, ~0ull);
		for (idx_t i = 0; i < m; i++) {
			if (!((validity[(off + i) >> 6] >> ((off + i) & 63)) & 1)) {
				local[i >> 6] &= ~(1ull << (i & 63));
			}
		}
		col.AppendData(values.data() + off, nulls ? local.data() : nullptr, m);
	}
	CHECK(col.rows == n);
	idx_t cap = BLOCK_SIZE / sizeof(T);
	CHECK(col.segments.size() == (n + cap - 1) / cap);
	// full segments were compacted by Append (column_segment.cpp:266-268); the last one compacts lazily on scan
	for (size_t g = 0; g + 1 < col.segments.size(); g++) {
		CHECK(col.segments[g]->IsBitCompressed());
	}
	// full scan in row-aligned vectors, as TableScan would
	ColumnScanState state;
	Vector result(type);
	for (idx_t row = 0; row < n; row += STANDARD_VECTOR_SIZE) {
		idx_t got = col.ScanVector(state, row, result);
		auto out = (T *)FlatVector::GetData(result);
		for (idx_t i = 0; i < got; i++) {
			bool valid = (validity[(row + i) >> 6] >> ((row + i) & 63)) & 1;
			if (valid) {
				CHECK(out[i] == values[row + i]);
			}
		}
	}
	CHECK(col.segments.back()->IsBitCompressed());
	// sizes: sum of 9 + 8*ceil(n*w/64)
	idx_t total = 0;
	for (auto &s : col.segments) {
		idx_t w = s->Width();
		idx_t expect = w < 8 * sizeof(T) ? 9 + 8 * ((s->count * w + 63) / 64) : 9 + 8 * ((s->SegmentSize() * 8 + 63) / 64);
		CHECK(s->GetDataSize() == expect);
		total += expect;
	}
	CHECK(db.column_segment_catalog.GetTotalDataSize() >= total);
	// fetch_row (intended semantics)
	ColumnFetchState fstate;
	Vector one(type);
	for (int q = 0; q < 200; q++) {
		idx_t row = rng() % n;
		if (!((validity[row >> 6] >> (row & 63)) & 1)) {
			continue;
		}
		col.GetSegment(row)->FetchRow(fstate, (row_t)row, one, 3);
		CHECK(((T *)FlatVector::GetData(one))[3] == values[row]);
	}
	// Uncompact + re-scan + append after uncompact
	col.segments[0]->Uncompact();
	CHECK(!col.segments[0]->IsBitCompressed());
	CHECK(col.segments[0]->GetDataSize() == col.segments[0]->SegmentSize());
	ColumnScanState state2;
	idx_t got = col.ScanVector(state2, 0, result);
	for (idx_t i = 0; i < got; i++) {
		if ((validity[i >> 6] >> (i & 63)) & 1) {
			CHECK(((T *)FlatVector::GetData(result))[i] == values[i]);
		}
	}
	CHECK(col.segments[0]->IsBitCompressed());  // lazily compacted again by the scan (non-adaptive mode)
}

// ColumnData: the segments of a column share one device column -- batched compaction, one scan stream across segment
// borders (zero-copy whole vectors), batched policy epochs, batched fetches, NULLs, appends after Uncompact.
template <class T>
void RunColumnData(PhysicalType type, idx_t n, uint64_t span, int64_t base, bool nulls, bool adaptive, idx_t pending_limit) {
	DBConfig cfg;
	cfg.adaptive_succinct_compression_enabled = adaptive;
	DatabaseInstance db(&cfg);
	std::mt19937_64 rng(7 + n + span);
	std::vector<T> values(n);
	for (auto &v : values) {
		v = (T)(base + (int64_t)(rng() % span));
	}
	auto valid = [&](idx_t row) { return !nulls || row % 7 != 3; };
	ColumnData col(db, type);
	col.pending_limit = pending_limit;
	ColumnAppendState astate;
	col.InitializeAppend(astate);
	for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
		idx_t m = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - off);
		std::vector<uint64_t> local((m + 63) / 64, ~0ull);
		for (idx_t i = 0; i < m; i++) {
			if (!valid(off + i)) {
				local[i >> 6] &= ~(1ull << (i & 63));
			}
		}
		UnifiedVectorFormat vdata;
		vdata.data = (data_ptr_t)(values.data() + off);
		vdata.validity.mask = nulls ? local.data() : nullptr;
		col.AppendData(astate, vdata, m);
	}
	CHECK(col.GetMaxEntry() == n);
	idx_t cap = BLOCK_SIZE / sizeof(T);
	CHECK(col.segments.size() == (n + cap - 1) / cap);
	if (adaptive) {
		for (auto &s : col.segments) {
			CHECK(!s->IsBitCompressed());
		}
		CHECK(db.column_segment_catalog.GetTotalDataSize() == col.segments.size() * BLOCK_SIZE);
		db.column_segment_catalog.CompactAllSegments();  // one launch set for the whole column
		CHECK(col.batches.size() == 1);
	}
	// full scan, row-aligned vectors; whole vectors inside a segment must arrive zero-copy
	auto full_scan = [&]() {
		ColumnScanState state;
		col.InitializeScan(state);
		Vector result(type);
		idx_t zero_copy = 0;
		for (idx_t row = 0; row < n;) {
			idx_t got = col.ScanVector(state, result, std::min<idx_t>(STANDARD_VECTOR_SIZE, n - row));
			CHECK(got == std::min<idx_t>(STANDARD_VECTOR_SIZE, n - row));
			zero_copy += result.external != nullptr;
			auto out = (T *)FlatVector::GetData(result);
			for (idx_t i = 0; i < got; i++) {
				if (valid(row + i)) {
					CHECK(out[i] == values[row + i]);
				}
			}
			row += got;
		}
		return zero_copy;
	};
	idx_t zero_copy = full_scan();
	idx_t vectors = (n + STANDARD_VECTOR_SIZE - 1) / STANDARD_VECTOR_SIZE;
	CHECK(zero_copy + 2 * col.segments.size() >= vectors);  // all but the vectors that straddle a segment border
	for (auto &s : col.segments) {
		CHECK(s->IsBitCompressed() && s->gpu_column && !s->owns_gpu_column && s->host_block->empty());
	}
	// sizes: 9 + 8*ceil(n*w/64) per segment; catalog and buffer manager agree
	idx_t total = 0;
	for (auto &s : col.segments) {
		idx_t w = s->Width();
		idx_t expect = w < 8 * sizeof(T) ? 9 + 8 * ((s->count * w + 63) / 64) : 9 + 8 * ((s->SegmentSize() * 8 + 63) / 64);
		CHECK(s->GetDataSize() == expect);
		total += expect;
	}
	CHECK(db.column_segment_catalog.GetTotalDataSize() == total);
	CHECK(db.buffer_manager.GetDataSize() == total);
	auto per_device = db.column_segment_catalog.GetDataSizePerDevice();
	CHECK(per_device.size() == 1 && per_device[0] == total);
	// a scan from the middle (InitializeScanWithOffset), crossing batches when pending_limit split the column
	{
		ColumnScanState state;
		idx_t row = (n / 2) / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE;
		col.InitializeScanWithOffset(state, row);
		Vector result(type);
		for (int k = 0; k < 40 && row < n; k++) {
			idx_t got = col.ScanVector(state, result, std::min<idx_t>(STANDARD_VECTOR_SIZE, n - row));
			auto out = (T *)FlatVector::GetData(result);
			for (idx_t i = 0; i < got; i++) {
				if (valid(row + i)) {
					CHECK(out[i] == values[row + i]);
				}
			}
			row += got;
		}
	}
	// fetch_row per row and batched
	ColumnFetchState fstate;
	Vector one(type);
	std::vector<row_t> ids;
	for (int q = 0; q < 300; q++) {
		idx_t row = rng() % n;
		if (!valid(row)) {
			continue;
		}
		ids.push_back((row_t)row);
		if (q < 40) {
			col.FetchRow(fstate, (row_t)row, one, 5);
			CHECK(((T *)FlatVector::GetData(one))[5] == values[row]);
		}
	}
	Vector many(type);
	col.FetchRows(ids.data(), ids.size(), many);
	for (idx_t k = 0; k < ids.size(); k++) {
		CHECK(((T *)FlatVector::GetData(many))[k] == values[ids[k]]);
	}
	// policy epoch: hot segments (read above by the fetches / scans) are uncompacted in ONE call, cold re-compacted
	if (col.segments.size() >= 10) {
		for (int rep = 0; rep < 50; rep++) {
			db.column_segment_catalog.AddReadAccess(col.segments[1].get());
		}
		std::vector<idx_t> sizes_before;
		for (auto &s : col.segments) {
			sizes_before.push_back(s->GetDataSize());
		}
		auto res = db.column_segment_catalog.RunCompactionEpoch();
		CHECK(res.first + res.second == col.segments.size());
		CHECK(!col.segments[1]->IsBitCompressed());
		CHECK(col.segments[1]->GetDataSize() == col.segments[1]->SegmentSize());
		idx_t after = db.column_segment_catalog.GetTotalDataSize();
		// every uncompacted segment is back at the flat block size (262136 B; its succinct image may have been
		// LARGER: a slot-width int_vector carries a 9-byte header -- INT16 with a negative base under the adaptive
		// encoder's zero-extended min/max does not shrink, quirk Q2b)
		idx_t expect_after = 0;
		for (size_t g = 0; g < col.segments.size(); g++) {
			expect_after += col.segments[g]->IsBitCompressed() ? sizes_before[g] : col.segments[g]->SegmentSize();
		}
		CHECK(after == expect_after && after != total);
		CHECK(db.buffer_manager.GetDataSize() == after);
		full_scan();  // values unchanged; non-adaptive mode re-compacts lazily, adaptive serves the flat image
		if (adaptive) {
			CHECK(!col.segments[1]->IsBitCompressed());
			db.column_segment_catalog.CompactAllSegments();
		}
		CHECK(col.segments[1]->IsBitCompressed());
		CHECK(db.column_segment_catalog.GetTotalDataSize() == total);
		CHECK(db.buffer_manager.GetDataSize() == total);
	}
	// append after the column was compacted: the open last segment leaves the shared column, continues on the host
	{
		idx_t extra = 3000;
		std::vector<T> more(extra);
		for (auto &v : more) {
			v = (T)(base + (int64_t)(rng() % span));
		}
		UnifiedVectorFormat vdata;
		vdata.data = (data_ptr_t)more.data();
		for (idx_t off = 0; off < extra; off += STANDARD_VECTOR_SIZE) {
			UnifiedVectorFormat part = vdata;
			part.data = (data_ptr_t)(more.data() + off);
			col.AppendData(astate, part, std::min<idx_t>(STANDARD_VECTOR_SIZE, extra - off));
		}
		values.insert(values.end(), more.begin(), more.end());
		idx_t n_old = n;
		n += extra;
		CHECK(col.GetMaxEntry() == n);
		ColumnScanState state;
		col.InitializeScanWithOffset(state, n_old / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE);
		Vector result(type);
		for (idx_t row = n_old / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE; row < n;) {
			idx_t got = col.ScanVector(state, result, std::min<idx_t>(STANDARD_VECTOR_SIZE, n - row));
			auto out = (T *)FlatVector::GetData(result);
			for (idx_t i = 0; i < got; i++) {
				if (row + i >= n_old || (row + i) % 7 != 3) {
					CHECK(out[i] == values[row + i]);
				}
			}
			row += got;
		}
	}
	// table filters pushed to the column: FilterSelection + COUNT / SUM over the approved rows, on the shared device
	// columns AND the detached last segment (appended to above), against a scalar loop over the inserted values
	{
		const idx_t first_appended = n - 3000;
		auto row_valid = [&](idx_t row) { return row >= first_appended || valid(row); };
		const int64_t lo = base + (int64_t)(span / 4), hi = base + (int64_t)(span / 2);
		auto check = [&](const TableFilter &filter, const std::function<bool(bool, int64_t)> &pass) {
			FilterAggregateResult got = col.FilterAggregate(filter);
			idx_t approved = 0, nvalid = 0;
			unsigned __int128 sum = 0;
			for (idx_t row = 0; row < n; row++) {
				const bool v = row_valid(row);
				if (pass(v, (int64_t)values[row])) {
					approved++;
					if (v) {
						nvalid++;
						sum += (unsigned __int128)(__int128)(int64_t)values[row];
					}
				}
			}
			CHECK(got.approved_tuple_count == approved);
			CHECK(got.valid_count == nvalid);
			CHECK(got.sum_lo == (uint64_t)sum && got.sum_hi == (uint64_t)(sum >> 64));
		};
		check(ConstantFilter(ExpressionType::COMPARE_LESSTHAN, hi), [&](bool v, int64_t x) { return v && x < hi; });
		check(ConstantFilter(ExpressionType::COMPARE_NOTEQUAL, lo), [&](bool v, int64_t x) { return v && x != lo; });
		ConjunctionAndFilter between;
		between.child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_GREATERTHANOREQUALTO, lo)));
		between.child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_LESSTHANOREQUALTO, hi)));
		check(between, [&](bool v, int64_t x) { return v && x >= lo && x <= hi; });
		ConjunctionOrFilter outside;
		outside.child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_LESSTHAN, lo)));
		outside.child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_GREATERTHAN, hi)));
		check(outside, [&](bool v, int64_t x) { return v && (x < lo || x > hi); });
		check(IsNullFilter(), [&](bool v, int64_t) { return !v; });
		check(IsNotNullFilter(), [&](bool v, int64_t) { return v; });
		ConjunctionOrFilter or_null;
		or_null.child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_EQUAL, lo)));
		or_null.child_filters.push_back(unique_ptr<TableFilter>(new IsNullFilter()));
		check(or_null, [&](bool v, int64_t x) { return !v || x == lo; });
		// x >= lo AND (x = a OR x = b OR x > hi) AND x IS NOT NULL: the OR group refines the selection of the leaves
		{
			const int64_t a = (int64_t)values[n / 3], b = (int64_t)values[n / 2];
			ConjunctionAndFilter in_list;
			in_list.child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_GREATERTHANOREQUALTO, lo)));
			auto group = new ConjunctionOrFilter();
			group->child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_EQUAL, a)));
			group->child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_EQUAL, b)));
			group->child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_GREATERTHAN, hi)));
			in_list.child_filters.push_back(unique_ptr<TableFilter>(group));
			in_list.child_filters.push_back(unique_ptr<TableFilter>(new IsNotNullFilter()));
			check(in_list, [&](bool v, int64_t x) { return v && x >= lo && (x == a || x == b || x > hi); });
			// only OR groups under the AND
			ConjunctionAndFilter two_groups;
			auto g1 = new ConjunctionOrFilter(), g2 = new ConjunctionOrFilter();
			g1->child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_LESSTHAN, hi)));
			g1->child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_EQUAL, b)));
			g2->child_filters.push_back(unique_ptr<TableFilter>(new ConstantFilter(ExpressionType::COMPARE_GREATERTHAN, lo)));
			g2->child_filters.push_back(unique_ptr<TableFilter>(new IsNullFilter()));
			two_groups.child_filters.push_back(unique_ptr<TableFilter>(g1));
			two_groups.child_filters.push_back(unique_ptr<TableFilter>(g2));
			check(two_groups, [&](bool v, int64_t x) { return v && (x < hi || x == b) && x > lo; });
		}
		// an AND group under an OR (a union of selections) is refused, not mis-evaluated
		ConjunctionOrFilter mixed;
		mixed.child_filters.push_back(unique_ptr<TableFilter>(new ConjunctionAndFilter()));
		bool threw = false;
		try {
			col.FilterAggregate(mixed);
		} catch (const InternalException &) {
			threw = true;
		}
		CHECK(threw);
	}
	// hot-vector cache: point reads (scan WithOffset + one pull, FetchRow) that come back to a vector are served from
	// host memory after the second touch, with the same values; a tail vector that grows is refreshed
	{
		auto before = col.GetVectorCacheStats();
		Vector result(type);
		ColumnFetchState fstate2;
		const idx_t probes[] = {0, 2048 * 3, n / 2 / 2048 * 2048, (n - 1) / 2048 * 2048};
		for (int rep = 0; rep < 4; rep++) {
			for (idx_t p : probes) {
				const idx_t want = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - p);
				ColumnScanState state;
				col.InitializeScanWithOffset(state, p);
				CHECK(col.ScanVector(state, result, want) == want);
				auto out = (T *)FlatVector::GetData(result);
				for (idx_t i = 0; i < want; i++) {
					if (p + i >= n - 3000 || valid(p + i)) {
						CHECK(out[i] == values[p + i]);
					}
				}
				// the scan continues correctly behind a vector that came from the cache
				if (p + want < n) {
					const idx_t more = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - p - want);
					CHECK(col.ScanVector(state, result, more) == more);
					out = (T *)FlatVector::GetData(result);
					for (idx_t i = 0; i < more; i++) {
						if (p + want + i >= n - 3000 || valid(p + want + i)) {
							CHECK(out[i] == values[p + want + i]);
						}
					}
				}
				const idx_t row = p + 17 < n ? p + 17 : p;
				if (row >= n - 3000 || valid(row)) {
					col.FetchRow(fstate2, (row_t)row, result, 3);
					CHECK(((T *)FlatVector::GetData(result))[3] == values[row]);
				}
			}
		}
		auto after = col.GetVectorCacheStats();
		CHECK(after.fills - before.fills >= 4 && after.hits - before.hits >= 8);
		// scans that start WithOffset and keep pulling (a parallel table scan starts every row group like that): the
		// column learns that they are not point reads and stops paying the one-launch first touch for them
		{
			auto s0 = col.GetVectorCacheStats();
			const idx_t n_vec = n / STANDARD_VECTOR_SIZE;
			for (int k = 0; k < 60 && n_vec > 8; k++) {
				const idx_t p = ((idx_t)(k * 37 + 5) % (n_vec - 4)) * STANDARD_VECTOR_SIZE;
				ColumnScanState state;
				col.InitializeScanWithOffset(state, p);
				for (int pull = 0; pull < 3; pull++) {
					CHECK(col.ScanVector(state, result, STANDARD_VECTOR_SIZE) == STANDARD_VECTOR_SIZE);
					auto out = (T *)FlatVector::GetData(result);
					const idx_t row = p + pull * STANDARD_VECTOR_SIZE + 11;
					if (row >= n - 3000 || valid(row)) {
						CHECK(out[11] == values[row]);
					}
				}
			}
			auto s1 = col.GetVectorCacheStats();
			CHECK(n_vec <= 8 || s1.lean_reads - s0.lean_reads <= 16);
		}
		// the tail vector grows: rows behind what the entry held are read correctly
		std::vector<T> more(700);
		for (auto &v : more) {
			v = (T)(base + (int64_t)(rng() % span));
		}
		UnifiedVectorFormat vdata;
		vdata.data = (data_ptr_t)more.data();
		col.AppendData(astate, vdata, more.size());
		values.insert(values.end(), more.begin(), more.end());
		const idx_t n_new = n + more.size();
		for (int rep = 0; rep < 3; rep++) {
			col.FetchRow(fstate2, (row_t)(n_new - 1), result, 0);
			CHECK(((T *)FlatVector::GetData(result))[0] == values[n_new - 1]);
			col.FetchRow(fstate2, (row_t)(n + 5), result, 1);
			CHECK(((T *)FlatVector::GetData(result))[1] == values[n + 5]);
		}
		CHECK(col.GetMaxEntry() == n_new);
	}
}

// Readers (full scans with zero-copy vectors, row fetches) against a policy thread that keeps moving segments between
// "hot: flat block in host memory" and "cold: packed in HBM": the values never change, no reader may see a freed block.
void RunConcurrentPolicy() {
	DBConfig cfg;
	cfg.adaptive_succinct_compression_enabled = true;
	DatabaseInstance db(&cfg);
	const idx_t n = 32767 * 14 + 123;
	std::vector<int64_t> values(n);
	std::mt19937_64 rng(99);
	for (auto &v : values) {
		v = 5000000000ll + (int64_t)(rng() % (1ull << 21));
	}
	ColumnData col(db, PhysicalType::INT64);
	ColumnAppendState astate;
	col.InitializeAppend(astate);
	for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
		UnifiedVectorFormat vdata;
		vdata.data = (data_ptr_t)(values.data() + off);
		col.AppendData(astate, vdata, std::min<idx_t>(STANDARD_VECTOR_SIZE, n - off));
	}
	std::atomic<bool> stop {false};
	std::atomic<int> errors {0};
	std::atomic<idx_t> scans {0};
	std::vector<std::thread> readers;
	for (int t = 0; t < 3; t++) {
		readers.emplace_back([&, t]() {
			std::mt19937_64 r(t);
			Vector result(PhysicalType::INT64), one(PhysicalType::INT64);
			ColumnFetchState fstate;
			while (!stop.load()) {
				ColumnScanState state;
				col.InitializeScan(state);
				for (idx_t row = 0; row < n;) {
					idx_t got = col.ScanVector(state, result, std::min<idx_t>(STANDARD_VECTOR_SIZE, n - row));
					auto out = (const int64_t *)FlatVector::GetData(result);
					for (idx_t i = 0; i < got; i += 97) {
						errors += out[i] != values[row + i];
					}
					row += got;
				}
				for (int q = 0; q < 50; q++) {
					idx_t row = r() % n;
					col.FetchRow(fstate, (row_t)row, one, 0);
					errors += ((int64_t *)FlatVector::GetData(one))[0] != values[row];
				}
				scans++;
			}
		});
	}
	for (int epoch = 0; epoch < 40; epoch++) {
		for (int rep = 0; rep < 2000; rep++) {
			db.column_segment_catalog.AddReadAccess(col.segments[(epoch * 3 + rep % 2) % col.segments.size()].get());
		}
		db.column_segment_catalog.RunCompactionEpoch();
	}
	stop = true;
	for (auto &t : readers) {
		t.join();
	}
	CHECK(errors.load() == 0);
	CHECK(scans.load() > 0);
}

int main() {
	RunConcurrentPolicy();
	RunColumnData<int64_t>(PhysicalType::INT64, 32767 * 12 + 500, 1ull << 33, -40000000000ll, false, false, 64);
	RunColumnData<int64_t>(PhysicalType::INT64, 32767 * 12 + 500, 1ull << 33, -40000000000ll, false, false, 5);
	RunColumnData<uint32_t>(PhysicalType::UINT32, 65534 * 11 + 77, 1ull << 13, 1000, true, false, 4);
	RunColumnData<int16_t>(PhysicalType::INT16, 131068 * 10, 1ull << 7, -50, false, true, 64);
	RunColumnData<uint8_t>(PhysicalType::UINT8, 262136 * 3 + 9, 1ull << 3, 2, true, true, 64);
	DatabaseInstance db;
	idx_t base_size = db.buffer_manager.GetDataSize();
	CHECK(base_size == 0);
	RunColumn<int64_t>(db, PhysicalType::INT64, 100000, 1ull << 40, -5000000000000ll, false);
	RunColumn<uint32_t>(db, PhysicalType::UINT32, 200000, 1ull << 16, 1000000, false);
	RunColumn<int32_t>(db, PhysicalType::INT32, 150000, 1ull << 12, -100000, true);
	RunColumn<int16_t>(db, PhysicalType::INT16, 300000, 1ull << 9, 100, false);
	RunColumn<uint8_t>(db, PhysicalType::UINT8, 300000, 1ull << 5, 3, false);

	// unsupported types raise InternalException exactly like SuccinctFun::GetFunction
	bool threw = false;
	try {
		SuccinctGpuFun::GetFunction(PhysicalType::DOUBLE);
	} catch (const InternalException &) {
		threw = true;
	}
	CHECK(threw);
	CHECK(!SuccinctGpuFun::TypeIsSupported(PhysicalType::INT128));
	CHECK(SuccinctGpuFun::TypeIsSupported(PhysicalType::UINT16));

	// adaptive mode: segments start UNCOMPRESSED, the policy compacts the cold 90 % and leaves the hot ones flat
	{
		DBConfig cfg;
		cfg.adaptive_succinct_compression_enabled = true;
		DatabaseInstance adb(&cfg);
		Column<uint32_t> col(adb, PhysicalType::UINT32);
		idx_t n = 65534 * 10;
		std::vector<uint32_t> values(n);
		for (idx_t i = 0; i < n; i++) {
			values[i] = (uint32_t)i;
		}
		for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
			col.AppendData(values.data() + off, nullptr, std::min<idx_t>(STANDARD_VECTOR_SIZE, n - off));
		}
		CHECK(col.segments.size() == 10);
		size_t before = adb.column_segment_catalog.GetTotalDataSize();
		CHECK(before == 10 * BLOCK_SIZE);
		ColumnScanState state;
		Vector result(PhysicalType::UINT32);
		for (int rep = 0; rep < 5; rep++) {  // segment 9 is hot
			col.ScanVector(state, 65534 * 9 + 2, result);
		}
		for (auto &s : col.segments) {
			CHECK(!s->IsBitCompressed());  // background mode: no lazy compaction on scan
		}
		auto res = adb.column_segment_catalog.RunCompactionEpoch();
		CHECK(res.first == 9 && res.second == 1);  // 9/10 = 0.9f < 0.90 (reference float arithmetic)
		CHECK(!col.segments[9]->IsBitCompressed());
		size_t after = adb.column_segment_catalog.GetTotalDataSize();
		CHECK(after == 9 * (9 + 8 * ((65534 * 16 + 63) / 64)) + 1 * BLOCK_SIZE);
		// values survive (the build stores the frame of reference; reference quirk Q2 loses it)
		for (idx_t row = 0; row < n; row += 50000) {
			col.ScanVector(state, row, result);
			CHECK(((uint32_t *)FlatVector::GetData(result))[0] == values[row]);
		}
	}

	// checkpoint-style slots with real meaning: exact size from analyze, segments from compress
	{
		DatabaseInstance cdb;
		BindAnalyzeDatabase(&cdb);
		CompressionFunction fn = SuccinctGpuFun::GetFunction(PhysicalType::UINT32);
		idx_t n = 200000;
		std::vector<uint32_t> values(n);
		for (idx_t i = 0; i < n; i++) {
			values[i] = (uint32_t)i;
		}
		ColumnData *no_column_data = nullptr;
		auto astate = fn.init_analyze(*no_column_data, PhysicalType::UINT32);
		Vector v(PhysicalType::UINT32);
		for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
			idx_t m = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - off);
			memcpy(v.buffer.data(), values.data() + off, m * 4);
			CHECK(fn.analyze(*astate, v, m));
		}
		idx_t score = fn.final_analyze(*astate);
		// 65534 x3 at 16/17/18 bits + 3398: 9 + 8*ceil(n*w/64) each
		idx_t expect = 0;
		idx_t counts[4] = {65534, 65534, 65534, 3398};
		idx_t first = 0;
		for (idx_t c : counts) {
			uint64_t range = c - 1;
			idx_t w = 1;
			while ((range >> w) != 0) {
				w++;
			}
			expect += 9 + 8 * ((c * w + 63) / 64);
			first += c;
		}
		CHECK(score == expect);
		ColumnDataCheckpointer *cp = NewCheckpointer(cdb, PhysicalType::UINT32, 0);
		auto cstate = fn.init_compression(*cp, std::move(astate));
		for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
			idx_t m = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - off);
			memcpy(v.buffer.data(), values.data() + off, m * 4);
			fn.compress(*cstate, v, m);
		}
		fn.compress_finalize(*cstate);
		auto &segs = CheckpointerSegments(*cp);
		CHECK(segs.size() == 4);
		idx_t bytes = 0;
		for (auto &s : segs) {
			CHECK(s->IsBitCompressed());
			bytes += s->GetDataSize();
		}
		CHECK(bytes == expect);
		cstate.reset();
		DeleteCheckpointer(cp);
		BindAnalyzeDatabase(nullptr);
	}
	printf("HOST_TESTS_OK\n");
	return 0;
}
