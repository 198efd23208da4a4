// host_resident_test.cpp -- ColumnData over segments that live on the host (adaptive mode before the first policy epoch:
// flat blocks, nothing uploaded), driven like DuckDB drives the reference: AppendData with and without NULLs and with a
// selection vector, InitializeScan / InitializeScanWithOffset / ScanVector across segment borders, FetchRow, FetchRows,
// the catalog's accounting.  No GPU is touched: this is the part of the shim that must behave exactly like the
// reference's uncompressed segments (src/storage/table/column_data.cpp:73-139, 249-315).
#include "../adacom_host.hpp"

#include <cstdio>
#include <cstdlib>
#include <limits>
#include <random>

using namespace adacom_host;

#define CHECK(cond)                                                       \
	do {                                                                  \
		if (!(cond)) {                                                    \
			printf("CHECK failed at line %d: %s\n", __LINE__, #cond);     \
			exit(1);                                                      \
		}                                                                 \
	} while (0)

template <class T>
static void Run(PhysicalType type, idx_t n, bool nulls, bool with_sel) {
	DBConfig cfg;
	cfg.adaptive_succinct_compression_enabled = true;
	DatabaseInstance db(&cfg);
	std::mt19937_64 rng(11 + n);
	std::vector<T> values(n);
	for (auto &v : values) {
		v = (T)(rng() % 200) - (std::numeric_limits<T>::is_signed ? (T)100 : (T)0);
	}
	auto valid = [&](idx_t row) { return !nulls || row % 5 != 2; };
	ColumnData col(db, type);
	col.vector_cache_bytes = 0;  // the cache has its own test
	ColumnAppendState astate;
	col.InitializeAppend(astate);
	std::vector<uint32_t> sel(STANDARD_VECTOR_SIZE);
	std::vector<T> shuffled(STANDARD_VECTOR_SIZE);
	for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
		const idx_t m = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - off);
		std::vector<uint64_t> mask((STANDARD_VECTOR_SIZE + 63) / 64, ~0ull);
		UnifiedVectorFormat vdata;
		SelectionVector sv;
		if (with_sel) {
			// the vector arrives reversed behind a selection vector: row i of the append is source slot m - 1 - i
			for (idx_t i = 0; i < m; i++) {
				sel[i] = (uint32_t)(m - 1 - i);
				shuffled[m - 1 - i] = values[off + i];
				if (!valid(off + i)) {
					mask[(m - 1 - i) >> 6] &= ~(1ull << ((m - 1 - i) & 63));
				}
			}
			sv.sel = sel.data();
			vdata.sel = &sv;
			vdata.data = (data_ptr_t)shuffled.data();
		} else {
			for (idx_t i = 0; i < m; i++) {
				if (!valid(off + i)) {
					mask[i >> 6] &= ~(1ull << (i & 63));
				}
			}
			vdata.data = (data_ptr_t)(values.data() + off);
		}
		vdata.validity.mask = nulls ? mask.data() : nullptr;
		col.AppendData(astate, vdata, m);
	}
	const idx_t cap = BLOCK_SIZE / sizeof(T);
	CHECK(col.GetMaxEntry() == n);
	CHECK(col.segments.size() == (n + cap - 1) / cap);
	for (auto &s : col.segments) {
		CHECK(!s->IsBitCompressed() && s->gpu_column == nullptr);
	}
	CHECK(db.column_segment_catalog.GetTotalDataSize() == col.segments.size() * BLOCK_SIZE);
	CHECK(db.buffer_manager.GetDataSize() == col.segments.size() * BLOCK_SIZE);
	// NULL rows hold NullValue<T>() (succinct.cpp:288-292)
	auto expect = [&](idx_t row) { return valid(row) ? values[row] : std::numeric_limits<T>::min(); };
	// full scan, then scans from the middle of a segment and from the last vector
	Vector result(type);
	for (idx_t from : {(idx_t)0, (cap / 2) / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE, (n - 1) / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE}) {
		if (from >= n) {
			continue;
		}
		ColumnScanState state;
		if (from == 0) {
			col.InitializeScan(state);
		} else {
			col.InitializeScanWithOffset(state, from);
		}
		for (idx_t row = from; row < n;) {
			const idx_t want = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - row);
			CHECK(col.ScanVector(state, result, want) == want);
			auto out = (const T *)FlatVector::GetData(result);
			for (idx_t i = 0; i < want; i++) {
				CHECK(out[i] == expect(row + i));
			}
			row += want;
		}
	}
	// point fetches, one by one and as a batch
	ColumnFetchState fstate;
	std::vector<row_t> ids;
	for (int q = 0; q < 200; q++) {
		const idx_t row = rng() % n;
		ids.push_back((row_t)row);
		col.FetchRow(fstate, (row_t)row, result, (idx_t)(q % 7));
		CHECK(((const T *)FlatVector::GetData(result))[q % 7] == expect(row));
	}
	Vector many(type);
	col.FetchRows(ids.data(), ids.size(), many);
	for (idx_t k = 0; k < ids.size(); k++) {
		CHECK(((const T *)FlatVector::GetData(many))[k] == expect((idx_t)ids[k]));
	}
}

int main() {
	Run<int64_t>(PhysicalType::INT64, 32767 * 2 + 1234, false, false);
	Run<int32_t>(PhysicalType::INT32, 65534 + 4000, true, false);
	Run<uint16_t>(PhysicalType::UINT16, 131068 + 77, true, true);
	Run<int8_t>(PhysicalType::INT8, 262136 + 2049, false, true);
	Run<uint8_t>(PhysicalType::UINT8, 5000, true, false);
	printf("HOST_RESIDENT_TESTS_OK\n");
	return 0;
}
