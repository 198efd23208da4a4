// race_test.cpp -- the host shim under ThreadSanitizer, no GPU: reader threads (point reads through the hot-vector cache,
// FetchRow, full scans with zero-copy vectors) against a writer that keeps appending to the same ColumnData and a thread
// that counts read accesses in the catalog, all over host-resident segments (adaptive mode, nothing uploaded).  Built by
// tests/test_host_shim.py with -fsanitize=thread over the shim's own sources; a data race fails the test.
// (The reference ships a suppression file for its own races on this path, .sanitizer-thread-suppressions.txt.)
#include "../adacom_host.hpp"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <thread>

using namespace adacom_host;

int main() {
	DBConfig cfg;
	cfg.adaptive_succinct_compression_enabled = true;
	DatabaseInstance db(&cfg);
	const idx_t n0 = 32767 * 3 + 999, n_max = n0 + 1200 * 1000;  // the writer opens ~36 new segments
	std::vector<int64_t> values(n_max);
	for (idx_t i = 0; i < n_max; i++) {
		values[i] = (int64_t)(i * 2654435761u % 1000003) - 7;
	}
	ColumnData col(db, PhysicalType::INT64);
	ColumnAppendState astate;
	col.InitializeAppend(astate);
	for (idx_t off = 0; off < n0; off += STANDARD_VECTOR_SIZE) {
		UnifiedVectorFormat vdata;
		vdata.data = (data_ptr_t)(values.data() + off);
		col.AppendData(astate, vdata, std::min<idx_t>(STANDARD_VECTOR_SIZE, n0 - off));
	}
	std::atomic<idx_t> published {n0};  // rows the writer has finished appending
	std::atomic<bool> stop {false};
	std::atomic<int> errors {0};
	std::thread writer([&] {
		idx_t n = n0;
		while (n + 1000 <= n_max) {
			UnifiedVectorFormat vdata;
			vdata.data = (data_ptr_t)(values.data() + n);
			col.AppendData(astate, vdata, 1000);
			n += 1000;
			published.store(n, std::memory_order_release);
			std::this_thread::yield();
		}
	});
	std::vector<std::thread> readers;
	for (int t = 0; t < 3; t++) {
		readers.emplace_back([&, t] {
			std::mt19937_64 rng(100 + t);
			Vector result(PhysicalType::INT64);
			ColumnFetchState fstate;
			while (!stop.load(std::memory_order_acquire)) {
				const idx_t n = n0;  // readers stay inside the sealed prefix; the tail is the writer's
				// skewed point reads: a few hot vectors come back again and again (fills and hits race with each other)
				const idx_t vec = (rng() % 4 == 0 ? rng() % (n / STANDARD_VECTOR_SIZE) : rng() % 6) * STANDARD_VECTOR_SIZE;
				ColumnScanState state;
				col.InitializeScanWithOffset(state, vec);
				const idx_t want = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - vec);
				if (col.ScanVector(state, result, want) != want) {
					errors++;
				}
				auto out = (const int64_t *)FlatVector::GetData(result);
				for (idx_t i = 0; i < want; i += 97) {
					errors += out[i] != values[vec + i];
				}
				const idx_t row = rng() % n;
				col.FetchRow(fstate, (row_t)row, result, 1);
				errors += ((const int64_t *)FlatVector::GetData(result))[1] != values[row];
				db.column_segment_catalog.AddReadAccess(col.GetSegment(row));
				if ((rng() & 7) == 0) {  // a batch of row ids (snapshot of the segment vector while the writer grows it)
					row_t ids[16];
					for (auto &id : ids) {
						id = (row_t)(rng() % n);
					}
					Vector many(PhysicalType::INT64);
					col.FetchRows(ids, 16, many);
					for (int k = 0; k < 16; k++) {
						errors += ((const int64_t *)FlatVector::GetData(many))[k] != values[ids[k]];
					}
				}
			}
		});
	}
	// a full scan (zero-copy vectors out of the host blocks) over the sealed prefix, again and again
	readers.emplace_back([&] {
		Vector result(PhysicalType::INT64);
		while (!stop.load(std::memory_order_acquire)) {
			ColumnScanState state;
			col.InitializeScan(state);
			for (idx_t row = 0; row < n0;) {
				const idx_t want = std::min<idx_t>(STANDARD_VECTOR_SIZE, n0 - row);
				if (col.ScanVector(state, result, want) != want) {
					errors++;
					break;
				}
				auto out = (const int64_t *)FlatVector::GetData(result);
				errors += out[0] != values[row] || out[want - 1] != values[row + want - 1];
				row += want;
			}
		}
	});
	writer.join();
	stop.store(true, std::memory_order_release);
	for (auto &r : readers) {
		r.join();
	}
	const idx_t n = published.load();
	if (errors.load() || col.GetMaxEntry() != n) {
		printf("RACE_TEST_FAILED errors=%d\n", errors.load());
		return 1;
	}
	auto st = col.GetVectorCacheStats();
	printf("RACE_TEST_OK rows=%llu hits=%llu fills=%llu\n", (unsigned long long)n, (unsigned long long)st.hits, (unsigned long long)st.fills);
	return 0;
}
