// vtable_bench.cpp -- the drop-in path timed the way DuckDB drives it (needs a B200):
//   load   ColumnData::AppendData in 2048-value vectors (host buffers) + the batched compaction it triggers
//   scan   ColumnData::ScanVector pulls of 2048 values through SuccinctGpuFun's scan_vector / scan_partial slots,
//          T scan threads over disjoint row ranges (DuckDB's parallel table scan hands out row groups the same way);
//          the timed loop only pulls (the operator above the scan is not part of the codec); a second, untimed pass
//          checks every value
//   fetch  ColumnData::FetchRow per row id (the fetch_row slot) and ColumnData::FetchRows (batched)
// usage: vtable_bench <log2 values> <bits> <scan threads> [pending_limit]      -> one JSON line on stdout
#include "../adacom_host.hpp"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <thread>

using namespace adacom_host;
using Clock = std::chrono::steady_clock;

static inline uint64_t splitmix(uint64_t x) {
	x += 0x9E3779B97F4A7C15ull;
	x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
	x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
	return x ^ (x >> 31);
}

static double seconds(Clock::time_point a, Clock::time_point b) {
	return std::chrono::duration<double>(b - a).count();
}

int main(int argc, char **argv) {
	const int logn = argc > 1 ? atoi(argv[1]) : 27;
	const int bits = argc > 2 ? atoi(argv[2]) : 17;
	const int threads = argc > 3 ? atoi(argv[3]) : 1;
	const idx_t pending_limit = argc > 4 ? (idx_t)atoll(argv[4]) : 512;
	const idx_t n = 1ull << logn;
	const uint64_t mask = bits >= 64 ? ~0ull : (1ull << bits) - 1;
	const int64_t base = 1000003;

	DatabaseInstance db;
	ColumnData col(db, PhysicalType::INT64);
	col.pending_limit = pending_limit;
	db.Gpu();  // context creation is not part of the load

	// ---- load ------------------------------------------------------------------------------------------
	// the source vectors are minted before the clock starts: a pool of 1021 distinct vectors used round-robin
	// (value of row r = pool[(r / 2048) % 1021][r % 2048]; 1021 is prime, so segments do not repeat)
	const idx_t kPool = 1021;
	std::vector<int64_t> pool(kPool * STANDARD_VECTOR_SIZE);
	std::vector<unsigned __int128> pool_sum(kPool, 0);
	for (idx_t v = 0; v < kPool; v++) {
		for (idx_t i = 0; i < STANDARD_VECTOR_SIZE; i++) {
			pool[v * STANDARD_VECTOR_SIZE + i] = base + (int64_t)(splitmix(v * STANDARD_VECTOR_SIZE + i) & mask);
			pool_sum[v] += (unsigned __int128)(uint64_t)pool[v * STANDARD_VECTOR_SIZE + i];
		}
	}
	auto value_of = [&](idx_t row) {
		return pool[((row / STANDARD_VECTOR_SIZE) % kPool) * STANDARD_VECTOR_SIZE + row % STANDARD_VECTOR_SIZE];
	};
	unsigned __int128 expect_sum = 0;
	for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
		expect_sum += pool_sum[(off / STANDARD_VECTOR_SIZE) % kPool];
	}
	auto t0 = Clock::now();
	{
		ColumnAppendState astate;
		col.InitializeAppend(astate);
		for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
			UnifiedVectorFormat vdata;
			vdata.data = (data_ptr_t)(pool.data() + ((off / STANDARD_VECTOR_SIZE) % kPool) * STANDARD_VECTOR_SIZE);
			col.AppendData(astate, vdata, STANDARD_VECTOR_SIZE);
		}
		col.CompactPending(true);
		adacom_ctx_synchronize(db.Gpu());
	}
	auto t1 = Clock::now();
	const double load_s = seconds(t0, t1);
	const idx_t data_size = db.column_segment_catalog.GetTotalDataSize();

	// ---- scan (timed: pulls only) --------------------------------------------------------------------
	auto scan_range = [&](idx_t row_begin, idx_t row_end, bool verify, unsigned __int128 *sum_out, idx_t *zero_copy_out) {
		ColumnScanState state;
		col.InitializeScanWithOffset(state, row_begin);
		Vector result(PhysicalType::INT64);
		unsigned __int128 sum = 0;
		idx_t zero_copy = 0;
		for (idx_t row = row_begin; row < row_end;) {
			idx_t got = col.ScanVector(state, result, std::min<idx_t>(STANDARD_VECTOR_SIZE, row_end - row));
			auto out = (const int64_t *)FlatVector::GetData(result);
			zero_copy += result.external != nullptr;
			if (verify) {
				for (idx_t i = 0; i < got; i++) {
					sum += (unsigned __int128)(uint64_t)out[i];
				}
			} else {
				sum += (uint64_t)out[0];  // touch the vector
			}
			row += got;
		}
		*sum_out = sum;
		*zero_copy_out = zero_copy;
	};
	auto run_scan = [&](int T, bool verify, unsigned __int128 *total, idx_t *zero_copy) {
		std::vector<std::thread> pool;
		std::vector<unsigned __int128> sums(T, 0);
		std::vector<idx_t> zc(T, 0);
		const idx_t vectors = n / STANDARD_VECTOR_SIZE;
		auto a = Clock::now();
		for (int t = 0; t < T; t++) {
			idx_t b = vectors * t / T * STANDARD_VECTOR_SIZE, e = vectors * (t + 1) / T * STANDARD_VECTOR_SIZE;
			pool.emplace_back(scan_range, b, e, verify, &sums[t], &zc[t]);
		}
		for (auto &th : pool) {
			th.join();
		}
		auto b = Clock::now();
		*total = 0;
		*zero_copy = 0;
		for (int t = 0; t < T; t++) {
			*total += sums[t];
			*zero_copy += zc[t];
		}
		return seconds(a, b);
	};
	unsigned __int128 sum = 0;
	idx_t zero_copy = 0;
	run_scan(threads, false, &sum, &zero_copy);  // warm-up (pinned rings, first launches)
	double scan_s = 1e30, scan1_s = 1e30;
	for (int rep = 0; rep < 3; rep++) {
		scan_s = std::min(scan_s, run_scan(threads, false, &sum, &zero_copy));
	}
	for (int rep = 0; rep < 2; rep++) {
		scan1_s = std::min(scan1_s, threads == 1 ? scan_s : run_scan(1, false, &sum, &zero_copy));
	}
	const double verify_s = run_scan(threads, true, &sum, &zero_copy);
	const bool ok = sum == expect_sum;

	// ---- fetch_row ---------------------------------------------------------------------------------------
	const idx_t n_single = 20000, n_batch = 1000000;
	std::vector<row_t> ids(n_batch);
	for (idx_t i = 0; i < n_batch; i++) {
		ids[i] = (row_t)(splitmix(0xFE7C000 + i) % n);
	}
	ColumnFetchState fstate;
	Vector one(PhysicalType::INT64);
	bool fetch_ok = true;
	auto f0 = Clock::now();
	for (idx_t i = 0; i < n_single; i++) {
		col.FetchRow(fstate, ids[i], one, 0);
		fetch_ok &= ((int64_t *)FlatVector::GetData(one))[0] == value_of((idx_t)ids[i]);
	}
	auto f1 = Clock::now();
	Vector many(PhysicalType::INT64, n_batch);
	col.FetchRows(ids.data(), n_batch, many);  // warm-up
	auto f2 = Clock::now();
	col.FetchRows(ids.data(), n_batch, many);
	auto f3 = Clock::now();
	for (idx_t i = 0; i < n_batch; i += 997) {
		fetch_ok &= ((int64_t *)FlatVector::GetData(many))[i] == value_of((idx_t)ids[i]);
	}

	// ---- point reads under the access-frequency policy: the three arms of run_diff_zipfian.sh through the vtable ----
	// A look-up = the scan of the 2048-row vector that holds the key (what a zone-map-pruned point filter pulls), keys
	// Zipf-like with q = 1 (row = N^u - 1).  non-adaptive succinct: every segment packed in HBM (the column above);
	// non-succinct: adaptive mode before the first epoch, every segment a flat host block; adaptive: after one epoch
	// of ColumnSegmentCatalog (coldest 90 % packed in HBM by ONE launch set, hottest 10 % flat in host memory).
	const idx_t n_point = 4000;
	std::vector<idx_t> keys(n_point * 8);
	for (idx_t i = 0; i < keys.size(); i++) {
		const double u = (double)(splitmix(0x21BF000 + i) >> 11) / 9007199254740992.0;
		keys[i] = std::min<idx_t>(n - 1, (idx_t)std::pow((double)n, u) - 0);
		keys[i] = keys[i] ? keys[i] - 1 : 0;
	}
	auto point_reads = [&](ColumnData &c, idx_t first, idx_t count, bool *ok_out) {
		Vector result(PhysicalType::INT64);
		bool all_ok = true;
		auto a = Clock::now();
		for (idx_t i = first; i < first + count; i++) {
			const idx_t row = keys[i], vec0 = row / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE;
			ColumnScanState state;
			c.InitializeScanWithOffset(state, vec0);
			c.ScanVector(state, result, std::min<idx_t>(STANDARD_VECTOR_SIZE, n - vec0));
			all_ok &= ((const int64_t *)FlatVector::GetData(result))[row - vec0] == value_of(row);
		}
		auto b = Clock::now();
		*ok_out = all_ok;
		return count / seconds(a, b);
	};
	bool ok_succinct = true, ok_flat = true, ok_adaptive = true;
	// first without the hot-vector cache (every read is a launch + a PCIe round trip), then with it
	const idx_t cache_bytes = col.vector_cache_bytes;
	col.vector_cache_bytes = 0;
	point_reads(col, 0, 200, &ok_succinct);  // warm-up
	const double point_uncached = point_reads(col, 200, n_point, &ok_succinct);
	col.vector_cache_bytes = cache_bytes;
	point_reads(col, 0, 200, &ok_succinct);
	const double point_succinct = point_reads(col, 200, n_point, &ok_succinct);
	const auto cache_stats = col.GetVectorCacheStats();
	double point_flat = 0, point_adaptive = 0, epoch_s = 0;
	idx_t adaptive_bytes = 0, hot_segments = 0;
	{
		DBConfig acfg;
		acfg.adaptive_succinct_compression_enabled = true;
		DatabaseInstance adb(&acfg);
		ColumnData acol(adb, PhysicalType::INT64);
		acol.pending_limit = pending_limit;
		adb.Gpu();
		ColumnAppendState astate;
		acol.InitializeAppend(astate);
		for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
			UnifiedVectorFormat vdata;
			vdata.data = (data_ptr_t)(pool.data() + ((off / STANDARD_VECTOR_SIZE) % kPool) * STANDARD_VECTOR_SIZE);
			acol.AppendData(astate, vdata, STANDARD_VECTOR_SIZE);
		}
		point_reads(acol, 0, 200, &ok_flat);
		point_flat = point_reads(acol, 200, 4 * n_point, &ok_flat);  // also the reads the policy learns from
		auto e0 = Clock::now();
		adb.column_segment_catalog.RunCompactionEpoch();
		adacom_ctx_synchronize(adb.Gpu());
		epoch_s = seconds(e0, Clock::now());
		adaptive_bytes = adb.column_segment_catalog.GetTotalDataSize();
		for (auto &sgm : acol.segments) {
			hot_segments += !sgm->IsBitCompressed();
		}
		point_reads(acol, 0, 200, &ok_adaptive);
		point_adaptive = point_reads(acol, 4 * n_point + 200, 3 * n_point, &ok_adaptive);
	}
	const bool point_ok = ok_succinct && ok_flat && ok_adaptive;

	const double gb = (double)n * 8 / 1e9;
	printf("{\"values\": %llu, \"bits\": %d, \"segments\": %zu, \"device_columns\": %zu, \"data_size\": %llu, "
	       "\"load_s\": %.4f, \"load_GBs\": %.2f, \"scan_threads\": %d, \"scan_s\": %.5f, \"scan_GBs\": %.2f, "
	       "\"scan_1thread_GBs\": %.2f, \"scan_verify_GBs\": %.2f, \"zero_copy_vectors\": %llu, \"vectors\": %llu, "
	       "\"scan_ok\": %s, \"fetch_row_per_s\": %.0f, \"fetch_rows_batched_per_s\": %.0f, \"fetch_ok\": %s, "
	       "\"point_reads_per_s\": {\"non_adaptive_succinct\": %.0f, \"non_adaptive_succinct_no_cache\": %.0f, "
	       "\"vector_cache\": {\"hits\": %llu, \"fills\": %llu, \"first_touches\": %llu}, \"non_succinct\": %.0f, \"adaptive\": %.0f}, "
	       "\"adaptive_epoch_s\": %.4f, \"adaptive_resident_bytes\": %llu, \"adaptive_hot_segments\": %llu, \"point_ok\": %s}\n",
	       (unsigned long long)n, bits, col.segments.size(), col.batches.size(), (unsigned long long)data_size, load_s,
	       gb / load_s, threads, scan_s, gb / scan_s, gb / scan1_s, gb / verify_s, (unsigned long long)zero_copy,
	       (unsigned long long)(n / STANDARD_VECTOR_SIZE), ok ? "true" : "false", n_single / seconds(f0, f1),
	       n_batch / seconds(f2, f3), fetch_ok ? "true" : "false", point_succinct, point_uncached,
	       (unsigned long long)cache_stats.hits, (unsigned long long)cache_stats.fills, (unsigned long long)cache_stats.misses,
	       point_flat, point_adaptive, epoch_s,
	       (unsigned long long)adaptive_bytes, (unsigned long long)hot_segments, point_ok ? "true" : "false");
	return ok && fetch_ok && point_ok ? 0 : 1;
}
