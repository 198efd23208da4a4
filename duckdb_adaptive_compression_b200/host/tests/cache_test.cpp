// cache_test.cpp -- the hot-vector cache of ColumnData (point reads: a scan WithOffset + one pull, FetchRow) without a
// GPU: in adaptive mode segments start as flat host blocks and are never uploaded unless the policy compacts them, so
// every read below is served from host memory and only the host logic is exercised -- second-touch fill, hits, the scan
// that continues behind a cached vector, vectors that cross a segment border, the tail vector that grows, and the
// counters that tell point reads from the row-group starts of a table scan.
#include "../adacom_host.hpp"

#include <cstdio>
#include <cstdlib>
#include <random>

using namespace adacom_host;

#define CHECK(cond)                                                       \
	do {                                                                  \
		if (!(cond)) {                                                    \
			printf("CHECK failed at line %d: %s\n", __LINE__, #cond);     \
			exit(1);                                                      \
		}                                                                 \
	} while (0)

int main() {
	DBConfig cfg;
	cfg.adaptive_succinct_compression_enabled = true;  // segments stay flat on the host until the policy says otherwise
	DatabaseInstance db(&cfg);
	const idx_t cap = BLOCK_SIZE / sizeof(int32_t);
	idx_t n = cap * 3 + 5000;  // vectors cross the segment borders (cap is not a multiple of 2048)
	std::vector<int32_t> values(n);
	std::mt19937 rng(3);
	for (auto &v : values) {
		v = (int32_t)(rng() % 100000) - 500;
	}
	ColumnData col(db, PhysicalType::INT32);
	ColumnAppendState astate;
	col.InitializeAppend(astate);
	for (idx_t off = 0; off < n; off += STANDARD_VECTOR_SIZE) {
		UnifiedVectorFormat vdata;
		vdata.data = (data_ptr_t)(values.data() + off);
		col.AppendData(astate, vdata, std::min<idx_t>(STANDARD_VECTOR_SIZE, n - off));
	}
	CHECK(col.GetMaxEntry() == n && col.segments.size() == 4);
	Vector result(PhysicalType::INT32);
	auto read_vector = [&](idx_t p, int pulls) {
		ColumnScanState state;
		col.InitializeScanWithOffset(state, p);
		for (int k = 0; k < pulls && p < n; k++) {
			const idx_t want = std::min<idx_t>(STANDARD_VECTOR_SIZE, n - p);
			CHECK(col.ScanVector(state, result, want) == want);
			auto out = (const int32_t *)FlatVector::GetData(result);
			for (idx_t i = 0; i < want; i++) {
				CHECK(out[i] == values[p + i]);
			}
			p += want;
		}
	};
	const idx_t border_vec = cap / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE;  // the vector that holds a segment border
	const idx_t probes[] = {0, 4096, border_vec, 2 * cap / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE, (n - 1) / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE};
	auto s0 = col.GetVectorCacheStats();
	CHECK(s0.hits == 0 && s0.fills == 0);
	for (idx_t p : probes) {
		read_vector(p, 1);  // first touch: the ordinary path
	}
	auto s1 = col.GetVectorCacheStats();
	CHECK(s1.fills == 0 && s1.hits == 0 && s1.misses == 5);
	for (idx_t p : probes) {
		read_vector(p, 1);  // second touch: the vector is decoded into the cache
	}
	auto s2 = col.GetVectorCacheStats();
	CHECK(s2.fills == 5 && s2.hits == 0);
	for (int rep = 0; rep < 3; rep++) {
		for (idx_t p : probes) {
			read_vector(p, 2);  // hit, and the scan continues correctly behind the cached vector (also across a border)
		}
	}
	auto s3 = col.GetVectorCacheStats();
	CHECK(s3.hits == 15 && s3.fills == 5);
	// FetchRow goes through the same entries; rows of untouched vectors take the ordinary path on their first touch
	ColumnFetchState fstate;
	for (idx_t p : probes) {
		for (idx_t d : {(idx_t)0, (idx_t)17, (idx_t)2047}) {
			if (p + d < n) {
				col.FetchRow(fstate, (row_t)(p + d), result, 2);
				CHECK(((const int32_t *)FlatVector::GetData(result))[2] == values[p + d]);
			}
		}
	}
	auto s4 = col.GetVectorCacheStats();
	CHECK(s4.hits > s3.hits && s4.fills == 5);
	col.FetchRow(fstate, (row_t)(10 * 2048 + 3), result, 0);
	CHECK(((const int32_t *)FlatVector::GetData(result))[0] == values[10 * 2048 + 3]);
	CHECK(col.GetVectorCacheStats().misses == s4.misses + 1);
	// the tail vector grows: the entry held fewer rows than the read needs and is refreshed
	const idx_t n_old = n;
	std::vector<int32_t> more(900);
	for (auto &v : more) {
		v = (int32_t)(rng() % 1000);
	}
	UnifiedVectorFormat vdata;
	vdata.data = (data_ptr_t)more.data();
	col.AppendData(astate, vdata, more.size());
	values.insert(values.end(), more.begin(), more.end());
	n += more.size();
	for (int rep = 0; rep < 3; rep++) {
		col.FetchRow(fstate, (row_t)(n - 1), result, 0);
		CHECK(((const int32_t *)FlatVector::GetData(result))[0] == values[n - 1]);
		col.FetchRow(fstate, (row_t)(n_old + 1), result, 1);
		CHECK(((const int32_t *)FlatVector::GetData(result))[1] == values[n_old + 1]);
	}
	read_vector((n - 1) / STANDARD_VECTOR_SIZE * STANDARD_VECTOR_SIZE, 1);
	// a disabled cache is never consulted
	col.vector_cache_bytes = 0;
	auto s5 = col.GetVectorCacheStats();
	read_vector(0, 2);
	col.FetchRow(fstate, 5, result, 0);
	auto s6 = col.GetVectorCacheStats();
	CHECK(s6.hits == s5.hits && s6.misses == s5.misses && s6.fills == s5.fills);
	printf("CACHE_TESTS_OK hits=%llu fills=%llu first_touches=%llu\n", (unsigned long long)s6.hits, (unsigned long long)s6.fills,
	       (unsigned long long)s6.misses);
	return 0;
}
