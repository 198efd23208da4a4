// multi_gpu_test.cpp -- one process, G GPUs (needs >= 1 B200; run with G = all visible devices, or argv[1]):
// the row groups of a table are dealt over the devices (row group g -> device gpu_device + g mod G, DESIGN.md
// section 7), every ColumnData lives entirely on its device, scans run with one host thread per device, the partial
// aggregates are combined on the host (exact 128-bit), a fetch is routed key -> row group -> owner device, and the
// catalog reports the resident bytes per device.  No data-path exchange between GPUs: segments never span row groups.
#include "../adacom_host.hpp"

#include <cstdio>
#include <cstdlib>
#include <thread>

using namespace adacom_host;

#define CHECK(cond)                                                           \
	do {                                                                      \
		if (!(cond)) {                                                        \
			fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
			exit(1);                                                          \
		}                                                                     \
	} while (0)

static inline uint64_t splitmix(uint64_t x) {
	x += 0x9E3779B97F4A7C15ull;
	x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
	x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
	return x ^ (x >> 31);
}

int main(int argc, char **argv) {
	const idx_t ROW_GROUP = 122880;  // storage_info.hpp:19
	const idx_t n_groups = 24;
	int G = argc > 1 ? atoi(argv[1]) : 0;
	if (G <= 0) {
		// probe: contexts 0, 1, ... until one fails
		DBConfig probe_cfg;
		DatabaseInstance probe(&probe_cfg);
		for (G = 0; G < 8; G++) {
			try {
				probe.Gpu(G);
			} catch (const InternalException &) {
				break;
			}
		}
	}
	CHECK(G >= 1);
	DBConfig cfg;
	cfg.gpu_devices = G;
	DatabaseInstance db(&cfg);
	std::vector<unique_ptr<ColumnData>> groups;
	std::vector<int64_t> vec(STANDARD_VECTOR_SIZE);
	unsigned __int128 expect = 0;
	for (idx_t g = 0; g < n_groups; g++) {
		const int device = db.DeviceForRowGroup(g);
		CHECK(device == cfg.gpu_device + (int)(g % (idx_t)G));
		groups.emplace_back(new ColumnData(db, PhysicalType::INT64, g * ROW_GROUP, device));
		ColumnAppendState astate;
		groups.back()->InitializeAppend(astate);
		for (idx_t off = 0; off < ROW_GROUP; off += STANDARD_VECTOR_SIZE) {
			for (idx_t i = 0; i < STANDARD_VECTOR_SIZE; i++) {
				vec[i] = -77 + (int64_t)(splitmix(g * ROW_GROUP + off + i) & 0xFFFFF);
				expect += (unsigned __int128)(uint64_t)vec[i];
			}
			UnifiedVectorFormat vdata;
			vdata.data = (data_ptr_t)vec.data();
			groups.back()->AppendData(astate, vdata, STANDARD_VECTOR_SIZE);
		}
	}
	db.column_segment_catalog.CompactAllSegments();
	// resident bytes per device: every device holds its share, the sum is the catalog's total
	auto per_device = db.column_segment_catalog.GetDataSizePerDevice();
	CHECK((int)per_device.size() == G);
	size_t total = 0;
	for (auto &kv : per_device) {
		CHECK(kv.second > 0);
		total += kv.second;
	}
	CHECK(total == db.column_segment_catalog.GetTotalDataSize());
	CHECK(total == db.buffer_manager.GetDataSize());
	// scan: one host thread per device over that device's row groups, partial sums combined on the host
	std::vector<unsigned __int128> partial(G, 0);
	std::vector<std::thread> pool;
	for (int d = 0; d < G; d++) {
		pool.emplace_back([&, d]() {
			Vector result(PhysicalType::INT64);
			for (idx_t g = (idx_t)d; g < n_groups; g += (idx_t)G) {
				ColumnScanState state;
				groups[g]->InitializeScan(state);
				for (idx_t row = 0; row < ROW_GROUP; row += STANDARD_VECTOR_SIZE) {
					idx_t got = groups[g]->ScanVector(state, result, STANDARD_VECTOR_SIZE);
					auto out = (const int64_t *)FlatVector::GetData(result);
					for (idx_t i = 0; i < got; i++) {
						partial[d] += (unsigned __int128)(uint64_t)out[i];
					}
				}
			}
		});
	}
	for (auto &t : pool) {
		t.join();
	}
	unsigned __int128 sum = 0;
	for (auto p : partial) {
		sum += p;
	}
	CHECK(sum == expect);
	// fetch: key -> row group -> owner device
	ColumnFetchState fstate;
	Vector one(PhysicalType::INT64);
	for (int q = 0; q < 500; q++) {
		idx_t row = splitmix(1000 + q) % (n_groups * ROW_GROUP);
		idx_t g = row / ROW_GROUP;
		CHECK(groups[g]->device == db.DeviceForRowGroup(g));
		groups[g]->FetchRow(fstate, (row_t)row, one, 0);
		CHECK(((int64_t *)FlatVector::GetData(one))[0] == -77 + (int64_t)(splitmix(row) & 0xFFFFF));
	}
	// a policy epoch spans the devices: one launch set per (column, decision) on each column's own device
	for (int rep = 0; rep < 30; rep++) {
		db.column_segment_catalog.AddReadAccess(groups[3]->segments[0].get());
	}
	auto res = db.column_segment_catalog.RunCompactionEpoch();
	CHECK(res.first + res.second == n_groups * groups[0]->segments.size());
	CHECK(!groups[3]->segments[0]->IsBitCompressed());
	printf("MULTI_GPU_TEST_OK devices=%d row_groups=%llu bytes_per_device=", G, (unsigned long long)n_groups);
	for (int d = 0; d < G; d++) {
		printf("%s%zu", d ? "," : "", per_device[cfg.gpu_device + d]);
	}
	printf("\n");
	return 0;
}
