// policy_test.cpp -- CPU-only test of the ColumnSegmentCatalog decision rule with fake segments.
// (no GPU, no codec: only the 90/10 split, counter reset, accounting and thread start/stop)
#include "../adacom_host.hpp"

#include <cstdio>
#include <cstdlib>

using namespace adacom_host;

struct FakeSegment : public CompactableSegment {
	int compacts = 0, uncompacts = 0;
	bool compacted = false;
	bool data = true;
	void Compact() override {
		compacts++;
		compacted = true;
	}
	void Uncompact() override {
		uncompacts++;
		compacted = false;
	}
	idx_t GetDataSize() const override {
		return compacted ? 100 : 1000;
	}
	bool IsDataSegment() const override {
		return data;
	}
};

#define CHECK(cond)                                                      \
	do {                                                                 \
		if (!(cond)) {                                                   \
			fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
			exit(1);                                                     \
		}                                                                \
	} while (0)

int main() {
	ColumnSegmentCatalog catalog(0.90, 0.05);
	std::vector<FakeSegment> segs(20);
	FakeSegment validity;
	validity.data = false;
	catalog.AddColumnSegment(&validity);  // non-data segments are ignored
	for (auto &s : segs) {
		catalog.AddColumnSegment(&s);
	}
	CHECK(catalog.GetTotalDataSize() == 20 * 1000);
	// segment i is read i times: 18 and 19 are the hottest
	for (size_t i = 0; i < segs.size(); i++) {
		for (size_t r = 0; r < i; r++) {
			catalog.AddReadAccess(&segs[i]);
		}
	}
	catalog.AddReadAccess(&validity);
	catalog.AddReadAccess(nullptr);
	CHECK(catalog.NumReads(&segs[7]) == 7);
	auto res = catalog.RunCompactionEpoch();
	// the reference compares float(cum)/n with the double 0.90: 18/20 = 0.9f = 0.89999998 < 0.90 holds, so the
	// 18 coldest are compacted and 2 stay flat (column_segment_catalog.cpp:88-105, same arithmetic here)
	CHECK(res.first == 18 && res.second == 2);
	for (size_t i = 0; i < 18; i++) {
		CHECK(segs[i].compacts == 1 && segs[i].uncompacts == 0);
	}
	for (size_t i = 18; i < 20; i++) {
		CHECK(segs[i].compacts == 0 && segs[i].uncompacts == 1);
	}
	CHECK(catalog.NumReads(&segs[19]) == 0);  // counters reset every epoch
	CHECK(catalog.GetTotalDataSize() == 18 * 100 + 2 * 1000);
	CHECK(validity.compacts == 0);
	// CompactAllSegments (what the benchmarks call) compacts everything
	catalog.CompactAllSegments();
	CHECK(catalog.GetTotalDataSize() == 20 * 100);
	// background thread: runs epochs on its own and stops cleanly
	for (int r = 0; r < 5; r++) {
		catalog.AddReadAccess(&segs[3]);
	}
	catalog.EnableBackgroundThreadCompaction();
	CHECK(catalog.BackgroundCompactionEnabled());
	std::this_thread::sleep_for(std::chrono::milliseconds(300));
	catalog.StopBackgroundThread();
	CHECK(segs[3].uncompacts >= 1);  // the only segment with reads is the hottest -> uncompacted
	catalog.RemoveColumnSegment(&segs[0]);
	CHECK(catalog.NumReads(&segs[0]) == 0);
	printf("POLICY_TESTS_OK\n");
	return 0;
}
