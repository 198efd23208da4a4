// column_segment_catalog.cpp -- the access-frequency policy that decides which segments get compacted.
//
// Same public surface and the same decision rule as the reference
// (src/catalog/catalog_entry/column_segment_catalog.cpp:13-135): per-segment read counters, and every
// `interval_s` seconds: sort by reads, Compact() the coldest `compression_rate` share, Uncompact() the rest,
// zero the counters.  Differences, all on the quirk list: the map is guarded by a mutex, the
// worker thread is joinable and stoppable, and one loop iteration is callable on its own (RunCompactionEpoch).
#include "adacom_host.hpp"

#include <algorithm>
#include <chrono>

namespace adacom_host {

ColumnSegmentCatalog::ColumnSegmentCatalog(double compression_rate_p, double interval_s_p)
    : compression_rate(compression_rate_p), interval_s(interval_s_p) {
}

ColumnSegmentCatalog::~ColumnSegmentCatalog() {
	StopBackgroundThread();
}

void ColumnSegmentCatalog::EnableBackgroundThreadCompaction() {
	std::lock_guard<std::mutex> guard(lock);
	if (!background_compaction_enabled) {
		background_compaction_enabled = true;
		stop = false;
		worker = std::thread(&ColumnSegmentCatalog::CompressLowestKSegments, this);
	}
}

void ColumnSegmentCatalog::StopBackgroundThread() {
	{
		std::lock_guard<std::mutex> guard(lock);
		stop = true;
	}
	cv.notify_all();
	if (worker.joinable()) {
		worker.join();
	}
	background_compaction_enabled = false;
}

void ColumnSegmentCatalog::AddColumnSegment(CompactableSegment *segment) {
	if (segment && segment->IsDataSegment()) {
		std::lock_guard<std::mutex> guard(lock);
		statistics[segment] = AccessStatistics {/* num_reads= */ 0};
		segment->catalog_reads.store(0, std::memory_order_relaxed);
		segment->catalog_registered.store(true, std::memory_order_release);
	}
}

void ColumnSegmentCatalog::RemoveColumnSegment(CompactableSegment *segment) {
	std::lock_guard<std::mutex> guard(lock);
	statistics.erase(segment);
	segment->catalog_registered.store(false, std::memory_order_release);
}

void ColumnSegmentCatalog::AddReadAccess(CompactableSegment *segment) {
	if (segment == nullptr || !segment->IsDataSegment()) {
		return;
	}
	if (segment->catalog_registered.load(std::memory_order_acquire)) {
		segment->catalog_reads.fetch_add(1, std::memory_order_relaxed);
		event_counter.fetch_add(1, std::memory_order_relaxed);
		return;
	}
	std::lock_guard<std::mutex> guard(lock);
	if (statistics.find(segment) == statistics.end()) {
		statistics[segment] = AccessStatistics {/* num_reads= */ 0};
		segment->catalog_reads.store(1, std::memory_order_relaxed);  // column_segment_catalog.cpp:44-46: first access
		segment->catalog_registered.store(true, std::memory_order_release);
	} else {
		segment->catalog_reads.fetch_add(1, std::memory_order_relaxed);
		event_counter.fetch_add(1, std::memory_order_relaxed);
	}
}

idx_t ColumnSegmentCatalog::NumReads(CompactableSegment *segment) {
	std::lock_guard<std::mutex> guard(lock);
	auto it = statistics.find(segment);
	return it == statistics.end() ? 0 : segment->catalog_reads.load(std::memory_order_relaxed);
}

void ColumnSegmentCatalog::CompactAllSegments() {
	std::vector<CompactableSegment *> all;
	{
		std::lock_guard<std::mutex> guard(lock);
		for (auto &kv : statistics) {
			all.push_back(kv.first);
		}
	}
	// segments of a ColumnData are compacted by their owner in one launch set; the others one by one
	std::unordered_map<SegmentBatchOwner *, std::vector<CompactableSegment *>> batches;
	for (auto *seg : all) {
		if (auto *owner = seg->BatchOwner()) {
			batches[owner].push_back(seg);
		} else {
			seg->Compact();
		}
	}
	for (auto &kv : batches) {
		kv.first->CompactMany(kv.second);
	}
}

std::pair<size_t, size_t> ColumnSegmentCatalog::RunCompactionEpoch() {
	std::vector<std::pair<CompactableSegment *, AccessStatistics>> v;
	{
		std::lock_guard<std::mutex> guard(lock);
		v.reserve(statistics.size());
		for (auto &kv : statistics) {
			v.emplace_back(kv.first, AccessStatistics {kv.first->catalog_reads.load(std::memory_order_relaxed)});
		}
	}
	// stable order for equal read counts (the reference's unordered_map order is unspecified)
	std::stable_sort(v.begin(), v.end(),
	                 [](const std::pair<CompactableSegment *, AccessStatistics> &l,
	                    const std::pair<CompactableSegment *, AccessStatistics> &r) {
		                 return l.second.num_reads < r.second.num_reads;
	                 });
	size_t compacted = 0, uncompacted = 0;
	float cum_sum = 0;
	const size_t curr_counter = v.size();
	// same decisions as the reference's loop; segments that share a device column (BatchOwner) are handed over as
	// two lists per owner -- one launch set each -- instead of one Compact()/Uncompact() call per segment
	std::unordered_map<SegmentBatchOwner *, std::pair<std::vector<CompactableSegment *>, std::vector<CompactableSegment *>>>
	    batches;
	for (auto &entry : v) {
		cum_sum += 1;
		SegmentBatchOwner *owner = entry.first->BatchOwner();
		if (cum_sum / curr_counter < compression_rate) {
			if (owner) {
				batches[owner].first.push_back(entry.first);
			} else {
				entry.first->Compact();  // least accessed share
			}
			compacted++;
		} else {
			if (owner) {
				batches[owner].second.push_back(entry.first);
			} else {
				entry.first->Uncompact();  // frequently accessed: leave / make flat
			}
			uncompacted++;
		}
	}
	for (auto &kv : batches) {
		if (!kv.second.first.empty()) {
			kv.first->CompactMany(kv.second.first);
		}
		if (!kv.second.second.empty()) {
			kv.first->UncompactMany(kv.second.second);
		}
	}
	{
		std::lock_guard<std::mutex> guard(lock);
		for (auto &entry : v) {
			auto it = statistics.find(entry.first);
			if (it != statistics.end()) {
				entry.first->catalog_reads.store(0, std::memory_order_relaxed);
			}
		}
		event_counter = 0;
	}
	return {compacted, uncompacted};
}

void ColumnSegmentCatalog::CompressLowestKSegments() {
	std::unique_lock<std::mutex> guard(lock);
	while (!stop) {
		cv.wait_for(guard, std::chrono::duration<double>(interval_s), [this] { return stop; });
		if (stop) {
			break;
		}
		guard.unlock();
		RunCompactionEpoch();
		guard.lock();
	}
}

size_t ColumnSegmentCatalog::GetTotalDataSize() {
	std::lock_guard<std::mutex> guard(lock);
	size_t data_size = 0;
	for (auto &kv : statistics) {
		data_size += kv.first->GetDataSize();
	}
	return data_size;
}

std::unordered_map<int, size_t> ColumnSegmentCatalog::GetDataSizePerDevice() {
	std::lock_guard<std::mutex> guard(lock);
	std::unordered_map<int, size_t> sizes;
	for (auto &kv : statistics) {
		sizes[kv.first->Device()] += kv.first->GetDataSize();
	}
	return sizes;
}

}  // namespace adacom_host
