// ref_harness.cpp -- L0 oracle: drives the REFERENCE'S OWN vendored SDSL (sdsl::int_vector<>,
// sdsl::bits::*, sdsl::size_in_bytes) in the order the reference's codec loops call it.
//
// TEST INFRASTRUCTURE ONLY (see succinct_oracle.h).  This file contains no reference source; it is
// compiled together with the SDSL sources where they lie under /root/reference/third_party by
// oracle/Makefile, and the result goes to oracle/_ref/libadacom_ref.so (git-ignored, travels to
// the GPU box).  It exists to pin the L1 restatement (succinct_oracle.c) and to mint the golden
// digests in tests/golden/.
//
// Loops mirrored (paths relative to /root/reference):
//   ColumnSegment ctor ............... src/storage/table/column_segment.cpp:101-105
//   SuccinctAppend / AppendLoop ...... src/storage/compression/succinct.cpp:271-322
//   BitCompressFromSuccinct .......... src/storage/table/column_segment.cpp:348-383
//   Compact (uncompressed branch) .... src/storage/table/column_segment.cpp:304-316
//   BitCompressFromUncompressed ...... src/storage/table/column_segment.cpp:385-456
//   SuccinctScanPartial .............. src/storage/compression/succinct.cpp:123-144
//   UncompressSuccinct ............... src/storage/table/column_segment.cpp:458-506
//   GetDataSize ...................... src/storage/table/column_segment.cpp:204-214
#include <sdsl/vectors.hpp>

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <limits>
#include <sstream>
#include <thread>
#include <vector>

namespace {

enum { FLAG_PADDED = 1, FLAG_LITERAL = 2 };
enum { FUNC_UNCOMPRESSED = 1, FUNC_SUCCINCT = 10 };

struct RefSegment {
	sdsl::int_vector<> succinct_vec;
	std::vector<uint8_t> block;
	uint64_t segment_size = 0;
	uint64_t count = 0;
	uint64_t num_elements = 0;
	uint64_t min_factor = UINT64_MAX;
	uint64_t max_factor = 0;
	uint8_t type_size = 8;
	bool is_signed = false;
	bool compacted = false;
	bool repacked = false;
	int function = FUNC_SUCCINCT;
};

template <class T>
uint64_t AppendLoop(RefSegment &seg, const T *sdata, const uint32_t *sel, const uint64_t *validity, uint64_t offset,
                    uint64_t copy_count) {
	uint64_t umin = UINT64_MAX, umax = 0;
	auto &target = seg.succinct_vec;
	for (uint64_t i = 0; i < copy_count; i++) {
		uint64_t source_idx = sel ? sel[offset + i] : offset + i;
		uint64_t target_idx = seg.count + i;
		bool valid = !validity || ((validity[source_idx >> 6] >> (source_idx & 63)) & 1);
		if (valid) {
			target[target_idx] = sdata[source_idx];
			umin = std::min(umin, uint64_t(sdata[source_idx]));
			umax = std::max(umax, uint64_t(sdata[source_idx]));
		} else {
			target[target_idx] = std::numeric_limits<T>::min();
		}
	}
	seg.min_factor = std::min(seg.min_factor, umin);
	seg.max_factor = std::max(seg.max_factor, umax);
	return copy_count;
}

uint8_t PickWidth(uint64_t range, int flags) {
	uint8_t min_width = sdsl::bits::hi(range) + 1;
	if (flags & FLAG_PADDED) {
		min_width = ((min_width + 7) & (-8));
	}
	return min_width;
}

bool AddsMin(const RefSegment &seg, int flags) {
	if (seg.min_factor == UINT64_MAX) {
		return false;
	}
	return (flags & FLAG_LITERAL) ? true : seg.repacked;
}

} // namespace

extern "C" {

uint32_t ref_bits_hi(uint64_t x) {
	return sdsl::bits::hi(x);
}

void *ref_segment_new_succinct(uint8_t type_size, int is_signed, uint64_t segment_size) {
	auto seg = new RefSegment();
	seg->type_size = type_size;
	seg->is_signed = is_signed != 0;
	seg->segment_size = segment_size;
	seg->function = FUNC_SUCCINCT;
	seg->succinct_vec.width(type_size * 8);
	seg->succinct_vec.resize(segment_size / type_size);
	// realloc'd memory is undefined in the reference; zero it so runs are deterministic
	sdsl::util::set_to_value(seg->succinct_vec, 0);
	return seg;
}

void *ref_segment_new_uncompressed(uint8_t type_size, int is_signed, uint64_t segment_size) {
	auto seg = new RefSegment();
	seg->type_size = type_size;
	seg->is_signed = is_signed != 0;
	seg->segment_size = segment_size;
	seg->function = FUNC_UNCOMPRESSED;
	seg->block.assign(segment_size ? segment_size : 1, 0);
	return seg;
}

void ref_segment_delete(void *h) {
	delete (RefSegment *)h;
}

uint64_t ref_segment_append(void *h, const void *values, const uint32_t *sel, const uint64_t *validity,
                            uint64_t offset, uint64_t count) {
	auto &seg = *(RefSegment *)h;
	uint64_t max_tuple_count = seg.segment_size / seg.type_size;
	uint64_t copy_count = std::min<uint64_t>(count, max_tuple_count - seg.count);
	switch (seg.type_size * 2 + (seg.is_signed ? 1 : 0)) {
	case 2: AppendLoop<uint8_t>(seg, (const uint8_t *)values, sel, validity, offset, copy_count); break;
	case 3: AppendLoop<int8_t>(seg, (const int8_t *)values, sel, validity, offset, copy_count); break;
	case 4: AppendLoop<uint16_t>(seg, (const uint16_t *)values, sel, validity, offset, copy_count); break;
	case 5: AppendLoop<int16_t>(seg, (const int16_t *)values, sel, validity, offset, copy_count); break;
	case 8: AppendLoop<uint32_t>(seg, (const uint32_t *)values, sel, validity, offset, copy_count); break;
	case 9: AppendLoop<int32_t>(seg, (const int32_t *)values, sel, validity, offset, copy_count); break;
	case 16: AppendLoop<uint64_t>(seg, (const uint64_t *)values, sel, validity, offset, copy_count); break;
	default: AppendLoop<int64_t>(seg, (const int64_t *)values, sel, validity, offset, copy_count); break;
	}
	seg.count += copy_count;
	seg.num_elements += count;
	return copy_count;
}

uint64_t ref_segment_append_uncompressed(void *h, const void *values, uint64_t count) {
	auto &seg = *(RefSegment *)h;
	uint64_t max_tuple_count = seg.segment_size / seg.type_size;
	uint64_t copy_count = std::min<uint64_t>(count, max_tuple_count - seg.count);
	memcpy(seg.block.data() + seg.count * seg.type_size, values, copy_count * seg.type_size);
	seg.count += copy_count;
	seg.num_elements += count;
	return copy_count;
}

void ref_bit_compress_from_succinct(void *h, int flags) {
	auto &seg = *(RefSegment *)h;
	uint64_t min = seg.min_factor;
	uint64_t max = seg.max_factor - seg.min_factor;
	uint8_t min_width = PickWidth(max, flags);
	uint8_t old_width = seg.succinct_vec.width();
	if (old_width > min_width) {
		const uint64_t *read_data = seg.succinct_vec.data();
		uint64_t *write_data = seg.succinct_vec.data();
		uint8_t read_offset = 0, write_offset = 0;
		for (uint64_t i = 0; i < seg.count; ++i) {
			uint64_t x = sdsl::bits::read_int_and_move(read_data, read_offset, old_width);
			if (min != UINT64_MAX) {
				x -= min;
			}
			sdsl::bits::write_int_and_move(write_data, x, write_offset, min_width);
		}
		seg.succinct_vec.bit_resize(seg.count * min_width);
		seg.succinct_vec.width(min_width);
		seg.repacked = true;
	}
	seg.compacted = true;
}

void ref_bit_compress_from_uncompressed(void *h, int flags) {
	auto &seg = *(RefSegment *)h;
	seg.succinct_vec.width(seg.type_size * 8);
	seg.succinct_vec.resize(seg.segment_size / seg.type_size);
	sdsl::util::set_to_value(seg.succinct_vec, 0);
	const uint8_t *ptr = seg.block.data();
	uint64_t min = UINT64_MAX, max = 0;
	for (uint64_t i = 0; i < seg.count; ++i) {
		uint64_t curr = 0;
		memcpy(&curr, ptr + i * seg.type_size, seg.type_size);
		min = std::min(min, curr);
		max = std::max(max, curr);
	}
	uint64_t local_max = max;
	if (max != 0 && min != UINT64_MAX && max > min) {
		max -= min;
	}
	uint8_t min_width = PickWidth(max, flags);
	uint8_t old_width = seg.succinct_vec.width();
	if (old_width > min_width) {
		uint64_t *write_data = seg.succinct_vec.data();
		uint8_t write_offset = 0;
		for (uint64_t i = 0; i < seg.count; ++i) {
			int64_t x_flat = 0;
			memcpy(&x_flat, ptr + i * seg.type_size, seg.type_size);
			if (min != UINT64_MAX) {
				x_flat -= min;
			}
			sdsl::bits::write_int_and_move(write_data, (uint64_t)x_flat, write_offset, min_width);
		}
		seg.succinct_vec.bit_resize(seg.count * min_width);
		seg.succinct_vec.width(min_width);
		seg.repacked = true;
	}
	if (!(flags & FLAG_LITERAL)) {
		seg.min_factor = min;
		seg.max_factor = local_max;
	}
	seg.function = FUNC_SUCCINCT;
	seg.compacted = true;
}

void ref_scan_partial(void *h, uint64_t start, uint64_t scan_count, void *result, uint64_t result_offset, int flags) {
	auto &seg = *(RefSegment *)h;
	auto &source = seg.succinct_vec;
	uint8_t *target_ptr = (uint8_t *)result + result_offset * seg.type_size;
	bool add = AddsMin(seg, flags);
	for (uint64_t i = 0; i < scan_count; ++i) {
		auto entry_at_i = uint64_t(source[start + i]);
		if (add) {
			entry_at_i += seg.min_factor;
		}
		memcpy(target_ptr + i * seg.type_size, &entry_at_i, seg.type_size);
	}
}

void ref_uncompress_succinct(void *h, int flags) {
	auto &seg = *(RefSegment *)h;
	seg.block.assign(seg.segment_size ? seg.segment_size : 1, 0);
	bool add = (flags & FLAG_LITERAL) ? true : AddsMin(seg, flags);
	for (size_t i = 0; i < seg.succinct_vec.size(); ++i) {
		uint64_t curr = uint64_t(seg.succinct_vec[i]);
		if (add) {
			curr += seg.min_factor;
		}
		memcpy(seg.block.data() + i * seg.type_size, &curr, seg.type_size);
	}
	seg.function = FUNC_UNCOMPRESSED;
	seg.compacted = false;
	seg.repacked = false;
	seg.succinct_vec.resize(0);
}

uint64_t ref_segment_data_size(void *h) {
	auto &seg = *(RefSegment *)h;
	if (seg.function == FUNC_SUCCINCT) {
		return sdsl::size_in_bytes(seg.succinct_vec);
	}
	return seg.segment_size;
}

// field ids shared with the L1 oracle's orc_segment_get
uint64_t ref_segment_get(void *h, int field) {
	auto &seg = *(RefSegment *)h;
	switch (field) {
	case 0: return seg.count;
	case 1: return seg.succinct_vec.width();
	case 2: return seg.succinct_vec.bit_size();
	case 3: return seg.min_factor;
	case 4: return seg.max_factor;
	case 5: return seg.compacted;
	case 6: return seg.repacked;
	case 7: return (uint64_t)seg.function;
	case 8: return seg.succinct_vec.capacity() >> 6; // payload words
	case 9: return seg.num_elements;
	default: return 0;
	}
}

const uint64_t *ref_segment_words(void *h) {
	return ((RefSegment *)h)->succinct_vec.data();
}

const uint8_t *ref_segment_block(void *h) {
	return ((RefSegment *)h)->block.data();
}

// int_vector<0>::serialize through a string stream; returns bytes written (== size_in_bytes)
uint64_t ref_segment_serialize(void *h, uint8_t *out, uint64_t out_capacity) {
	auto &seg = *(RefSegment *)h;
	std::ostringstream os;
	uint64_t n = seg.succinct_vec.serialize(os);
	const std::string &s = os.str();
	if (s.size() <= out_capacity) {
		memcpy(out, s.data(), s.size());
	}
	return n;
}

// Timed CPU baseline on the reference's SDSL: one segment per task, segments dealt round-robin.
struct ref_bench_result {
	double analyze_s, encode_s, scan_s, sum_s, uncompact_s;
	uint64_t values;
	uint64_t sum_lo, sum_hi, sum_count;
	uint64_t packed_bytes;
};

int ref_bench_column(const void *values, uint64_t n, uint8_t type_size, int is_signed, uint64_t seg_capacity,
                     int threads, int copy_per_call, ref_bench_result *out) {
	if (threads < 1 || seg_capacity == 0 || n == 0) {
		return -1;
	}
	uint64_t n_segments = (n + seg_capacity - 1) / seg_capacity;
	std::vector<void *> segs(n_segments);
	for (auto &s : segs) {
		s = ref_segment_new_succinct(type_size, is_signed, seg_capacity * type_size);
	}
	const uint8_t *base = (const uint8_t *)values;
	double phase_s[5] = {0, 0, 0, 0, 0};
	std::vector<unsigned __int128> sums(threads, 0);
	std::vector<uint64_t> packed(threads, 0);
	for (int phase = 0; phase < 5; phase++) {
		std::vector<std::thread> pool;
		std::vector<double> secs(threads, 0.0);
		for (int tid = 0; tid < threads; tid++) {
			pool.emplace_back([&, tid]() {
				std::vector<uint8_t> sink(2048 * 8);
				auto t0 = std::chrono::steady_clock::now();
				for (uint64_t g = tid; g < n_segments; g += threads) {
					auto &seg = *(RefSegment *)segs[g];
					uint64_t first = g * seg_capacity;
					uint64_t cnt = std::min<uint64_t>(n - first, seg_capacity);
					if (phase == 0) {
						for (uint64_t off = 0; off < cnt; off += 2048) {
							uint64_t m = std::min<uint64_t>(cnt - off, 2048);
							ref_segment_append(&seg, base + (first + off) * type_size, nullptr, nullptr, 0, m);
						}
					} else if (phase == 1) {
						ref_bit_compress_from_succinct(&seg, 0);
						packed[tid] += seg.succinct_vec.capacity() >> 3;
					} else if (phase == 4) {
						ref_uncompress_succinct(&seg, 0);
					} else {
						for (uint64_t off = 0; off < seg.count; off += 2048) {
							uint64_t m = std::min<uint64_t>(seg.count - off, 2048);
							if (copy_per_call) {
								RefSegment copy = seg; // quirk Q4: deep copy per scan call
								ref_scan_partial(&copy, off, m, sink.data(), 0, 0);
							} else {
								ref_scan_partial(&seg, off, m, sink.data(), 0, 0);
							}
							if (phase == 3) {
								for (uint64_t i = 0; i < m; i++) {
									uint64_t v = 0;
									memcpy(&v, sink.data() + i * type_size, type_size);
									if (is_signed && type_size < 8) {
										unsigned sh = 64 - 8 * type_size;
										v = (uint64_t)(((int64_t)(v << sh)) >> sh);
									}
									sums[tid] += is_signed ? (unsigned __int128)(__int128)(int64_t)v : (unsigned __int128)v;
								}
							}
						}
					}
				}
				secs[tid] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
			});
		}
		for (auto &t : pool) {
			t.join();
		}
		phase_s[phase] = *std::max_element(secs.begin(), secs.end());
	}
	unsigned __int128 total = 0;
	uint64_t packed_bytes = 0;
	for (int t = 0; t < threads; t++) {
		total += sums[t];
		packed_bytes += packed[t];
	}
	out->analyze_s = phase_s[0];
	out->encode_s = phase_s[1];
	out->scan_s = phase_s[2];
	out->sum_s = phase_s[3];
	out->uncompact_s = phase_s[4];
	out->values = n;
	out->sum_lo = (uint64_t)total;
	out->sum_hi = (uint64_t)(total >> 64);
	out->sum_count = n;
	out->packed_bytes = packed_bytes;
	for (auto s : segs) {
		ref_segment_delete(s);
	}
	return 0;
}

} // extern "C"

// ---- block checksum: the reference's own duckdb::Checksum (src/common/checksum.cpp:6-25), compiled from the
// reference tree where it lies (oracle/Makefile: duckdb_checksum.o).  Its string-hash tail is only reached for sizes
// that are not a multiple of 8; DuckDB blocks never are, so the symbol is satisfied by a trap.
namespace duckdb {
uint64_t Checksum(uint8_t *buffer, size_t size);
uint64_t Hash(uint8_t *, size_t) {
	abort();
}
} // namespace duckdb

extern "C" uint64_t ref_checksum(const uint8_t *buffer, uint64_t size) {
	if (size % 8 != 0) {
		abort();
	}
	return duckdb::Checksum(const_cast<uint8_t *>(buffer), (size_t)size);
}
