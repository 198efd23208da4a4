// oracle/l2/codecs_ref.cpp -- TEST INFRASTRUCTURE ONLY.  Oracle for SURVEY 8(f) row 4, the neighbour codecs.
//
// DuckDB's bitpacking and RLE CompressionFunctions, driven WITHOUT a database: this translation unit includes the
// reference's own src/storage/compression/bitpacking.cpp and rle.cpp where they lie (read-only), so the encoder decisions
// are the reference's BitpackingState<T>::Update / Flush (bitpacking.cpp:103-296: CONSTANT / CONSTANT_DELTA / DELTA_FOR /
// FOR per 2048-value metadata group) and RLEState<T>::Update (rle.cpp:24-77), the bit layout is the reference's
// BitpackingPrimitives::PackBuffer / UnPackBlock over its vendored third_party/fastpforlib, and the decode arithmetic is
// its ApplyFrameOfReference / DeltaDecode (bitpacking.cpp:543-578).  What is restated here (because the reference's
// versions need a ColumnDataCheckpointer / BufferManager / ColumnSegment) is only the glue around them:
//   * the segment writer   = BitpackingCompressState::BitpackingWriter + CreateEmptySegment / FlushSegment
//                            (bitpacking.cpp:376-513) and RLECompressState::WriteValue / FlushSegment (rle.cpp:170-215),
//   * the scan loop        = BitpackingScanState::LoadNextGroup / Skip + BitpackingScanPartial + BitpackingFetchRow
//                            (bitpacking.cpp:583-840) and RLEScanPartial / RLEFetchRow (rle.cpp:298-331).
// Built by oracle/l2/build_drivers.sh into oracle/_ref/libadacom_codecs_ref.so (linked against the L2 library for the
// exception / string symbols the included sources reference).
#include "duckdb/common/types/validity_mask.hpp"

#include "src/storage/compression/bitpacking.cpp"
#include "src/storage/compression/rle.cpp"

#include <cstdint>
#include <cstring>
#include <vector>

namespace {

using namespace duckdb;

constexpr idx_t kBlockSize = Storage::BLOCK_SIZE;
constexpr idx_t kGroup = 2048; // BITPACKING_METADATA_GROUP_SIZE

struct RefSegmentImage {
	std::vector<uint8_t> data; // the segment's block image (total_segment_size bytes)
	uint64_t count = 0;        // tuples
};

struct RefCodecColumn {
	std::vector<RefSegmentImage> segments;
	int type_size = 4;
	bool is_signed = false;
	bool rle = false;
};

//===--------------------------------------------------------------------===//
// bitpacking: segment writer (glue) around the reference's BitpackingState
//===--------------------------------------------------------------------===//
template <class T, class T_S = typename std::make_signed<T>::type>
struct RefBitpackingCompress {
	RefCodecColumn *column;
	std::vector<uint8_t> block;
	data_ptr_t base = nullptr, data_ptr = nullptr, metadata_ptr = nullptr;
	uint64_t count = 0;
	BitpackingState<T> state;

	explicit RefBitpackingCompress(RefCodecColumn *column_p, BitpackingMode mode) : column(column_p) {
		state.data_ptr = (void *)this;
		state.mode = mode;
		CreateEmptySegment();
	}

	void CreateEmptySegment() {
		block.assign(kBlockSize, 0);
		base = block.data();
		data_ptr = base + BitpackingPrimitives::BITPACKING_HEADER_SIZE;
		metadata_ptr = base + kBlockSize;
		count = 0;
	}
	idx_t RemainingSize() {
		return metadata_ptr - data_ptr;
	}
	void FlushSegment() {
		idx_t metadata_offset = AlignValue(data_ptr - base);
		idx_t metadata_size = base + kBlockSize - metadata_ptr;
		idx_t total_segment_size = metadata_offset + metadata_size;
		memmove(base + metadata_offset, metadata_ptr, metadata_size);
		Store<idx_t>(metadata_offset + metadata_size, base);
		RefSegmentImage image;
		image.data.assign(base, base + total_segment_size);
		image.count = count;
		column->segments.push_back(std::move(image));
	}
	void ReserveSpace(idx_t data_bytes) {
		idx_t required = data_bytes + sizeof(bitpacking_metadata_encoded_t);
		if (RemainingSize() < required) {
			FlushSegment();
			CreateEmptySegment();
		}
	}
	void WriteMetaData(BitpackingMode mode) {
		bitpacking_metadata_t metadata {mode, (uint32_t)(data_ptr - base)};
		metadata_ptr -= sizeof(bitpacking_metadata_encoded_t);
		Store<bitpacking_metadata_encoded_t>(EncodeMeta(metadata), metadata_ptr);
	}
	template <class V>
	void WriteData(V val) {
		Store<V>(val, data_ptr);
		data_ptr += sizeof(V);
	}

	struct Writer {
		static void WriteConstant(T constant, idx_t count, void *p, bool all_invalid) {
			auto s = (RefBitpackingCompress<T> *)p;
			s->ReserveSpace(sizeof(T));
			s->WriteMetaData(BitpackingMode::CONSTANT);
			s->WriteData(constant);
			s->count += count;
		}
		static void WriteConstantDelta(T_S constant, T frame_of_reference, idx_t count, T *values, bool *validity, void *p) {
			auto s = (RefBitpackingCompress<T> *)p;
			s->ReserveSpace(2 * sizeof(T));
			s->WriteMetaData(BitpackingMode::CONSTANT_DELTA);
			s->WriteData(frame_of_reference);
			s->WriteData(constant);
			s->count += count;
		}
		static void WriteDeltaFor(T *values, bool *validity, bitpacking_width_t width, T frame_of_reference, T_S delta_offset,
		                          T *original_values, idx_t count, void *p) {
			auto s = (RefBitpackingCompress<T> *)p;
			auto bp_size = BitpackingPrimitives::GetRequiredSize(count, width);
			s->ReserveSpace(bp_size + 3 * sizeof(T));
			s->WriteMetaData(BitpackingMode::DELTA_FOR);
			s->WriteData(frame_of_reference);
			s->WriteData((T)width);
			s->WriteData(delta_offset);
			BitpackingPrimitives::PackBuffer<T, false>(s->data_ptr, values, count, width);
			s->data_ptr += bp_size;
			s->count += count;
		}
		static void WriteFor(T *values, bool *validity, bitpacking_width_t width, T frame_of_reference, idx_t count, void *p) {
			auto s = (RefBitpackingCompress<T> *)p;
			auto bp_size = BitpackingPrimitives::GetRequiredSize(count, width);
			s->ReserveSpace(bp_size + 2 * sizeof(T));
			s->WriteMetaData(BitpackingMode::FOR);
			s->WriteData(frame_of_reference);
			s->WriteData((T)width);
			BitpackingPrimitives::PackBuffer<T, false>(s->data_ptr, values, count, width);
			s->data_ptr += bp_size;
			s->count += count;
		}
	};

	bool Run(const T *values, const uint8_t *validity, uint64_t n) {
		for (uint64_t i = 0; i < n; i++) {
			if (!state.template Update<Writer>(values[i], validity ? validity[i] != 0 : true)) {
				return false;
			}
		}
		if (!state.template Flush<Writer>()) {
			return false;
		}
		FlushSegment();
		return true;
	}
};

//===--------------------------------------------------------------------===//
// bitpacking: scan state (glue) -- LoadNextGroup / Skip / ScanPartial / FetchRow on a block image
//===--------------------------------------------------------------------===//
template <class T, class T_S = typename std::make_signed<T>::type>
struct RefBitpackingScan {
	data_ptr_t base;
	T decompression_buffer[kGroup];
	bitpacking_metadata_t current_group;
	bitpacking_width_t current_width = 0;
	T current_frame_of_reference = 0, current_constant = 0, current_delta_offset = 0;
	idx_t current_group_offset = 0;
	data_ptr_t current_group_ptr = nullptr, bitpacking_metadata_ptr = nullptr;

	explicit RefBitpackingScan(const RefSegmentImage &seg) : base((data_ptr_t)seg.data.data()) {
		auto off = Load<idx_t>(base);
		bitpacking_metadata_ptr = base + off - sizeof(bitpacking_metadata_encoded_t);
		LoadNextGroup();
	}
	void LoadNextGroup() {
		current_group_offset = 0;
		current_group = DecodeMeta((bitpacking_metadata_encoded_t *)bitpacking_metadata_ptr);
		bitpacking_metadata_ptr -= sizeof(bitpacking_metadata_encoded_t);
		current_group_ptr = base + current_group.offset;
		switch (current_group.mode) {
		case BitpackingMode::CONSTANT:
			current_constant = Load<T>(current_group_ptr);
			current_group_ptr += sizeof(T);
			break;
		default:
			current_frame_of_reference = Load<T>(current_group_ptr);
			current_group_ptr += sizeof(T);
			break;
		}
		switch (current_group.mode) {
		case BitpackingMode::CONSTANT_DELTA:
			current_constant = Load<T>(current_group_ptr);
			current_group_ptr += sizeof(T);
			break;
		case BitpackingMode::FOR:
		case BitpackingMode::DELTA_FOR:
			current_width = (bitpacking_width_t)Load<T>(current_group_ptr);
			current_group_ptr += MaxValue(sizeof(T), sizeof(bitpacking_width_t));
			break;
		default:
			break;
		}
		if (current_group.mode == BitpackingMode::DELTA_FOR) {
			current_delta_offset = Load<T>(current_group_ptr);
			current_group_ptr += sizeof(T);
		}
	}
	void Skip(idx_t skip_count) {
		while (skip_count > 0) {
			if (current_group_offset + skip_count < kGroup) {
				if (current_group.mode == BitpackingMode::DELTA_FOR) {
					idx_t extra_count = current_group_offset % BitpackingPrimitives::BITPACKING_ALGORITHM_GROUP_SIZE;
					idx_t base_decompress_count = BitpackingPrimitives::RoundUpToAlgorithmGroupSize(skip_count);
					idx_t decompress_count = base_decompress_count + extra_count;
					idx_t decompress_offset = current_group_offset - extra_count;
					// literal (bitpacking.cpp:685-687): the reference adds the VALUE offset to a byte pointer; harmless
					// here because this glue only ever skips on a fresh state (decompress_offset == 0), like FetchRow
					BitpackingPrimitives::UnPackBuffer<T>((data_ptr_t)decompression_buffer, current_group_ptr + decompress_offset,
					                                      decompress_count, current_width, true);
					ApplyFrameOfReference<T_S>((T_S *)&decompression_buffer[extra_count], current_frame_of_reference, skip_count);
					DeltaDecode<T_S>((T_S *)&decompression_buffer[extra_count], (T_S)current_delta_offset, (idx_t)skip_count);
					current_delta_offset = decompression_buffer[extra_count + skip_count - 1];
				}
				current_group_offset += skip_count;
				break;
			} else {
				auto left_in_this_group = kGroup - current_group_offset;
				auto number_of_groups_to_skip = (skip_count - left_in_this_group) / kGroup;
				current_group_offset = 0;
				bitpacking_metadata_ptr -= number_of_groups_to_skip * sizeof(bitpacking_metadata_encoded_t);
				LoadNextGroup();
				skip_count -= left_in_this_group;
				skip_count -= number_of_groups_to_skip * kGroup;
			}
		}
	}
	void ScanPartial(idx_t scan_count, T *result_data) {
		idx_t scanned = 0;
		while (scanned < scan_count) {
			if (current_group_offset >= kGroup) {
				LoadNextGroup();
			}
			idx_t offset_in_compression_group = current_group_offset % BitpackingPrimitives::BITPACKING_ALGORITHM_GROUP_SIZE;
			if (current_group.mode == BitpackingMode::CONSTANT) {
				idx_t to_scan = MinValue(scan_count - scanned, kGroup - current_group_offset);
				std::fill(result_data + scanned, result_data + scanned + to_scan, current_constant);
				scanned += to_scan;
				current_group_offset += to_scan;
				continue;
			}
			if (current_group.mode == BitpackingMode::CONSTANT_DELTA) {
				idx_t to_scan = MinValue(scan_count - scanned, kGroup - current_group_offset);
				for (idx_t i = 0; i < to_scan; i++) {
					result_data[scanned + i] = ((current_group_offset + i) * current_constant) + current_frame_of_reference;
				}
				scanned += to_scan;
				current_group_offset += to_scan;
				continue;
			}
			idx_t to_scan = MinValue<idx_t>(scan_count - scanned,
			                                BitpackingPrimitives::BITPACKING_ALGORITHM_GROUP_SIZE - offset_in_compression_group);
			data_ptr_t current_position_ptr = current_group_ptr + current_group_offset * current_width / 8;
			data_ptr_t decompression_group_start_pointer = current_position_ptr - offset_in_compression_group * current_width / 8;
			T *current_result_ptr = result_data + scanned;
			BitpackingPrimitives::UnPackBlock<T>((data_ptr_t)decompression_buffer, decompression_group_start_pointer,
			                                     current_width, true);
			memcpy(current_result_ptr, decompression_buffer + offset_in_compression_group, to_scan * sizeof(T));
			if (current_group.mode == BitpackingMode::DELTA_FOR) {
				ApplyFrameOfReference<T_S>((T_S *)current_result_ptr, (T_S)current_frame_of_reference, to_scan);
				DeltaDecode<T_S>((T_S *)current_result_ptr, (T_S)current_delta_offset, to_scan);
				current_delta_offset = ((T *)current_result_ptr)[to_scan - 1];
			} else {
				ApplyFrameOfReference<T>(current_result_ptr, current_frame_of_reference, to_scan);
			}
			scanned += to_scan;
			current_group_offset += to_scan;
		}
	}
	T FetchRow(idx_t row_id) {
		Skip(row_id);
		idx_t offset_in_compression_group = current_group_offset % BitpackingPrimitives::BITPACKING_ALGORITHM_GROUP_SIZE;
		if (current_group.mode == BitpackingMode::CONSTANT) {
			return current_constant;
		}
		if (current_group.mode == BitpackingMode::CONSTANT_DELTA) {
			return (T)(current_group_offset * current_constant + current_frame_of_reference);
		}
		data_ptr_t start = current_group_ptr + (current_group_offset - offset_in_compression_group) * current_width / 8;
		BitpackingPrimitives::UnPackBlock<T>((data_ptr_t)decompression_buffer, start, current_width, true);
		T result = decompression_buffer[offset_in_compression_group];
		result += current_frame_of_reference;
		if (current_group.mode == BitpackingMode::DELTA_FOR) {
			result += current_delta_offset;
		}
		return result;
	}
};

//===--------------------------------------------------------------------===//
// RLE: writer + scan glue around the reference's RLEState
//===--------------------------------------------------------------------===//
template <class T>
struct RefRLECompress {
	RefCodecColumn *column;
	std::vector<uint8_t> block;
	RLEState<T> state;
	idx_t entry_count = 0, max_rle_count, count = 0;

	explicit RefRLECompress(RefCodecColumn *column_p) : column(column_p) {
		auto entry_size = sizeof(T) + sizeof(rle_count_t);
		auto entries = (kBlockSize - RLEConstants::RLE_HEADER_SIZE) / entry_size;
		max_rle_count = (entries / STANDARD_VECTOR_SIZE) * STANDARD_VECTOR_SIZE; // RLECompressState::MaxRLECount
		state.dataptr = (void *)this;
		block.assign(kBlockSize, 0);
	}
	struct Writer {
		template <class VALUE_TYPE>
		static void Operation(VALUE_TYPE value, rle_count_t run, void *p, bool is_null) {
			((RefRLECompress<T> *)p)->WriteValue(value, run);
		}
	};
	void WriteValue(T value, rle_count_t run) {
		auto handle_ptr = block.data() + RLEConstants::RLE_HEADER_SIZE;
		((T *)handle_ptr)[entry_count] = value;
		((rle_count_t *)(handle_ptr + max_rle_count * sizeof(T)))[entry_count] = run;
		entry_count++;
		count += run;
		if (entry_count == max_rle_count) {
			FlushSegment();
			block.assign(kBlockSize, 0);
			entry_count = 0;
			count = 0;
		}
	}
	void FlushSegment() {
		idx_t counts_size = sizeof(rle_count_t) * entry_count;
		idx_t original_rle_offset = RLEConstants::RLE_HEADER_SIZE + max_rle_count * sizeof(T);
		idx_t minimal_rle_offset = AlignValue(RLEConstants::RLE_HEADER_SIZE + sizeof(T) * entry_count);
		idx_t total_segment_size = minimal_rle_offset + counts_size;
		memmove(block.data() + minimal_rle_offset, block.data() + original_rle_offset, counts_size);
		Store<uint64_t>(minimal_rle_offset, block.data());
		RefSegmentImage image;
		image.data.assign(block.data(), block.data() + total_segment_size);
		image.count = count;
		column->segments.push_back(std::move(image));
	}
	void Run(T *values, const uint8_t *validity, uint64_t n) {
		ValidityMask mask;
		if (validity) {
			mask.Initialize(n);
			for (uint64_t i = 0; i < n; i++) {
				mask.Set(i, validity[i] != 0);
			}
		}
		// RLEState::Update takes a ValidityMask sized like a vector; walk the column in vectors like RLECompress does
		for (uint64_t i = 0; i < n; i++) {
			state.template Update<Writer>(values, mask, i);
		}
		state.template Flush<Writer>();
		FlushSegment();
	}
};

template <class T>
void RefRLEScan(const RefSegmentImage &seg, uint64_t start, uint64_t scan_count, T *out) {
	// RLEScanState::Skip + RLEScanPartial (rle.cpp:262-322)
	auto data = seg.data.data();
	auto rle_count_offset = Load<uint64_t>((data_ptr_t)data);
	auto data_pointer = (const T *)(data + RLEConstants::RLE_HEADER_SIZE);
	auto index_pointer = (const rle_count_t *)(data + rle_count_offset);
	idx_t entry_pos = 0, position_in_entry = 0;
	for (idx_t i = 0; i < start; i++) {
		position_in_entry++;
		if (position_in_entry >= index_pointer[entry_pos]) {
			entry_pos++;
			position_in_entry = 0;
		}
	}
	for (idx_t i = 0; i < scan_count; i++) {
		out[i] = data_pointer[entry_pos];
		position_in_entry++;
		if (position_in_entry >= index_pointer[entry_pos]) {
			entry_pos++;
			position_in_entry = 0;
		}
	}
}

template <class T>
void *CompressT(const void *values, const uint8_t *validity, uint64_t n, bool is_signed, int mode, bool rle) {
	auto column = new RefCodecColumn();
	column->type_size = sizeof(T);
	column->is_signed = is_signed;
	column->rle = rle;
	if (rle) {
		RefRLECompress<T> compress(column);
		compress.Run((T *)values, validity, n);
	} else {
		RefBitpackingCompress<T> compress(column, (BitpackingMode)mode);
		if (!compress.Run((const T *)values, validity, n)) {
			delete column;
			return nullptr;
		}
	}
	return column;
}

template <class T>
void ScanT(RefCodecColumn &column, uint64_t seg, uint64_t start, uint64_t count, void *out) {
	if (column.rle) {
		RefRLEScan<T>(column.segments[seg], start, count, (T *)out);
		return;
	}
	RefBitpackingScan<T> scan(column.segments[seg]);
	scan.Skip(start);
	scan.ScanPartial(count, (T *)out);
}

template <class T>
void FetchT(RefCodecColumn &column, uint64_t seg, uint64_t row, void *out) {
	if (column.rle) {
		RefRLEScan<T>(column.segments[seg], row, 1, (T *)out);
		return;
	}
	RefBitpackingScan<T> scan(column.segments[seg]);
	*(T *)out = scan.FetchRow(row);
}

#define DISPATCH(fn, ...)                                                                                             \
	switch (type_size * 2 + (is_signed ? 1 : 0)) {                                                                    \
	case 2: return fn<uint8_t>(__VA_ARGS__);                                                                          \
	case 3: return fn<int8_t>(__VA_ARGS__);                                                                           \
	case 4: return fn<uint16_t>(__VA_ARGS__);                                                                         \
	case 5: return fn<int16_t>(__VA_ARGS__);                                                                          \
	case 8: return fn<uint32_t>(__VA_ARGS__);                                                                         \
	case 9: return fn<int32_t>(__VA_ARGS__);                                                                          \
	case 16: return fn<uint64_t>(__VA_ARGS__);                                                                        \
	default: return fn<int64_t>(__VA_ARGS__);                                                                         \
	}

} // namespace

extern "C" {

// mode: BitpackingMode (0 auto, 1 constant, 2 constant_delta, 3 delta_for, 4 for); codec 0 = bitpacking, 1 = rle
void *codecref_compress(const void *values, const uint8_t *validity, uint64_t n, int type_size, int is_signed, int mode,
                        int codec) {
	DISPATCH(CompressT, values, validity, n, is_signed != 0, mode, codec == 1)
}
void codecref_free(void *h) {
	delete (RefCodecColumn *)h;
}
uint64_t codecref_num_segments(void *h) {
	return ((RefCodecColumn *)h)->segments.size();
}
uint64_t codecref_segment_count(void *h, uint64_t s) {
	return ((RefCodecColumn *)h)->segments[s].count;
}
uint64_t codecref_segment_size(void *h, uint64_t s) {
	return ((RefCodecColumn *)h)->segments[s].data.size();
}
const uint8_t *codecref_segment_data(void *h, uint64_t s) {
	return ((RefCodecColumn *)h)->segments[s].data.data();
}
void codecref_scan(void *h, uint64_t seg, uint64_t start, uint64_t count, void *out) {
	auto &column = *(RefCodecColumn *)h;
	int type_size = column.type_size;
	bool is_signed = column.is_signed;
	DISPATCH(ScanT, column, seg, start, count, out)
}
void codecref_fetch_row(void *h, uint64_t seg, uint64_t row, void *out) {
	auto &column = *(RefCodecColumn *)h;
	int type_size = column.type_size;
	bool is_signed = column.is_signed;
	DISPATCH(FetchT, column, seg, row, out)
}

} // extern "C"
