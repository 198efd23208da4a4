// swar8.hpp -- word-parallel packing of one-byte slots (four slots per 32-bit register), shared by pack_staged_kernel<1>
// (kernels_pack.cu), decode_kernel<1, ..., flat> (kernels_decode.cu) and the CPU test csrc/tests/swar8_test.cpp, which
// checks it against the scalar definition of the codec:
//     field = (slot - min_factor) & lo_set[w],  fields LSB first          (column_segment.cpp:365-378, bits.hpp:478-499)
//     slot  = low byte of (field + min_factor)                            (succinct.cpp:123-144)
// Plain C++ with host / device qualifiers under nvcc.
#pragma once

#include <cstdint>

#ifdef __CUDACC__
#define ADACOM_SWAR_HD __host__ __device__ __forceinline__
#else
#define ADACOM_SWAR_HD inline
#endif

namespace adacom {

// four slots -> four w-bit fields at byte positions (w <= 7).  (x | 0x80..) - (min & 0x7f..) cannot borrow across bytes
// and its low 7 bits per byte are those of slot - min; fmask keeps the low w bits of every byte.
ADACOM_SWAR_HD uint32_t swar8_fields(uint32_t slots, uint32_t sub7, uint32_t fmask) {
	return ((slots | 0x80808080u) - sub7) & fmask;
}

// four w-bit fields at byte positions -> 4w contiguous bits (two shift-and-merge steps)
ADACOM_SWAR_HD uint32_t swar8_squeeze(uint32_t x, uint32_t w) {
	const uint32_t p = (x & 0x00ff00ffu) | ((x & 0xff00ff00u) >> (8u - w));
	return (p & 0xffffu) | ((p >> 16) << (2u * w));
}

// 4w contiguous bits -> four w-bit fields at byte positions; m2 = 2^(2w) - 1, fm_even / fm_odd = the field mask at the
// even / odd bytes
ADACOM_SWAR_HD uint32_t swar8_spread(uint32_t q, uint32_t w, uint32_t m2, uint32_t fm_even, uint32_t fm_odd) {
	const uint32_t p = (q & m2) | ((q >> (2u * w)) << 16);
	return (p & fm_even) | ((p << (8u - w)) & fm_odd);
}

// four slots + the frame of reference, modulo 256 per byte: the low 7 bits cannot carry into the next byte, the top bit
// is an XOR.  add7 = (min & 0x7f) in every byte, add80 = (min & 0x80) in every byte.
ADACOM_SWAR_HD uint32_t swar8_add(uint32_t x, uint32_t add7, uint32_t add80) {
	return ((x & 0x7f7f7f7fu) + add7) ^ ((x & 0x80808080u) ^ add80);
}

}  // namespace adacom
