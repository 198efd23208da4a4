// swar8_test.cpp -- the word-parallel one-byte packing (csrc/swar8.hpp, used by pack_staged_kernel<1> and
// decode_kernel<1, .., flat>) against the scalar definition of the codec, exhaustively: every width 1..7, every frame of
// reference 0..255, every slot value in every byte lane (the other lanes hold different values, so a borrow or carry
// that crossed a byte would show).  Plain C++: g++ -std=c++17, no GPU.
#include "../swar8.hpp"

#include <cstdio>
#include <cstdlib>

using namespace adacom;

static void fail(const char *what, unsigned w, unsigned mn, unsigned v) {
	printf("SWAR8 MISMATCH %s w=%u min=%u v=%u\n", what, w, mn, v);
	exit(1);
}

int main() {
	unsigned long long checked = 0;
	for (unsigned w = 1; w <= 8; w++) {
		const uint32_t fm = ((1u << w) - 1u) * 0x01010101u, fm_even = fm & 0x00ff00ffu, fm_odd = fm & 0xff00ff00u;
		const uint32_t m2 = (uint32_t)((1ull << (2u * w)) - 1u);
		for (unsigned mn = 0; mn < 256; mn++) {
			const uint32_t sub7 = (mn & 0x7fu) * 0x01010101u, add7 = sub7, add80 = (mn & 0x80u) * 0x01010101u;
			for (unsigned v = 0; v < 256; v++) {
				uint8_t slot[8];
				for (int j = 0; j < 8; j++) {
					slot[j] = (uint8_t)(v + 37u * (unsigned)j * (unsigned)j + 11u * (unsigned)j);
				}
				uint32_t x0 = 0, x1 = 0;
				for (int j = 0; j < 4; j++) {
					x0 |= (uint32_t)slot[j] << (8 * j);
					x1 |= (uint32_t)slot[4 + j] << (8 * j);
				}
				// decode side: slot = low byte of (field + min), for ANY byte value (w = 8 is the slot-width segment)
				const uint32_t y0 = swar8_add(x0, add7, add80), y1 = swar8_add(x1, add7, add80);
				for (int j = 0; j < 4; j++) {
					if (((y0 >> (8 * j)) & 0xffu) != ((slot[j] + mn) & 0xffu) || ((y1 >> (8 * j)) & 0xffu) != ((slot[4 + j] + mn) & 0xffu)) {
						fail("add", w, mn, v);
					}
				}
				checked += 8;
				if (w == 8) {
					continue;  // re-packed payloads of a one-byte type are 1..7 bits wide
				}
				// encode side: field = (slot - min) & lo_set[w], fields LSB first
				const uint32_t f0 = swar8_fields(x0, sub7, fm), f1 = swar8_fields(x1, sub7, fm);
				uint64_t want = 0;
				for (int j = 0; j < 8; j++) {
					const uint32_t field = (uint32_t)(slot[j] - mn) & ((1u << w) - 1u);
					const uint32_t got = ((j < 4 ? f0 : f1) >> (8 * (j & 3))) & 0xffu;
					if (got != field) {
						fail("fields", w, mn, v);
					}
					want |= (uint64_t)field << (w * (unsigned)j);
				}
				const uint32_t q0 = swar8_squeeze(f0, w), q1 = swar8_squeeze(f1, w);
				const uint64_t c = (uint64_t)q0 | ((uint64_t)q1 << (4u * w));
				if (c != want || (q0 >> (4u * w)) != 0 || (q1 >> (4u * w)) != 0) {
					fail("squeeze", w, mn, v);
				}
				// and back: 4w bits -> four fields at byte positions -> slots
				if (swar8_spread(q0, w, m2, fm_even, fm_odd) != f0 || swar8_spread(q1, w, m2, fm_even, fm_odd) != f1) {
					fail("spread", w, mn, v);
				}
				const uint32_t d0 = swar8_add(swar8_spread(q0, w, m2, fm_even, fm_odd), add7, add80);
				for (int j = 0; j < 4; j++) {
					const uint32_t field = (uint32_t)(slot[j] - mn) & ((1u << w) - 1u);
					if (((d0 >> (8 * j)) & 0xffu) != ((field + mn) & 0xffu)) {
						fail("decode", w, mn, v);
					}
				}
				checked += 8;
			}
		}
	}
	printf("SWAR8_OK checked=%llu\n", checked);
	return 0;
}
