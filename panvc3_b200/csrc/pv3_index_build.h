// panvc3_b200 — host-side index construction shared by the library (pv3_api.cu) and the CPU emulation harness of the
// rewrite core (tests/cpp/k1w_emul.cc): packed target codes and the rank/select directory of one gap row.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstring>

namespace pv3 {

// dna5::assign_char (lib/seqan3 dna5.hpp:99-110) rank in bits 0-2, bit 3 set unless the byte is exactly one of "ACGNT":
// (query rank == nibble) ⇔ (dna5::to_char(query) == byte), the raw byte compare of rewrite_cigar.hh:104-110.
inline uint8_t code4_of_char(char ch)
{
	switch (ch) {
		case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'N': return 3; case 'T': return 4;
		case 'a': return 8; case 'c': return 9; case 'g': return 10;
		case 't': case 'U': case 'u': return 12;
		default: return 11;
	}
}

constexpr size_t CODES4_PAD = 512;
inline size_t codes4_bytes(size_t n) { return (n + 1) / 2 + CODES4_PAD; }

// two bases per byte, even base in the low nibble; CODES4_PAD readable bytes of 0xff follow
inline void pack_codes4(const char *bases, size_t n, uint8_t *codes4)
{
	for (size_t i = 0; i + 1 < n; i += 2) codes4[i >> 1] = (uint8_t) (code4_of_char(bases[i]) | (code4_of_char(bases[i + 1]) << 4));
	if (n & 1) codes4[n >> 1] = (uint8_t) (code4_of_char(bases[n - 1]) | 0xf0);
	memset(codes4 + (n + 1) / 2, 0xff, CODES4_PAD);
}

} // namespace pv3
