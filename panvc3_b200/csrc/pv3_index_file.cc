// panvc3_b200 — MSA index file: dependency-free reader and writer of the reference's on-disk format (host code).
//
// The file is panvc3::msa_index (reference include/panvc3/msa_index.hh:40-192) serialised through
// cereal::PortableBinaryOutputArchive (index-msa/input_processor.cc:117,197-200; read back in
// project-alignments/project_alignments.cc:1458-1465). Framing (lib/cereal/include/cereal/archives/portable_binary.hpp,
// lib/cereal/include/cereal/cereal.hpp class versions; pinned by tests/golden/reference_written.msa_index, a file the
// reference's own code wrote):
//   u8   1 = little-endian data
//   u32  class version of msa_index (first and only occurrence)
//   u64  number of chr entries, sorted by chr_id; per entry:
//     u32  class version of chr_entry (first occurrence only)      string chr_id (u64 length + bytes)
//     u64  number of sequence entries, sorted by seq_id; per entry:
//       u32  class version of sequence_entry (first occurrence only)   string seq_id
//       sdsl::rrr_vector<15>: u64 size, then four int_vectors m_bt, m_btnr, m_btnrp, m_rank
//         (lib/sdsl-lite/include/sdsl/rrr_vector_15.hpp:389-407), each: u8 width, f32 growth factor, u64 size in BITS,
//         ((bits + 63) >> 6) u64 words (int_vector.hpp:1853-1858); the rank/select supports serialise nothing (:577-593, :721-736)
// rrr<15> (rrr_vector_15.hpp:50-80, 197-273): block b covers bits [15b, 15b+15); m_bt[b] (4 bits) = popcount of the block;
// its offset among the C(15, bt) patterns of that class takes space_for_bt[bt] bits of m_btnr; patterns are enumerated by
// std::next_permutation starting from the bt HIGHEST positions of a 15-element bool vector. m_btnrp / m_rank hold the bit
// position in m_btnr and the ones so far at every 32nd block, m_rank one more element with the total.
//
// The device index keeps plain words (pv3_device.cuh), so the reader expands every vector once; the writer re-creates the
// compressed form byte for byte (checked against the golden file), which makes this the format half of `index_msa`.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/panvc3_b200.h"

namespace {

std::string g_error;

int fail(const char *fmt, const char *a = "", unsigned long long b = 0)
{
	char buf[512];
	snprintf(buf, sizeof buf, fmt, a, b);
	g_error = buf;
	return PV3_ERR_INVALID;
}

struct Binomial15 {
	uint8_t  space[16];
	uint32_t c_prefix[17];
	uint16_t nr_to_bin[1 << 15];
	uint16_t bin_to_nr[1 << 15];
	Binomial15()
	{
		uint32_t cnt = 0;
		for (int i = 0; i <= 15; ++i) {
			c_prefix[i] = cnt;
			uint32_t class_cnt = 0;
			std::vector <bool> b(15, false);
			for (int j = 0; j < i; ++j) b[15 - j - 1] = true;
			do {
				uint32_t x = 0;
				for (int k = 0; k < 15; ++k) x |= (uint32_t) b[15 - k - 1] << (15 - 1 - k);
				nr_to_bin[cnt] = (uint16_t) x;
				bin_to_nr[x] = (uint16_t) class_cnt;
				++cnt; ++class_cnt;
			} while (std::next_permutation(b.begin(), b.end()));
			space[i] = class_cnt == 1 ? 0 : (uint8_t) (hi(class_cnt) + 1);
		}
		c_prefix[16] = cnt;
	}
	static uint32_t hi(uint64_t x) { uint32_t r = 0; while (x >>= 1) ++r; return r; }   // sdsl::bits::hi, hi(0) == 0
};

Binomial15 const &binomial() { static Binomial15 const b; return b; }

// `len` (<= 64) bits at bit position pos of a word array (sdsl int_vector::get_int)
uint64_t get_bits(const uint64_t *w, uint64_t pos, unsigned len)
{
	if (!len) return 0;
	uint64_t const i = pos >> 6; unsigned const o = (unsigned) (pos & 63);
	uint64_t v = w[i] >> o;
	if (o + len > 64) v |= w[i + 1] << (64 - o);
	return len == 64 ? v : v & ((1ull << len) - 1);
}

void set_bits(uint64_t *w, uint64_t pos, uint64_t x, unsigned len)
{
	if (!len) return;
	uint64_t const i = pos >> 6; unsigned const o = (unsigned) (pos & 63);
	w[i] |= x << o;
	if (o + len > 64) w[i + 1] |= x >> (64 - o);
}

struct IntVector { uint8_t width = 0; float growth = 1.5f; uint64_t bits = 0; std::vector <uint64_t> words; };

struct Entry { std::string chr_id, seq_id; uint64_t n_bits = 0; std::vector <uint64_t> words; };

} // namespace

struct pv3_index_file { std::vector <Entry> entries; };

namespace {

struct Reader {
	const uint8_t *p, *end;
	bool ok = true;
	bool take(void *dst, size_t n) { if ((size_t) (end - p) < n) { ok = false; memset(dst, 0, n); return false; } memcpy(dst, p, n); p += n; return true; }
	uint8_t u8() { uint8_t v; take(&v, 1); return v; }
	uint32_t u32() { uint32_t v; take(&v, 4); return v; }
	uint64_t u64() { uint64_t v; take(&v, 8); return v; }
	float f32() { float v; take(&v, 4); return v; }
	std::string str() { uint64_t n = u64(); if (!ok || (uint64_t) (end - p) < n) { ok = false; return {}; } std::string s((const char *) p, (size_t) n); p += n; return s; }
	IntVector int_vector()
	{
		IntVector v;
		v.width = u8(); v.growth = f32(); v.bits = u64();
		uint64_t const n_words = (v.bits + 63) >> 6;
		if (!ok || (uint64_t) (end - p) / 8 < n_words) { ok = false; return v; }
		v.words.resize(n_words + 1, 0);                 // one spare word: get_bits may touch w[i + 1]
		memcpy(v.words.data(), p, n_words * 8); p += n_words * 8;
		return v;
	}
};

// rrr<15> → plain words (bit = 1 ⇔ gap). Checks the sampled pointers and ranks on the way.
int decode_rrr(uint64_t size, IntVector const &bt, IntVector const &btnr, IntVector const &btnrp, IntVector const &rank, Entry &e)
{
	Binomial15 const &B = binomial();
	uint64_t const n_blocks_stored = size / 15 + 1, n_blocks = (size + 14) / 15;
	if (bt.width != 4 || bt.bits != n_blocks_stored * 4) return fail("msa index: rrr block-type vector has the wrong shape%s (%llu bits)", "", bt.bits);
	if (btnr.width != 1) return fail("msa index: rrr offset vector is not a bit vector%s", "");
	uint64_t const n_super = (n_blocks_stored + 31) / 32;
	if (!btnrp.width || btnrp.bits != n_super * btnrp.width || !rank.width || rank.bits < n_super * rank.width)
		return fail("msa index: rrr sample vectors have the wrong shape%s", "");
	e.n_bits = size;
	e.words.assign((size + 63) / 64 + 1, 0);
	uint64_t pos = 0, ones = 0;
	for (uint64_t b = 0; b < n_blocks; ++b) {
		if (b % 32 == 0) {
			if (get_bits(btnrp.words.data(), (b / 32) * btnrp.width, btnrp.width) != pos || get_bits(rank.words.data(), (b / 32) * rank.width, rank.width) != ones)
				return fail("msa index: rrr samples disagree with the blocks%s at block %llu", "", b);
		}
		unsigned const k = (unsigned) get_bits(bt.words.data(), b * 4, 4);
		unsigned const sp = B.space[k];
		if (pos + sp > btnr.bits) return fail("msa index: rrr offsets run past the end%s", "");
		uint32_t const nr = (uint32_t) get_bits(btnr.words.data(), pos, sp);
		if (B.c_prefix[k] + nr >= B.c_prefix[k + 1]) return fail("msa index: rrr offset out of range%s at block %llu", "", b);
		uint64_t bits = B.nr_to_bin[B.c_prefix[k] + nr];
		unsigned const len = (unsigned) std::min <uint64_t>(15, size - 15 * b);
		if (len < 15 && (bits >> len)) return fail("msa index: rrr last block has bits beyond the size%s", "");
		set_bits(e.words.data(), 15 * b, bits, len);
		pos += sp; ones += k;
	}
	e.words.resize((size + 63) / 64);
	return PV3_OK;
}

struct Writer {
	std::vector <uint8_t> out;
	void put(const void *src, size_t n) { const uint8_t *s = (const uint8_t *) src; out.insert(out.end(), s, s + n); }
	void u8(uint8_t v) { put(&v, 1); }
	void u32(uint32_t v) { put(&v, 4); }
	void u64(uint64_t v) { put(&v, 8); }
	void str(std::string const &s) { u64(s.size()); put(s.data(), s.size()); }
	void int_vector(uint8_t width, uint64_t bits, std::vector <uint64_t> const &words)
	{
		float const growth = 1.5f;
		u8(width); put(&growth, 4); u64(bits);
		put(words.data(), ((bits + 63) >> 6) * 8);
	}
};

// plain words → the four vectors of sdsl::rrr_vector<15> (constructor, rrr_vector_15.hpp:197-273), serialised
void encode_rrr(Writer &w, const uint64_t *words_in, uint64_t size)
{
	Binomial15 const &B = binomial();
	std::vector <uint64_t> src((size + 63) / 64 + 1, 0);
	if (size) memcpy(src.data(), words_in, ((size + 63) / 64) * 8);
	if (size % 64) src[(size - 1) / 64] &= ~0ull >> (64 - size % 64);
	uint64_t const n_stored = size / 15 + 1, n_blocks = (size + 14) / 15, n_super = (n_stored + 31) / 32;
	std::vector <uint8_t> bt(n_stored, 0);
	uint64_t btnr_bits = 0, total_ones = 0;
	for (uint64_t b = 0; b < n_blocks; ++b) {
		unsigned const len = (unsigned) std::min <uint64_t>(15, size - 15 * b);
		bt[b] = (uint8_t) __builtin_popcountll(get_bits(src.data(), 15 * b, len));
		btnr_bits += B.space[bt[b]]; total_ones += bt[b];
	}
	uint8_t const wp = (uint8_t) (Binomial15::hi(btnr_bits) + 1), wr = (uint8_t) (Binomial15::hi(total_ones) + 1);
	uint64_t const btnr_size = std::max <uint64_t>(btnr_bits, 64);
	uint64_t const n_rank = n_super + ((size % (32 * 15)) > 0 ? 1 : 0);
	std::vector <uint64_t> v_bt((n_stored * 4 + 63) / 64 + 1, 0), v_btnr((btnr_size + 63) / 64 + 1, 0),
		v_btnrp((n_super * wp + 63) / 64 + 1, 0), v_rank((n_rank * wr + 63) / 64 + 1, 0);
	uint64_t pos = 0, ones = 0;
	for (uint64_t b = 0; b < n_blocks; ++b) {
		if (b % 32 == 0) { set_bits(v_btnrp.data(), (b / 32) * wp, pos, wp); set_bits(v_rank.data(), (b / 32) * wr, ones, wr); }
		unsigned const len = (unsigned) std::min <uint64_t>(15, size - 15 * b);
		unsigned const sp = B.space[bt[b]];
		set_bits(v_bt.data(), b * 4, bt[b], 4);
		if (sp) set_bits(v_btnr.data(), pos, B.bin_to_nr[get_bits(src.data(), 15 * b, len)], sp);
		pos += sp; ones += bt[b];
	}
	// a trailing dummy block (size a multiple of 15) that opens a new superblock gets no samples: the constructor leaves zeros
	{
		// m_rank[m_rank.size() - 1] = total (overwrites whatever the loop stored there)
		uint64_t const at = (n_rank - 1) * wr;
		uint64_t const i = at >> 6; unsigned const o = (unsigned) (at & 63);
		uint64_t const mask = wr == 64 ? ~0ull : ((1ull << wr) - 1);
		v_rank[i] &= ~(mask << o);
		if (o + wr > 64) v_rank[i + 1] &= ~(mask >> (64 - o));
		set_bits(v_rank.data(), at, total_ones, wr);
	}
	w.u64(size);
	w.int_vector(4, n_stored * 4, v_bt);
	w.int_vector(1, btnr_size, v_btnr);
	w.int_vector(wp, n_super * wp, v_btnrp);
	w.int_vector(wr, n_rank * wr, v_rank);
}

} // namespace

extern "C" {

const char *pv3_index_file_error(void) { return g_error.c_str(); }

int pv3_index_file_open(const char *path, pv3_index_file **out)
{
	if (!path || !out) return fail("pv3_index_file_open: NULL argument%s", "");
	*out = nullptr;
	FILE *f = fopen(path, "rb");
	if (!f) return fail("msa index: cannot open %s", path);
	std::vector <uint8_t> data;
	uint8_t buf[1 << 16];
	for (size_t n; (n = fread(buf, 1, sizeof buf, f)) > 0; ) data.insert(data.end(), buf, buf + n);
	fclose(f);
	Reader r{ data.data(), data.data() + data.size() };
	if (r.u8() != 1) return fail("msa index: %s is not a little-endian cereal PortableBinary archive", path);
	r.u32();                                                     // msa_index class version
	uint64_t const n_chr = r.u64();
	auto file = new pv3_index_file;
	bool chr_version_seen = false, seq_version_seen = false;
	for (uint64_t c = 0; c < n_chr && r.ok; ++c) {
		if (!chr_version_seen) { r.u32(); chr_version_seen = true; }
		std::string const chr_id = r.str();
		uint64_t const n_seq = r.u64();
		for (uint64_t s = 0; s < n_seq && r.ok; ++s) {
			if (!seq_version_seen) { r.u32(); seq_version_seen = true; }
			Entry e;
			e.chr_id = chr_id; e.seq_id = r.str();
			uint64_t const size = r.u64();
			IntVector const bt = r.int_vector(), btnr = r.int_vector(), btnrp = r.int_vector(), rank = r.int_vector();
			if (!r.ok) break;
			if (decode_rrr(size, bt, btnr, btnrp, rank, e) != PV3_OK) { delete file; return PV3_ERR_INVALID; }
			file->entries.emplace_back(std::move(e));
		}
	}
	if (!r.ok) { delete file; return fail("msa index: %s is truncated", path); }
	if (r.p != r.end) { delete file; return fail("msa index: %s has %llu trailing bytes", path, (unsigned long long) (r.end - r.p)); }
	*out = file;
	return PV3_OK;
}

uint64_t pv3_index_file_entries(const pv3_index_file *f) { return f ? f->entries.size() : 0; }

int pv3_index_file_entry(const pv3_index_file *f, uint64_t i, const char **chr_id, const char **seq_id, const uint64_t **gap_words, uint64_t *n_bits)
{
	if (!f || i >= f->entries.size()) return fail("pv3_index_file_entry: no such entry%s", "");
	Entry const &e = f->entries[i];
	if (chr_id) *chr_id = e.chr_id.c_str();
	if (seq_id) *seq_id = e.seq_id.c_str();
	if (gap_words) *gap_words = e.words.data();
	if (n_bits) *n_bits = e.n_bits;
	return PV3_OK;
}

void pv3_index_file_close(pv3_index_file *f) { delete f; }

int pv3_index_file_write(const char *path, uint64_t n, const char *const *chr_ids, const char *const *seq_ids,
	const uint64_t *const *gap_words, const uint64_t *n_bits)
{
	if (!path || (n && (!chr_ids || !seq_ids || !gap_words || !n_bits))) return fail("pv3_index_file_write: NULL argument%s", "");
	// chr entries sorted by id, sequence entries sorted by id inside each (index-msa/input_processor.cc:186,197)
	std::vector <uint64_t> order(n);
	for (uint64_t i = 0; i < n; ++i) order[i] = i;
	std::stable_sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) {
		int const c = strcmp(chr_ids[a], chr_ids[b]);
		return c ? c < 0 : strcmp(seq_ids[a], seq_ids[b]) < 0;
	});
	Writer w;
	w.u8(1);
	w.u32(0);
	uint64_t n_chr = 0;
	for (uint64_t k = 0; k < n; ++k) if (!k || strcmp(chr_ids[order[k]], chr_ids[order[k - 1]])) ++n_chr;
	w.u64(n_chr);
	bool chr_version_written = false, seq_version_written = false;
	for (uint64_t k = 0; k < n; ) {
		uint64_t k1 = k;
		while (k1 < n && !strcmp(chr_ids[order[k1]], chr_ids[order[k]])) ++k1;
		if (!chr_version_written) { w.u32(0); chr_version_written = true; }
		w.str(chr_ids[order[k]]);
		w.u64(k1 - k);
		for (uint64_t j = k; j < k1; ++j) {
			if (!seq_version_written) { w.u32(0); seq_version_written = true; }
			w.str(seq_ids[order[j]]);
			encode_rrr(w, gap_words[order[j]], n_bits[order[j]]);
		}
		k = k1;
	}
	FILE *f = fopen(path, "wb");
	if (!f) return fail("msa index: cannot create %s", path);
	bool const ok = fwrite(w.out.data(), 1, w.out.size(), f) == w.out.size();
	if (fclose(f) != 0 || !ok) return fail("msa index: short write to %s", path);
	return PV3_OK;
}

} // extern "C"
