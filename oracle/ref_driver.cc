/*
 * TEST INFRASTRUCTURE ONLY — driver that exposes the REFERENCE's own hot path behind oracle/pv3_oracle_api.h.
 *
 * This file is ours; it calls (does not restate) the reference code compiled in place from /root/reference:
 *   panvc3::alignment_projector::project_alignment   libpanvc3/alignment_projector.cc:30
 *   panvc3::rewrite_cigar                            include/panvc3/rewrite_cigar.hh:19
 *   panvc3::indel_run_checker                        libpanvc3/indel_run_checker.cc:13-121
 *   panvc3::align_global                             include/panvc3/align.hh:66
 *   msa_index::sequence_entry (rrr_vector<15>)       include/panvc3/msa_index.hh:47-74,239-258
 * Built by oracle/Makefile into oracle/_ref/libpv3_ref.so (git-ignored). Never linked into the product.
 */
#include "pv3_oracle_api.h"

#include <panvc3/align.hh>
#include <panvc3/alignment_projector.hh>
#include <panvc3/indel_run_checker.hh>
#include <panvc3/msa_index.hh>
#include <panvc3/rewrite_cigar.hh>
#include <range/v3/view/transform.hpp>
#include <seqan3/alphabet/quality/qualified.hpp>

#include <cstring>
#include <memory>
#include <thread>
#include <vector>

namespace {
	char const bam_ops[] = "MIDNSHP=X";

	inline int bam_op_of_char(char c)
	{
		for (int i = 0; i < 9; ++i) if (bam_ops[i] == c) return i;
		return 15;
	}

	inline seqan3::cigar make_cigar(std::uint32_t packed)
	{
		seqan3::cigar cc;
		seqan3::cigar::operation op;
		op.assign_char(bam_ops[packed & 0xf]); // caller guarantees op < 9
		cc = op;
		cc = std::uint32_t(packed >> 4);
		return cc;
	}

	inline std::uint32_t pack_cigar(seqan3::cigar const &cc)
	{
		using seqan3::get;
		std::uint32_t const count(get <0>(cc));
		char const op(get <1>(cc).to_char());
		return (count << 4) | std::uint32_t(bam_op_of_char(op));
	}
}

struct pv3o_index
{
	// unique_ptr: sequence_entry holds rank/select supports with raw back-pointers (msa_index.hh:195-199).
	std::vector <std::unique_ptr <panvc3::msa_index::sequence_entry>>	rows;
	std::vector <std::vector <char>>										refs;
};

extern "C" {

const char *pv3o_kind(void) { return "reference"; }

pv3o_index *pv3o_index_create(void) { return new pv3o_index; }
void pv3o_index_destroy(pv3o_index *idx) { delete idx; }

int64_t pv3o_index_add_row(pv3o_index *idx, const uint64_t *gap_words, uint64_t n_bits)
{
	sdsl::bit_vector bv(n_bits, 0);
	if (n_bits)
		std::memcpy(bv.data(), gap_words, ((n_bits + 63) / 64) * 8);
	// Clear the padding bits of the last word (sdsl expects them to be zero).
	if (n_bits % 64)
		bv.data()[(n_bits - 1) / 64] &= (~std::uint64_t(0)) >> (64 - n_bits % 64);
	auto entry(std::make_unique <panvc3::msa_index::sequence_entry>(std::string(), bv));
	entry->fix_rank_select_pointers();
	idx->rows.emplace_back(std::move(entry));
	return std::int64_t(idx->rows.size()) - 1;
}

int64_t pv3o_index_add_ref(pv3o_index *idx, const char *bases, uint64_t n)
{
	idx->refs.emplace_back(bases, bases + n);
	return std::int64_t(idx->refs.size()) - 1;
}

uint64_t pv3o_rank0(const pv3o_index *idx, uint32_t row, uint64_t pos) { return idx->rows[row]->gap_positions_rank0_support(pos); }
uint64_t pv3o_select0(const pv3o_index *idx, uint32_t row, uint64_t k) { return idx->rows[row]->gap_positions_select0_support(k); }
int pv3o_access(const pv3o_index *idx, uint32_t row, uint64_t pos) { return idx->rows[row]->gap_positions[pos]; }
uint64_t pv3o_project_position(const pv3o_index *idx, uint32_t src_row, uint32_t dst_row, uint64_t pos)
{
	return idx->rows[src_row]->project_position(pos, *idx->rows[dst_row]);
}

} // extern "C"

namespace {
	struct record_result
	{
		std::uint64_t				dst_pos{};
		std::vector <std::uint32_t>	ops;
		std::vector <std::uint64_t>	rr;	// 4 per run
	};

	struct batch_view
	{
		pv3o_index const *idx;
		std::uint32_t const *src_row, *dst_row, *dst_ref;
		std::uint64_t const *src_pos, *seq_off;
		std::uint8_t const *seq;
		std::uint64_t const *qual_off;
		std::uint8_t const *qual;
		std::uint64_t const *cig_off;
		std::uint32_t const *cig;
		std::int32_t go, ge;
		int stage;
	};

	void process_shard(batch_view const &bv, std::uint64_t begin, std::uint64_t end, record_result *results)
	{
		panvc3::alignment_projector projector;
		panvc3::cigar_buffer rewrite_buffer;
		std::vector <seqan3::dna5> query;
		std::vector <seqan3::phred42> quals;
		std::vector <seqan3::cigar> cigar;

		for (std::uint64_t i(begin); i < end; ++i)
		{
			query.clear(); quals.clear(); cigar.clear();
			for (auto j(bv.seq_off[i]); j < bv.seq_off[i + 1]; ++j)
			{
				seqan3::dna5 cc; cc.assign_rank(bv.seq[j]);
				query.push_back(cc);
			}
			if (bv.qual_off)
			{
				for (auto j(bv.qual_off[i]); j < bv.qual_off[i + 1]; ++j)
				{
					seqan3::phred42 qq; qq.assign_rank(bv.qual[j]);
					quals.push_back(qq);
				}
			}
			for (auto j(bv.cig_off[i]); j < bv.cig_off[i + 1]; ++j)
				cigar.push_back(make_cigar(bv.cig[j]));

			auto const &src(*bv.idx->rows[bv.src_row[i]]);
			auto const &dst(*bv.idx->rows[bv.dst_row[i]]);
			auto const &ref_seq(bv.idx->refs[bv.dst_ref[i]]);
			auto &res(results[i]);

			if (0 == bv.stage)
			{
				projector.reset();
				res.dst_pos = projector.project_alignment(bv.src_pos[i], src, dst, ref_seq, query, cigar, quals, bv.go, bv.ge);
				for (auto const &cc : projector.alignment())
					res.ops.push_back(pack_cigar(cc));
				auto const &rref(projector.realigned_reference_ranges());
				auto const &rqry(projector.realigned_query_ranges());
				for (std::size_t k(0); k < rref.size(); ++k)
				{
					res.rr.push_back(rref[k].location);
					res.rr.push_back(rref[k].length);
					res.rr.push_back(rqry[k].location);
					res.rr.push_back(rqry[k].length);
				}
			}
			else
			{
				// Same call as alignment_projector.cc:42-50.
				res.dst_pos = panvc3::rewrite_cigar(
					bv.src_pos[i], cigar, src, dst,
					query | ranges::views::transform([](auto const dna_cc){ return dna_cc.to_char(); }),
					ref_seq, rewrite_buffer
				);
				for (auto const &cc : rewrite_buffer.operations())
					res.ops.push_back(pack_cigar(cc));
			}
		}
	}
}

extern "C" {

int pv3o_project_batch(const pv3o_index *idx, uint64_t n,
	const uint32_t *src_row, const uint32_t *dst_row, const uint32_t *dst_ref, const uint64_t *src_pos,
	const uint64_t *seq_off, const uint8_t *seq,
	const uint64_t *qual_off, const uint8_t *qual,
	const uint64_t *cig_off, const uint32_t *cig,
	int32_t gap_open, int32_t gap_ext, int n_threads, int stage,
	uint64_t *dst_pos,
	uint64_t *out_cig_off, uint32_t *out_cig, uint64_t out_cig_cap,
	uint64_t *rr_off, uint64_t *rr, uint64_t rr_cap,
	uint32_t *flags)
{
	batch_view const bv{idx, src_row, dst_row, dst_ref, src_pos, seq_off, seq, qual_off, qual, cig_off, cig, gap_open, gap_ext, stage};
	std::vector <record_result> results(n);

	if (n_threads < 1) n_threads = 1;
	if (std::uint64_t(n_threads) > n) n_threads = n ? int(n) : 1;
	if (1 == n_threads)
		process_shard(bv, 0, n, results.data());
	else
	{
		std::vector <std::thread> threads;
		for (int t(0); t < n_threads; ++t)
		{
			auto const b(n * t / n_threads), e(n * (t + 1) / n_threads);
			threads.emplace_back([&bv, b, e, &results]{ process_shard(bv, b, e, results.data()); });
		}
		for (auto &th : threads) th.join();
	}

	std::uint64_t co(0), ro(0);
	int status(0);
	for (std::uint64_t i(0); i < n; ++i)
	{
		auto const &res(results[i]);
		dst_pos[i] = res.dst_pos;
		if (flags) flags[i] = 0;
		out_cig_off[i] = co;
		rr_off[i] = ro;
		if (co + res.ops.size() > out_cig_cap || ro + res.rr.size() / 4 > rr_cap) { status = -1; continue; }
		std::copy(res.ops.begin(), res.ops.end(), out_cig + co);
		std::copy(res.rr.begin(), res.rr.end(), rr + 4 * ro);
		co += res.ops.size();
		ro += res.rr.size() / 4;
	}
	out_cig_off[n] = co;
	rr_off[n] = ro;
	return status;
}

int64_t pv3o_align_global(const uint8_t *ref_ranks, uint64_t n, const uint8_t *qry_ranks, uint64_t m,
	const uint8_t *qual, int32_t gap_open, int32_t gap_ext, uint32_t *out_ops, uint64_t cap, int32_t *score)
{
	typedef seqan3::dna5 seq_alphabet;
	typedef seqan3::phred42 qual_alphabet;
	panvc3::cigar_buffer buffer;
	std::int32_t sc(0);

	if (!qual)
	{
		std::vector <seq_alphabet> rr(n), qq(m);
		for (std::uint64_t i(0); i < n; ++i) rr[i].assign_rank(ref_ranks[i]);
		for (std::uint64_t i(0); i < m; ++i) qq[i].assign_rank(qry_ranks[i]);
		sc = panvc3::align_global <false>(rr, qq, gap_open, gap_ext, buffer);
	}
	else
	{
		typedef seqan3::qualified <seq_alphabet, qual_alphabet> qualified_alphabet;
		std::vector <qualified_alphabet> rr(n), qq(m);
		for (std::uint64_t i(0); i < n; ++i)
		{
			seq_alphabet cc; cc.assign_rank(ref_ranks[i]);
			rr[i] = cc;
			rr[i] = panvc3::max_letter <qual_alphabet>(); // alignment_projector.cc:108
		}
		for (std::uint64_t i(0); i < m; ++i)
		{
			seq_alphabet cc; cc.assign_rank(qry_ranks[i]);
			qual_alphabet ql; ql.assign_rank(qual[i]);
			qq[i] = cc;
			qq[i] = ql;
		}
		sc = panvc3::align_global <true, seq_alphabet, qual_alphabet>(rr, qq, gap_open, gap_ext, buffer);
	}

	if (score) *score = sc;
	auto const &ops(buffer.operations());
	if (ops.size() > cap) return -1;
	for (std::size_t i(0); i < ops.size(); ++i) out_ops[i] = pack_cigar(ops[i]);
	return std::int64_t(ops.size());
}

int64_t pv3o_indel_runs(const uint32_t *ops, uint64_t n_ops, uint64_t ref_pos, uint64_t *out, uint64_t cap, uint64_t *final_pos)
{
	panvc3::cigar_vector cv;
	for (std::uint64_t i(0); i < n_ops; ++i) cv.push_back(make_cigar(ops[i]));
	panvc3::indel_run_checker checker;
	checker.reset(cv, ref_pos);
	std::int64_t count(0);
	while (checker.find_next_range_for_realigning())
	{
		if (std::uint64_t(count) >= cap) return -1;
		auto const rng(checker.cigar_realigned_range());
		auto const rr(checker.reference_range());
		auto const qr(checker.query_range());
		auto *dst(out + 6 * count);
		dst[0] = rng.first - cv.begin();
		dst[1] = rng.second - cv.begin();
		dst[2] = rr.location; dst[3] = rr.length;
		dst[4] = qr.location; dst[5] = qr.length;
		++count;
	}
	if (final_pos)
	{
		final_pos[0] = checker.query_position();
		final_pos[1] = checker.reference_position();
	}
	return count;
}

void pv3o_quality_lut(int32_t *lut84)
{
	panvc3::base_quality_scorer <seqan3::phred42> scorer;
	seqan3::phred42 ref_q(panvc3::max_letter <seqan3::phred42>());
	for (int q(0); q < 42; ++q)
	{
		seqan3::phred42 qq; qq.assign_rank(q);
		lut84[2 * q + 0] = std::int32_t(scorer.score(true, ref_q, qq));   // implicit double→int32 truncation, SURVEY F4
		lut84[2 * q + 1] = std::int32_t(scorer.score(false, ref_q, qq));
	}
}

} // extern "C"


/* ---- MSA index file, written by the REFERENCE's own serialisation (golden files for the index-format tests) -----------
 * panvc3::msa_index::serialize (include/panvc3/msa_index.hh:158-192) through cereal::PortableBinaryOutputArchive, as
 * index-msa/input_processor.cc:117,197-200 does (entries sorted by id before they are written). Reference build only. */
#include <cereal/archives/portable_binary.hpp>
#include <cereal/types/string.hpp>
#include <cereal/types/vector.hpp>
#include <algorithm>
#include <fstream>

extern "C" {

int pv3o_write_msa_index_file(const char *path, uint64_t n_entries, const char *const *chr_ids, const char *const *seq_ids,
	const uint64_t *const *gap_words, const uint64_t *n_bits)
{
	panvc3::msa_index index;
	for (uint64_t i = 0; i < n_entries; ++i) {
		sdsl::bit_vector bv(n_bits[i], 0);
		if (n_bits[i]) {
			std::memcpy(bv.data(), gap_words[i], ((n_bits[i] + 63) / 64) * 8);
			if (n_bits[i] % 64)
				bv.data()[(n_bits[i] - 1) / 64] &= (~std::uint64_t(0)) >> (64 - n_bits[i] % 64);
		}
		auto it(std::find_if(index.chr_entries.begin(), index.chr_entries.end(), [&](auto const &e){ return e.chr_id == chr_ids[i]; }));
		if (it == index.chr_entries.end()) {
			index.chr_entries.emplace_back(std::string(chr_ids[i]));
			it = index.chr_entries.end() - 1;
		}
		it->sequence_entries.emplace_back(std::string(seq_ids[i]), bv);
	}
	for (auto &chr : index.chr_entries)
		std::sort(chr.sequence_entries.begin(), chr.sequence_entries.end());
	std::sort(index.chr_entries.begin(), index.chr_entries.end());
	std::ofstream os(path, std::ios::binary);
	if (!os) return -1;
	{
		cereal::PortableBinaryOutputArchive archive(os);
		archive(index);
	}
	os.flush();
	return os ? 0 : -1;
}

} // extern "C"

/* ---- index-msa: gap bits and the unaligned sequence of one MSA row, by the REFERENCE's own builder ----------------------
 * panvc3::msa_indices::index_vector_builder (index-msa/index_vector_builder.cc:16-63: begin_sequence, update_sequence,
 * end_sequence), fed in `chunk`-byte pieces as build_index_vector_one_sequence (:66-97) does with st_blksize buffers. The
 * unaligned FASTA the builder prints to std::cout (">chr/seq\n" + non-gap bytes + "\n", wrap 0) is captured; only the
 * sequence bytes are returned. gap_words: (n + 63) / 64 words; ungapped: capacity n (may be NULL). Reference build only. */
#include <index_vector_builder.hh>
#include <iostream>
#include <sstream>

extern "C" {

int pv3o_msa_row_extract_ref(const char *text, uint64_t n, uint64_t chunk, uint64_t *gap_words, char *ungapped, uint64_t *n_ungapped)
{
	bool const want_seq = ungapped != nullptr;
	panvc3::msa_indices::index_vector_builder builder(want_seq, 0);
	std::ostringstream captured;
	std::streambuf *old = std::cout.rdbuf(captured.rdbuf());
	builder.begin_sequence("c", "s", n);
	if (!chunk) chunk = 4096;
	for (uint64_t at = 0; at < n; at += chunk)
		builder.update_sequence(std::span <char const>(text + at, std::min <uint64_t>(chunk, n - at)));
	builder.end_sequence();
	std::cout.rdbuf(old);
	auto const &bv = builder.destination_bit_vector();
	if (bv.size() != n) return -1;
	std::memcpy(gap_words, bv.data(), ((n + 63) / 64) * 8);
	if (want_seq) {
		std::string const s = captured.str();
		std::size_t const hdr = s.find('\n');
		if (hdr == std::string::npos || s.empty() || s.back() != '\n') return -2;
		std::size_t const len = s.size() - hdr - 2;
		if (len > n) return -3;
		std::memcpy(ungapped, s.data() + hdr + 1, len);
		if (n_ungapped) *n_ungapped = len;
	}
	return 0;
}

} // extern "C"
