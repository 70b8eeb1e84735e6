/*
 * TEST INFRASTRUCTURE ONLY (oracle/_ref build) — the REFERENCE's own recalculate-mapq code behind the oracle API.
 *
 * recalculate-mapq/main.cc is the CLI's translation unit (Boost.Iostreams, gengetopt output, seqan3 SAM I/O: not buildable
 * here), but the two per-record computations sit in its first part and depend on nothing of that:
 *   cigar_alignment_scorer <t_aln_rec>::calculate_score      main.cc:189-283 (fill_penalties :181-186)
 *   bowtie2_v2_score_calculator::calculate_mapq              main.cc:479-565 (tables :378-430)
 * The Makefile cuts lines 36-566 of the file WHERE IT LIES into oracle/_ref/shim/recalculate_mapq_excerpt.hh at build time
 * (generated, git-ignored — nothing of the reference is copied into the repository); this file includes the excerpt and
 * instantiates the scorer with a record type of its own (the template only calls sequence(), cigar_sequence(),
 * base_qualities()). `private` is lifted for the include so that calculate_score can be called without the tag handling
 * of operator(). libbio_fail / libbio_always_assert throw here (PV3_SHIM_FAIL_THROWS) and become the record's flag.
 * Inputs on which the reference reads out of range (SEQ or QUAL shorter than the CIGAR consumes, next score above the
 * score) are screened out before the call and flagged like the port flags them.
 */
#include <algorithm>
#include <array>
#include <cstdint>
#include <iostream>
#include <limits>
#include <optional>
#include <variant>
#include <vector>
#include <libbio/assert.hh>
#include <panvc3/sam_tag.hh>
#include <panvc3/utility.hh>
#include <range/v3/algorithm/upper_bound.hpp>
#include <seqan3/alphabet/cigar/cigar.hpp>
#include <seqan3/alphabet/nucleotide/dna5.hpp>
#include <seqan3/alphabet/quality/phred42.hpp>
#include <seqan3/io/sam_file/sam_tag_dictionary.hpp>

namespace lb = libbio;

#define private public
#include <recalculate_mapq_excerpt.hh>
#undef private

namespace {
	struct plain_record
	{
		std::vector <seqan3::dna5>		m_sequence;
		std::vector <seqan3::cigar>		m_cigar;
		std::vector <seqan3::phred42>	m_qualities;
		seqan3::sam_tag_dictionary		m_tags;
		auto const &sequence() const { return m_sequence; }
		auto const &cigar_sequence() const { return m_cigar; }
		auto const &base_qualities() const { return m_qualities; }
		auto &tags() { return m_tags; }
		auto const &tags() const { return m_tags; }
	};
}

extern "C" {

void pv3o_rescore_alignments(const int32_t *scoring5, uint64_t n, const uint64_t *seq_off, const uint8_t *seq,
	const uint64_t *qual_off, const uint8_t *qual, const uint64_t *cig_off, const uint32_t *cig,
	const uint64_t *refn_off, const uint8_t *refn, double *score, uint32_t *flags)
{
	using seqan3::operator""_cigar_operation;
	cigar_alignment_scorer <plain_record> scorer(alignment_scoring{scoring5[0], scoring5[1], scoring5[2], scoring5[3], scoring5[4]}, 0);
	static char const op_chars[] = "MIDNSHP=X";
	for (uint64_t i = 0; i < n; ++i)
	{
		plain_record rec;
		uint64_t const nq = seq_off[i + 1] - seq_off[i], nl = qual_off[i + 1] - qual_off[i];
		rec.m_sequence.resize(nq); rec.m_qualities.resize(nl);
		for (uint64_t k = 0; k < nq; ++k) rec.m_sequence[k].assign_rank(seq[seq_off[i] + k]);
		for (uint64_t k = 0; k < nl; ++k) rec.m_qualities[k].assign_rank(std::min <uint8_t>(qual[qual_off[i] + k], 41));
		std::vector <std::uint8_t> ref_n;
		if (refn_off) ref_n.assign(refn + refn_off[i], refn + refn_off[i + 1]);
		score[i] = 0; flags[i] = 0;
		// screen: an op seqan3::cigar cannot hold, or reads past SEQ / QUAL (undefined behaviour in the reference)
		uint64_t qpos = 0;
		for (uint64_t c = cig_off[i]; c < cig_off[i + 1] && !flags[i]; ++c)
		{
			uint32_t const code = cig[c] & 15u, count = cig[c] >> 4;
			if (code > 8) { flags[i] = 1; break; }
			if (7 == code || 8 == code)
			{
				bool const need_qual_everywhere = (8 == code);
				for (uint32_t k = 0; k < count && !flags[i]; ++k)
				{
					if (qpos + k >= nq) flags[i] = 4;
					else if ((need_qual_everywhere || 3 == seq[seq_off[i] + qpos + k]) && qpos + k >= nl) flags[i] = 0x40;
				}
			}
			if (0 == code) break;                                          // the reference fails here before reading further
			if (7 == code || 8 == code || 1 == code || 4 == code) qpos += count;
			seqan3::cigar::operation op{};
			op.assign_char(op_chars[code]);
			rec.m_cigar.emplace_back(count, op);
		}
		if (flags[i] && 1 != flags[i]) continue;
		if (1 == flags[i]) continue;
		// an M: let the reference itself fail on it
		for (uint64_t c = cig_off[i]; c < cig_off[i + 1]; ++c)
			if (0 == (cig[c] & 15u)) { seqan3::cigar::operation op{}; op.assign_char('M'); rec.m_cigar.emplace_back(cig[c] >> 4, op); break; }
		try { score[i] = scorer.calculate_score(rec, ref_n); }
		catch (libbio::shim::failure const &) { flags[i] = 0xffffffffu; }
	}
}

void pv3o_calculate_mapq(uint64_t n, const uint32_t *read_len, const uint32_t *other_len, const double *score_in, const double *next_in,
	uint8_t *mapq, uint32_t *flags)
{
	bowtie2_v2_score_calculator calc;
	for (uint64_t i = 0; i < n; ++i)
	{
		mapq[i] = 0; flags[i] = 0;
		double const min_score(calc.calculate_read_min_score(read_len[i]) + calc.calculate_read_min_score(other_len[i]));
		if (!(score_in[i] < min_score) && score_in[i] < next_in[i]) { flags[i] = 0x10; continue; }   // (it - 1) before begin in the reference
		try { mapq[i] = calc.calculate_mapq(read_len[i], other_len[i], score_in[i], next_in[i]); }
		catch (...) { flags[i] = 0xffffffffu; }
	}
}

} // extern "C"
