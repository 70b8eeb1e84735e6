/*
 * TEST INFRASTRUCTURE ONLY (oracle/_ref build) — the REFERENCE's own tag filter and OA construction behind the oracle API.
 *
 * Both are statements inside project_task::process (project-alignments/project_alignments.cc), a member function of the
 * CLI's translation unit that cannot be built here. The Makefile cuts three pieces out of the file WHERE IT LIES into
 * oracle/_ref/shim/ at build time (generated, git-ignored — nothing of the reference is copied into the repository):
 *   project_alignments_oa_tag_type.inc      lines 38-46    seqan3::sam_tag_type <"OA"_tag>
 *   project_alignments_preserved_tags.inc   lines 53-87    the constinit table preserved_sam_tags
 *   project_alignments_tags_oa.inc          lines 762-845  `auto &tags(aln_rec.tags());` … the end of the OA block:
 *                                                          original NM, removal of the non-preserved tags, OA entry
 * and this file includes the last piece as the body of a function that provides the names it uses (aln_rec, ref_ids,
 * ref_id, src_pos, cigar_seq, oa_buffer, additional_preserved_tags, m_removed_tag_counts) with the reference's types:
 * src_pos is the int32 of seqan3's reference_position(), tags a seqan3::sam_tag_dictionary.
 */
#include <algorithm>
#include <array>
#include <cstdint>
#include <cstring>
#include <map>
#include <optional>
#include <sstream>
#include <string>
#include <utility>
#include <vector>
#include <seqan3/alphabet/cigar/cigar.hpp>
#include <seqan3/io/sam_file/detail/cigar.hpp>
#include <seqan3/io/sam_file/sam_flag.hpp>
#include <seqan3/io/sam_file/sam_tag_dictionary.hpp>

#include <project_alignments_oa_tag_type.inc>      // using seqan3::operator""_tag; sam_tag_type <"OA"_tag>

namespace {
#include <project_alignments_preserved_tags.inc>

	struct plain_record
	{
		seqan3::sam_tag_dictionary	m_tags;
		seqan3::sam_flag			m_flag{};
		std::uint8_t				m_mapq{};
		auto &tags() { return m_tags; }
		seqan3::sam_flag flag() const { return m_flag; }
		std::uint8_t mapping_quality() const { return m_mapq; }
	};

	// Runs the reference's statements on one record; returns the OA tag's value afterwards.
	std::string update_tags(
		plain_record &aln_rec,
		std::vector <std::string> const &ref_ids,
		std::int32_t const ref_id,
		std::int32_t const src_pos,
		std::vector <seqan3::cigar> const &cigar_seq,
		std::vector <std::uint16_t> const &additional_preserved_tags,
		std::map <std::uint16_t, std::size_t> &m_removed_tag_counts
	)
	{
		std::stringstream oa_buffer;
#include <project_alignments_tags_oa.inc>
		return aln_rec.tags().template get <"OA"_tag>();
	}
}

extern "C" {

uint64_t pv3o_format_oa(uint64_t n, const uint32_t *ref_id, const uint64_t *pos, const uint16_t *flag, const uint8_t *mapq,
	const int32_t *nm, const uint8_t *has_nm, const uint64_t *cig_off, const uint32_t *cig_len, const uint32_t *cig,
	uint32_t n_names, const uint64_t *name_off, const char *names, uint64_t *text_off, char *text, uint64_t cap)
{
	if (!ref_id) return ~0ull;                         // CIGAR-only text is not a piece of the reference
	std::vector <std::string> ref_ids;
	for (uint32_t k = 0; k < n_names; ++k) ref_ids.emplace_back(names + name_off[k], names + name_off[k + 1]);
	static char const op_chars[] = "MIDNSHP=X";
	std::map <std::uint16_t, std::size_t> removed;
	uint64_t at = 0;
	for (uint64_t i = 0; i < n; ++i)
	{
		plain_record rec;
		rec.m_flag = seqan3::sam_flag(flag[i]);
		rec.m_mapq = mapq[i];
		if (has_nm && has_nm[i]) rec.m_tags.template get <"NM"_tag>() = nm[i];
		std::vector <seqan3::cigar> cigar_seq;
		uint64_t const c0 = cig_off[i], c1 = cig_len ? c0 + cig_len[i] : cig_off[i + 1];
		for (uint64_t c = c0; c < c1; ++c)
		{
			seqan3::cigar::operation op{};
			op.assign_char(op_chars[cig[c] & 15u]);
			cigar_seq.emplace_back(cig[c] >> 4, op);
		}
		std::string const oa(update_tags(rec, ref_ids, std::int32_t(ref_id[i]), std::int32_t(pos[i]), cigar_seq, {}, removed));
		text_off[i] = at;
		if (at + oa.size() <= cap && text) std::memcpy(text + at, oa.data(), oa.size());
		at += oa.size();
	}
	text_off[n] = at;
	return at;
}

/* The tag filter alone (project_alignments.cc:774-800 with the table :53-87): kept[k] = 1 when tags[k] survives.
 * additional must be sorted, as the reference keeps it. removed_total receives the sum of the removal counters. */
void pv3o_filter_tags(uint64_t n_tags, const uint16_t *tags, uint64_t n_additional, const uint16_t *additional, uint8_t *kept,
	uint64_t *removed_total)
{
	plain_record rec;
	for (uint64_t k = 0; k < n_tags; ++k)
	{
		if ("OA"_tag == tags[k]) rec.m_tags[tags[k]] = std::string("x");   // the OA statements append to a string value
		else rec.m_tags[tags[k]] = std::int32_t(k);
	}
	std::vector <std::uint16_t> const add(additional, additional + n_additional);
	std::map <std::uint16_t, std::size_t> removed;
	std::vector <seqan3::cigar> cigar_seq;
	update_tags(rec, {"r"}, 0, 0, cigar_seq, add, removed);
	for (uint64_t k = 0; k < n_tags; ++k) kept[k] = rec.m_tags.count(tags[k]) ? 1 : 0;
	uint64_t total = 0;
	for (auto const &kv : removed) total += kv.second;
	*removed_total = total;
}

} // extern "C"
