/*
 * panvc3_b200 — host-side mirror of the record post-processing in project_task::process
 * (project-alignments/project_alignments.cc:762-918), SURVEY §8f N1: what happens to a record after
 * alignment_projector::project_alignment returned.
 *
 *   filter_tags()            :774-800   drop every tag that is neither in the fixed preserved set (:53-87) nor in the caller's
 *                                        additional list; count what was removed
 *   oa_batch / format_oa()   :803-845   the OA entries of a whole batch, built on the GPU (pv3_format_oa)
 *   realigned_range_tag()    :848-860   (location, location + length) pairs as the B:I array the XQ / XF style tags hold
 *   record_index()           :863-868   value of the record index tag, when it fits an int32
 *   project_mate()           :880-906   RNEXT remap + PNEXT projection (device_index::project_position) + YR / YP originals
 *
 * Templates over the caller's types: a tag dictionary is anything map-like keyed by the 16-bit tag value seqan3 uses
 * (first char * 256 + second char; seqan3::sam_tag_dictionary is std::map <uint16_t, variant>), a range has .location and
 * .length. The reference's own types fit without adaptation.
 */
#ifndef PANVC3_B200_RECORD_ANNOTATOR_HH
#define PANVC3_B200_RECORD_ANNOTATOR_HH

#include <algorithm>
#include <array>
#include <cstdint>
#include <limits>
#include <optional>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

#include "../panvc3_b200.h"

namespace panvc3_b200 {

	typedef std::uint16_t sam_tag_type;

	constexpr sam_tag_type make_sam_tag(char const c0, char const c1)
	{
		return sam_tag_type((unsigned char) c0) * 256u + sam_tag_type((unsigned char) c1);
	}

	// project_alignments.cc:53-87, sorted (upper case ASCII, so the source order already is).
	inline constexpr std::array <sam_tag_type, 25> preserved_sam_tags{
		make_sam_tag('A', 'M'), make_sam_tag('A', 'S'), make_sam_tag('B', 'C'), make_sam_tag('B', 'Z'), make_sam_tag('C', 'B'),
		make_sam_tag('C', 'O'), make_sam_tag('C', 'R'), make_sam_tag('C', 'S'), make_sam_tag('C', 'T'), make_sam_tag('C', 'Y'),
		make_sam_tag('E', '2'), make_sam_tag('F', 'Z'), make_sam_tag('L', 'B'), make_sam_tag('M', 'I'), make_sam_tag('M', 'L'),
		make_sam_tag('M', 'M'), make_sam_tag('O', 'A'), make_sam_tag('O', 'Q'), make_sam_tag('O', 'X'), make_sam_tag('P', 'G'),
		make_sam_tag('P', 'U'), make_sam_tag('Q', 'T'), make_sam_tag('Q', 'X'), make_sam_tag('R', 'X'), make_sam_tag('T', 'S')
	};

	inline bool is_preserved_tag(sam_tag_type const tag)
	{
		return std::binary_search(preserved_sam_tags.begin(), preserved_sam_tags.end(), tag);
	}


	// :774-800. additional_preserved must be sorted (the reference sorts it once at start-up). removed_counts: map-like tag → count.
	template <typename t_tags, typename t_additional, typename t_counts>
	void filter_tags(t_tags &tags, t_additional const &additional_preserved, t_counts &removed_counts)
	{
		auto it(tags.begin());
		while (it != tags.end())
		{
			auto const tag(it->first);
			if (is_preserved_tag(tag) || std::binary_search(additional_preserved.begin(), additional_preserved.end(), tag))
			{
				++it;
				continue;
			}
			it = tags.erase(it);
			++removed_counts[tag];
		}
	}


	// :848-860
	template <typename t_ranges>
	std::vector <std::uint32_t> realigned_range_tag(t_ranges const &ranges)
	{
		std::vector <std::uint32_t> retval;
		retval.reserve(2 * ranges.size());
		for (auto const &range : ranges)
		{
			retval.push_back(std::uint32_t(range.location));
			retval.push_back(std::uint32_t(range.location + range.length));
		}
		return retval;
	}


	// :863-868 — the tag is only written when the value fits.
	inline std::optional <std::int32_t> record_index(std::uint64_t const last_rec_idx, std::uint64_t const valid_records)
	{
		auto const rec_idx(last_rec_idx - valid_records + 1);
		if (rec_idx <= std::uint64_t(std::numeric_limits <std::int32_t>::max()))
			return std::int32_t(rec_idx);
		return std::nullopt;
	}


	// The OA entries of a batch (:803-845). Fill the per-record vectors (the CIGAR arrays are the ones already built for
	// pv3_batch_in), call format(); entry(i) is what the reference appends to record i's OA tag.
	class oa_batch
	{
	public:
		std::vector <std::uint32_t>	ref_id;
		std::vector <std::uint64_t>	pos;
		std::vector <std::uint16_t>	flag;
		std::vector <std::uint8_t>	mapq;
		std::vector <std::int32_t>	nm;
		std::vector <std::uint8_t>	has_nm;

	private:
		std::vector <std::uint64_t>	m_name_off{0};
		std::string					m_names;
		std::vector <std::uint64_t>	m_text_off;
		std::string					m_text;

	public:
		void set_reference_names(std::vector <std::string> const &names)
		{
			m_name_off.assign(1, 0);
			m_names.clear();
			for (auto const &name : names) { m_names += name; m_name_off.push_back(m_names.size()); }
		}

		void clear() { ref_id.clear(); pos.clear(); flag.clear(); mapq.clear(); nm.clear(); has_nm.clear(); }

		void push_back(std::uint32_t ref_id_, std::uint64_t pos_, std::uint16_t flag_, std::uint8_t mapq_, std::optional <std::int32_t> nm_)
		{
			ref_id.push_back(ref_id_); pos.push_back(pos_); flag.push_back(flag_); mapq.push_back(mapq_);
			nm.push_back(nm_.value_or(0)); has_nm.push_back(nm_.has_value());
		}

		// cig_off has size() + 1 entries.
		void format(pv3_ctx *ctx, std::uint64_t const *cig_off, std::uint32_t const *cig)
		{
			pv3_oa_in in{};
			in.n = ref_id.size();
			in.ref_id = ref_id.data(); in.pos = pos.data(); in.flag = flag.data(); in.mapq = mapq.data();
			in.nm = nm.data(); in.has_nm = has_nm.data();
			in.cig_off = cig_off; in.cig = cig;
			in.n_names = std::uint32_t(m_name_off.size() - 1); in.name_off = m_name_off.data(); in.names = m_names.data();
			m_text_off.assign(in.n + 1, 0);
			std::uint64_t used{};
			int st(pv3_format_oa(ctx, &in, m_text_off.data(), nullptr, 0, &used));
			if (PV3_OK != st && PV3_ERR_CAPACITY != st) throw std::runtime_error(pv3_last_error(ctx));
			m_text.resize(used);
			st = pv3_format_oa(ctx, &in, m_text_off.data(), m_text.data(), m_text.size(), &used);
			if (PV3_OK != st) throw std::runtime_error(pv3_last_error(ctx));
		}

		std::size_t size() const { return ref_id.size(); }
		std::string_view entry(std::size_t const i) const { return std::string_view(m_text).substr(m_text_off[i], m_text_off[i + 1] - m_text_off[i]); }
	};


	// :880-906 for one record. t_index: device_index (alignment_projector.hh), with src_rows / dst_rows the device row ids of
	// the reference's src_seq_entries / dst_seq_entries (indexed by reference id), or anything with project_position(src, dst, pos).
	struct mate_projection
	{
		std::uint32_t	dst_mate_ref_id{};
		std::uint64_t	dst_mate_position{};
		bool			has_position{};
	};

	template <typename t_index, typename t_ref_id_mapping, typename t_rows>
	mate_projection project_mate(
		t_index &index,
		t_ref_id_mapping const &ref_id_mapping,
		t_rows const &src_rows,
		t_rows const &dst_rows,
		std::uint32_t const mate_ref_id,
		std::optional <std::uint64_t> const mate_position
	)
	{
		mate_projection retval;
		retval.dst_mate_ref_id = ref_id_mapping[mate_ref_id];
		if (mate_position)
		{
			retval.dst_mate_position = index.project_position(src_rows[mate_ref_id], dst_rows[retval.dst_mate_ref_id], *mate_position);
			retval.has_position = true;
		}
		return retval;
	}
}

#endif
