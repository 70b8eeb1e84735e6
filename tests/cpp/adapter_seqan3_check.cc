// Compile-only check (build container only: needs /root/reference): the adapter instantiates with the reference's own
// types — seqan3::dna5 / phred42 / cigar, panvc3::msa_index::sequence_entry, panvc3::range — and exposes the same
// project_alignment signature as panvc3::alignment_projector (include/panvc3/alignment_projector.hh:58-68).
#include "../../include/panvc3_b200/alignment_projector.hh"

#include <type_traits>

template class panvc3_b200::basic_alignment_projector <panvc3_b200::seqan3_traits>;

using projector = panvc3_b200::alignment_projector;
static_assert(std::is_same_v <
	decltype(&projector::project_alignment),
	std::size_t (projector::*)(std::size_t, panvc3::msa_index::sequence_entry const &, panvc3::msa_index::sequence_entry const &,
		std::vector <char> const &, std::vector <seqan3::dna5> const &, std::vector <seqan3::cigar> const &,
		std::vector <seqan3::phred42> const &, std::int32_t, std::int32_t)>);
static_assert(std::is_same_v <decltype(std::declval <projector const &>().alignment()), std::vector <seqan3::cigar> const &>);
static_assert(std::is_same_v <decltype(std::declval <projector const &>().realigned_reference_ranges()), std::vector <panvc3::range> const &>);

int main() { return 0; }
