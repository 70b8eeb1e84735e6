// Test driver for include/panvc3_b200/alignment_projector.hh with plain types (runs on the GPU box, no seqan3 needed).
// Reads cases from stdin, one per line, tab-separated:
//   src_row_aligned  dst_row_aligned  query(ACGTN)  src_pos  cigar  qual(0|1: all phred 41)  gap_open  gap_ext
// and prints: dst_pos \t cigar \t ref ranges (loc:len,...) \t query ranges      or "ERROR flags" when the projector throws.
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/panvc3_b200/alignment_projector.hh"

namespace pb = panvc3_b200;

static std::uint8_t rank_of(char c)
{
	switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 4; default: return 3; }
}

struct counting_delegate : public pb::basic_alignment_projector_delegate <pb::plain_traits>
{
	int begins{}, ends{};
	void alignment_projector_begin_realignment(pb::plain_alignment_projector const &) override { ++begins; }
	void alignment_projector_end_realignment(pb::plain_alignment_projector const &) override { ++ends; }
};

int main()
{
	pb::device_index index(0);
	counting_delegate delegate;
	pb::plain_alignment_projector projector(index, delegate);
	// entries and contigs must outlive the projector calls (keyed by address, like the reference's back-pointers)
	std::vector <std::unique_ptr <pb::plain_sequence_entry>> entries;
	std::vector <std::unique_ptr <std::vector <char>>> contigs;

	std::string line;
	while (std::getline(std::cin, line))
	{
		if (line.empty()) continue;
		std::vector <std::string> f;
		std::stringstream ss(line);
		std::string tok;
		while (std::getline(ss, tok, '\t')) f.push_back(tok);
		if (f.size() < 8) { std::cout << "BAD INPUT\n"; continue; }
		auto make_entry = [&entries](std::string const &row) -> pb::plain_sequence_entry const & {
			auto e(std::make_unique <pb::plain_sequence_entry>());
			for (char c : row) e->gap_positions.push_back('-' == c);
			entries.emplace_back(std::move(e));
			return *entries.back();
		};
		auto const &src(make_entry(f[0]));
		auto const &dst(make_entry(f[1]));
		contigs.emplace_back(std::make_unique <std::vector <char>>());
		for (char c : f[1]) if ('-' != c) contigs.back()->push_back(c);
		std::vector <std::uint8_t> query, qual;
		for (char c : f[2]) query.push_back(rank_of(c));
		if ("1" == f[5]) qual.assign(query.size(), 41);
		std::vector <pb::plain_cigar> cigar;
		{
			std::uint32_t num(0);
			for (char c : f[4]) { if ('0' <= c && c <= '9') num = 10 * num + (c - '0'); else { cigar.push_back({num, c}); num = 0; } }
		}
		try
		{
			projector.reset();
			auto const pos(projector.project_alignment(std::stoull(f[3]), src, dst, *contigs.back(), query, cigar, qual, std::stoi(f[6]), std::stoi(f[7])));
			std::cout << pos << '\t';
			for (auto const &c : projector.alignment()) std::cout << c.count << c.op;
			std::cout << '\t';
			for (auto const &r : projector.realigned_reference_ranges()) std::cout << r.location << ':' << r.length << ',';
			std::cout << '\t';
			for (auto const &r : projector.realigned_query_ranges()) std::cout << r.location << ':' << r.length << ',';
			std::cout << '\n';
		}
		catch (pb::projection_error const &err) { std::cout << "ERROR " << err.flags << '\n'; }
	}
	std::cout << "DELEGATE " << delegate.begins << ' ' << delegate.ends << '\n';
	return 0;
}
