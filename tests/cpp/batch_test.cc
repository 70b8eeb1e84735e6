// Test driver for panvc3_b200::basic_batch_projector (include/panvc3_b200/alignment_projector.hh) with plain types.
// Same input lines and output format as adapter_test.cc, but every record of the input is packed into ONE batch
// (push_back into the pinned arena, one flush), optionally over several devices:
//   batch_test [device,device,...]     e.g. "0" (default), "0,1", or "0,0" (two shards on one GPU)
// All lines must carry the same gap costs (the costs are per flush, as they are per CLI run in the reference).
// After the records: "MATE <projected position>" for every record (its own start position projected again as a mate).
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

template <typename t_index>
static int run(t_index &index)
{
	pb::basic_batch_projector <pb::plain_traits, t_index> batch(index);
	std::vector <std::unique_ptr <pb::plain_sequence_entry>> entries;
	std::vector <std::unique_ptr <std::vector <char>>> contigs;
	int go(0), ge(-1);
	std::string line;
	while (std::getline(std::cin, line))
	{
		if (line.empty()) continue;
		std::vector <std::string> f;
		std::stringstream ss(line);
		std::string tok;
		while (std::getline(ss, tok, '\t')) f.push_back(tok);
		if (f.size() < 8) { std::cout << "BAD INPUT\n"; return 1; }
		auto make_entry = [&entries](std::string const &row) -> pb::plain_sequence_entry const & {
			auto e(std::make_unique <pb::plain_sequence_entry>());
			for (char c : row) e->gap_positions.push_back('-' == c);
			entries.emplace_back(std::move(e));
			return *entries.back();
		};
		auto const src_row(batch.register_entry(make_entry(f[0])));
		auto const dst_row(batch.register_entry(make_entry(f[1])));
		contigs.emplace_back(std::make_unique <std::vector <char>>());
		for (char c : f[1]) if ('-' != c) contigs.back()->push_back(c);
		auto const dst_ref(batch.register_reference(*contigs.back()));
		std::vector <std::uint8_t> query, qual;
		for (char c : f[2]) query.push_back(rank_of(c));
		if ("1" == f[5]) qual.assign(query.size(), 41);
		std::vector <pb::plain_cigar> cigar;
		{
			std::uint32_t num(0);
			for (char c : f[4]) { if ('0' <= c && c <= '9') num = 10 * num + (c - '0'); else { cigar.push_back({num, c}); num = 0; } }
		}
		go = std::stoi(f[6]); ge = std::stoi(f[7]);
		batch.push_back(std::stoull(f[3]), src_row, dst_row, dst_ref, query, cigar, qual);
		batch.set_mate(src_row, std::stoull(f[3]));
	}
	batch.flush(go, ge);
	std::vector <pb::plain_cigar> aln;
	std::vector <pb::plain_range> rr, qr;
	for (std::size_t i(0); i < batch.size(); ++i)
	{
		if (batch.flags(i)) { std::cout << "ERROR " << batch.flags(i) << '\n'; continue; }
		batch.alignment(i, aln);
		batch.realigned_ranges(i, rr, qr);
		std::cout << batch.position(i) << '\t';
		for (auto const &c : aln) std::cout << c.count << c.op;
		std::cout << '\t';
		for (auto const &r : rr) std::cout << r.location << ':' << r.length << ',';
		std::cout << '\t';
		for (auto const &r : qr) std::cout << r.location << ':' << r.length << ',';
		std::cout << '\n';
	}
	for (std::size_t i(0); i < batch.size(); ++i) std::cout << "MATE " << batch.mate_position(i) << '\n';
	return 0;
}

int main(int argc, char **argv)
{
	std::vector <int> devices;
	if (argc > 1)
	{
		std::stringstream ss(argv[1]);
		std::string tok;
		while (std::getline(ss, tok, ',')) devices.push_back(std::stoi(tok));
	}
	try
	{
		if (devices.size() > 1)
		{
			pb::multi_device_index index(int(devices.size()), devices.data());
			return run(index);
		}
		pb::device_index index(devices.empty() ? 0 : devices[0]);
		return run(index);
	}
	catch (std::exception const &e) { std::cout << "EXCEPTION " << e.what() << '\n'; return 2; }
}
