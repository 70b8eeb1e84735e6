// Host-side checks of include/panvc3_b200/record_annotator.hh (no GPU): tag filter, range tags, record index.
// With "oa" as argv[1] and a GPU present: reads "ref_id pos flag mapq nm|. cigar" lines, prints the OA entries.
#include <cstdio>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <panvc3_b200/record_annotator.hh>

namespace p = panvc3_b200;

struct range { std::size_t location; std::size_t length; };

static int check(bool ok, char const *what) { if (!ok) std::fprintf(stderr, "FAILED: %s\n", what); return ok ? 0 : 1; }

int main(int argc, char **argv)
{
	if (argc > 1 && std::string(argv[1]) == "oa")
	{
		pv3_ctx *ctx{};
		if (pv3_create(&ctx, 0) != PV3_OK) { std::fprintf(stderr, "%s\n", pv3_last_error(nullptr)); return 2; }
		p::oa_batch batch;
		batch.set_reference_names({"1", "chrB", "a_longer_name"});
		std::vector <std::uint64_t> cig_off{0};
		std::vector <std::uint32_t> cig;
		std::string line;
		while (std::getline(std::cin, line))
		{
			std::istringstream is(line);
			std::uint32_t ref_id, flag, mapq; std::uint64_t pos; std::string nm, cigar;
			is >> ref_id >> pos >> flag >> mapq >> nm >> cigar;
			batch.push_back(ref_id, pos, std::uint16_t(flag), std::uint8_t(mapq), nm == "." ? std::nullopt : std::optional <std::int32_t>(std::stoi(nm)));
			std::uint32_t count{};
			for (char const c : cigar)
			{
				if ('0' <= c && c <= '9') { count = 10 * count + (c - '0'); continue; }
				cig.push_back((count << 4) | std::uint32_t(std::string("MIDNSHP=X").find(c)));
				count = 0;
			}
			cig_off.push_back(cig.size());
		}
		batch.format(ctx, cig_off.data(), cig.data());
		for (std::size_t i(0); i < batch.size(); ++i) std::cout << batch.entry(i) << '\n';
		pv3_destroy(ctx);
		return 0;
	}

	int failed{};
	{
		std::map <p::sam_tag_type, int> tags{
			{p::make_sam_tag('N', 'M'), 1}, {p::make_sam_tag('A', 'S'), 2}, {p::make_sam_tag('X', 'S'), 3}, {p::make_sam_tag('O', 'A'), 4},
			{p::make_sam_tag('Y', 'T'), 5}, {p::make_sam_tag('M', 'D'), 6}, {p::make_sam_tag('Z', 'Z'), 7}, {p::make_sam_tag('T', 'S'), 8}
		};
		std::vector <p::sam_tag_type> additional{p::make_sam_tag('X', 'S'), p::make_sam_tag('Z', 'Z')};
		std::sort(additional.begin(), additional.end());
		std::map <p::sam_tag_type, std::size_t> removed;
		p::filter_tags(tags, additional, removed);
		failed += check(tags.size() == 5 && tags.count(p::make_sam_tag('A', 'S')) && tags.count(p::make_sam_tag('O', 'A'))
			&& tags.count(p::make_sam_tag('T', 'S')) && tags.count(p::make_sam_tag('X', 'S')) && tags.count(p::make_sam_tag('Z', 'Z')), "kept tags");
		failed += check(removed.size() == 3 && removed[p::make_sam_tag('N', 'M')] == 1 && removed[p::make_sam_tag('Y', 'T')] == 1
			&& removed[p::make_sam_tag('M', 'D')] == 1, "removed tag counts");
		p::filter_tags(tags, additional, removed);
		failed += check(tags.size() == 5 && removed.size() == 3, "filter is idempotent");
		failed += check(std::is_sorted(p::preserved_sam_tags.begin(), p::preserved_sam_tags.end()), "preserved set sorted");
	}
	{
		std::vector <range> ranges{{10, 5}, {100, 0}, {4000000000u, 7}};
		auto const tag(p::realigned_range_tag(ranges));
		failed += check(tag == std::vector <std::uint32_t>{10, 15, 100, 100, 4000000000u, 4000000007u}, "range tag");
		failed += check(p::realigned_range_tag(std::vector <range>{}).empty(), "empty range tag");
	}
	{
		failed += check(p::record_index(10, 3) == std::optional <std::int32_t>(8), "record index");
		failed += check(!p::record_index(std::uint64_t(1) << 31, 1).has_value(), "record index past int32");
		failed += check(p::record_index((std::uint64_t(1) << 31) - 1, 1) == std::optional <std::int32_t>(2147483647), "record index at int32 max");
	}
	{
		struct fake_index { std::uint64_t project_position(std::uint32_t s, std::uint32_t d, std::uint64_t pos) { return 1000 * s + 100 * d + pos; } } index;
		std::vector <std::uint32_t> mapping{2, 0, 1}, src_rows{5, 6, 7}, dst_rows{1, 2, 3};
		auto const m(p::project_mate(index, mapping, src_rows, dst_rows, 1, std::optional <std::uint64_t>(42)));
		failed += check(m.dst_mate_ref_id == 0 && m.has_position && m.dst_mate_position == 6000 + 100 + 42, "mate projection");
		auto const m2(p::project_mate(index, mapping, src_rows, dst_rows, 0, std::nullopt));
		failed += check(m2.dst_mate_ref_id == 2 && !m2.has_position, "mate without position");
	}
	std::puts(failed ? "FAILED" : "ok");
	return failed ? 1 : 0;
}
