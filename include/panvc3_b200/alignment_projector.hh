/*
 * panvc3_b200 — source-compatible C++ adapter over the C-ABI (include/panvc3_b200.h).
 *
 * Mirrors the reference's public interface for the hot path so that project_task::process
 * (project-alignments/project_alignments.cc:721-732, :736, :859-861, :900, :916) compiles unchanged against it:
 *
 *   panvc3::alignment_projector            include/panvc3/alignment_projector.hh:33-75
 *       reset(), project_alignment(...), alignment(), realigned_reference_ranges(), realigned_query_ranges()
 *   panvc3::alignment_projector_delegate   include/panvc3/alignment_projector.hh:25-30
 *   msa_index::sequence_entry::project_position   include/panvc3/msa_index.hh:253-258  → device_index::project_position
 *
 * The class is a template over a small traits type that says how to read the caller's element types, so the same code
 * serves (a) the reference's own types — seqan3::dna5 / phred42 / cigar, panvc3::msa_index::sequence_entry,
 * panvc3::range — through `seqan3_traits` (defined when the seqan3 headers are on the include path), and (b) plain
 * structs in this repository's tests (`plain_traits`).  A call is a batch of one; the batched entry
 * (`project_batch`) is what a pipeline should use.
 *
 * Ownership as in the reference: the caller owns all inputs; the projector owns its output vectors, valid until the
 * next reset()/project_alignment().  The gap rows and target contigs are registered with a `device_index` on first
 * use (keyed by object address, like the reference's raw back-pointers, msa_index.hh:195-199) and stay resident on the GPU.
 * Errors: where the reference aborts (libbio_fail / assert) or has UB, `projection_error` is thrown with the same message.
 */
#ifndef PANVC3_B200_ALIGNMENT_PROJECTOR_HH
#define PANVC3_B200_ALIGNMENT_PROJECTOR_HH

#include <cstdint>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../panvc3_b200.h"

namespace panvc3_b200 {

	struct projection_error : public std::runtime_error
	{
		std::uint32_t flags{};
		projection_error(std::uint32_t flags_, std::string const &msg): std::runtime_error(msg), flags(flags_) {}
	};

	inline std::string describe_flags(std::uint32_t const flags)
	{
		std::string msg;
		auto add = [&msg](char const *s){ if (!msg.empty()) msg += "; "; msg += s; };
		if (flags & PV3_FLAG_BAD_OP)      add("Unexpected CIGAR operation");                      // rewrite_cigar.hh:150
		if (flags & PV3_FLAG_POS_OOB)     add("source position past the end of the sequence");
		if (flags & PV3_FLAG_QUERY_OOB)   add("query_pos < query_seq.size() failed");              // rewrite_cigar.hh:102
		if (flags & PV3_FLAG_REF_OOB)     add("excess < dst_seq.size() failed");                   // rewrite_cigar.hh:103
		if (flags & PV3_FLAG_TRACE_UB)    add("traceback left the matrix (reference UB with gap_open < 0)");
		if (flags & PV3_FLAG_EMPTY_CIGAR) add("empty CIGAR");
		if (flags & PV3_FLAG_BAD_QUAL)    add("QUAL length differs from SEQ length");
		if (flags & PV3_FLAG_BAD_ID)      add("unknown sequence entry or contig");
		return msg;
	}


	// Owns the GPU context; maps the caller's sequence entries / contigs to device row and contig ids.
	class device_index
	{
		pv3_ctx												*m_ctx{};
		std::unordered_map <void const *, std::uint32_t>	m_rows;
		std::unordered_map <void const *, std::uint32_t>	m_refs;

		void check(int const st, char const *what) const
		{
			if (st < 0)
				throw std::runtime_error(std::string(what) + ": " + pv3_last_error(m_ctx));
		}

	public:
		explicit device_index(int const device = 0)
		{
			if (PV3_OK != pv3_create(&m_ctx, device))
				throw std::runtime_error(std::string("pv3_create: ") + pv3_last_error(nullptr));
		}

		~device_index() { pv3_destroy(m_ctx); }
		device_index(device_index const &) = delete;
		device_index &operator=(device_index const &) = delete;

		pv3_ctx *context() const { return m_ctx; }

		// gap_bit(i) must return 1 where column i of the aligned row is a gap (msa_index.hh:22-32).
		template <typename t_gap_bit>
		std::uint32_t register_row(void const *key, std::uint64_t const n_bits, t_gap_bit &&gap_bit)
		{
			auto const it(m_rows.find(key));
			if (m_rows.end() != it)
				return it->second;
			std::vector <std::uint64_t> words((n_bits + 63) / 64, 0);
			for (std::uint64_t i(0); i < n_bits; ++i)
				if (gap_bit(i)) words[i / 64] |= std::uint64_t(1) << (i % 64);
			int const id(pv3_index_add_row(m_ctx, words.data(), n_bits));
			check(id, "pv3_index_add_row");
			m_rows.emplace(key, std::uint32_t(id));
			return std::uint32_t(id);
		}

		// Words already expanded (e.g. rrr_vector<15>::get_int(64 * w, 64), rrr_vector_15.hpp:320-360).
		std::uint32_t register_row_words(void const *key, std::uint64_t const *words, std::uint64_t const n_bits)
		{
			auto const it(m_rows.find(key));
			if (m_rows.end() != it)
				return it->second;
			int const id(pv3_index_add_row(m_ctx, words, n_bits));
			check(id, "pv3_index_add_row");
			m_rows.emplace(key, std::uint32_t(id));
			return std::uint32_t(id);
		}

		std::uint32_t register_reference(std::vector <char> const &ref_seq)
		{
			auto const it(m_refs.find(&ref_seq));
			if (m_refs.end() != it)
				return it->second;
			int const id(pv3_index_add_ref(m_ctx, ref_seq.data(), ref_seq.size()));
			check(id, "pv3_index_add_ref");
			m_refs.emplace(&ref_seq, std::uint32_t(id));
			return std::uint32_t(id);
		}

		// Forget a contig buffer that is about to be reused for another contig (sequence_buffer_store recycles buffers).
		void forget_reference(std::vector <char> const &ref_seq) { m_refs.erase(&ref_seq); }

		void commit() { check(pv3_index_commit(m_ctx), "pv3_index_commit"); }

		// msa_index::sequence_entry::project_position (msa_index.hh:253-258)
		std::uint64_t project_position(std::uint32_t const src_row, std::uint32_t const dst_row, std::uint64_t const pos)
		{
			commit();
			std::uint64_t out{};
			check(pv3_project_positions(m_ctx, 1, &src_row, &dst_row, &pos, &out), "pv3_project_positions");
			return out;
		}
	};
}



namespace panvc3_b200 {

	// Traits for plain test types. A traits type provides:
	//   sequence_entry_type, sequence_alphabet, quality_alphabet, cigar_type, range_type
	//   n_bits(entry), gap_bit(entry, i)            the gap row (1 = gap)
	//   seq_rank(x) / qual_rank(x)                   dna5 rank 0..4 (A C G N T) / phred rank 0..41
	//   cigar_count(c), cigar_op_char(c), make_cigar(count, op_char), make_range(location, length)
	struct plain_cigar { std::uint32_t count; char op; };
	struct plain_range { std::size_t location; std::size_t length; };
	struct plain_sequence_entry { std::vector <bool> gap_positions; };   // true = gap

	struct plain_traits
	{
		typedef plain_sequence_entry	sequence_entry_type;
		typedef std::uint8_t			sequence_alphabet;		// dna5 rank
		typedef std::uint8_t			quality_alphabet;		// phred rank
		typedef plain_cigar				cigar_type;
		typedef plain_range				range_type;

		static std::uint64_t n_bits(sequence_entry_type const &e) { return e.gap_positions.size(); }
		static bool gap_bit(sequence_entry_type const &e, std::uint64_t const i) { return e.gap_positions[i]; }
		static std::uint8_t seq_rank(sequence_alphabet const x) { return x; }
		static std::uint8_t qual_rank(quality_alphabet const x) { return x; }
		static std::uint32_t cigar_count(cigar_type const &c) { return c.count; }
		static char cigar_op_char(cigar_type const &c) { return c.op; }
		static cigar_type make_cigar(std::uint32_t const count, char const op) { return cigar_type{count, op}; }
		static range_type make_range(std::size_t const loc, std::size_t const len) { return range_type{loc, len}; }
	};


	template <typename t_traits> class basic_alignment_projector; // Fwd.

	template <typename t_traits>
	struct basic_alignment_projector_delegate
	{
		virtual ~basic_alignment_projector_delegate() {}
		// On the GPU all runs of a record are realigned inside one call, so both callbacks fire once per realigned run
		// after the call returns; they cannot bracket the DP (documented deviation).
		virtual void alignment_projector_begin_realignment(basic_alignment_projector <t_traits> const &) = 0;
		virtual void alignment_projector_end_realignment(basic_alignment_projector <t_traits> const &) = 0;
	};


	template <typename t_traits>
	class basic_alignment_projector
	{
	public:
		typedef typename t_traits::sequence_entry_type	sequence_entry_type;
		typedef typename t_traits::sequence_alphabet	sequence_alphabet;
		typedef typename t_traits::quality_alphabet		quality_alphabet;
		typedef typename t_traits::cigar_type			cigar_type;
		typedef typename t_traits::range_type			range_type;
		typedef std::vector <cigar_type>				cigar_vector;
		typedef std::vector <range_type>				range_vector;
		typedef basic_alignment_projector_delegate <t_traits>	delegate_type;

	protected:
		device_index					*m_index{};
		delegate_type					*m_delegate{};
		cigar_vector					m_cigar_realigned;
		range_vector					m_realigned_reference_ranges;
		range_vector					m_realigned_query_ranges;
		// reusable host buffers (the reference reuses its vectors across records, alignment_projector.hh:41-47)
		std::vector <std::uint8_t>		m_seq, m_qual;
		std::vector <std::uint32_t>		m_cig_in, m_cig_out;
		std::vector <std::uint64_t>		m_rr;

		static std::uint32_t bam_op(char const c)
		{
			char const ops[] = "MIDNSHP=X";
			for (std::uint32_t i(0); i < 9; ++i) if (ops[i] == c) return i;
			return 15; // flagged by the kernel (PV3_FLAG_BAD_OP), like libbio_fail in rewrite_cigar.hh:150
		}

	public:
		explicit basic_alignment_projector(device_index &index): m_index(&index) {}
		basic_alignment_projector(device_index &index, delegate_type &delegate): m_index(&index), m_delegate(&delegate) {}

		void reset()                                                        // alignment_projector.cc:19-27
		{
			m_cigar_realigned.clear();
			m_realigned_reference_ranges.clear();
			m_realigned_query_ranges.clear();
		}

		std::uint32_t register_entry(sequence_entry_type const &entry)
		{
			return m_index->register_row(&entry, t_traits::n_bits(entry), [&entry](std::uint64_t const i){ return t_traits::gap_bit(entry, i); });
		}

		// Same parameter list as panvc3::alignment_projector::project_alignment (alignment_projector.hh:58-68).
		std::size_t project_alignment(
			std::size_t const src_pos,
			sequence_entry_type const &src_seq_entry,
			sequence_entry_type const &dst_seq_entry,
			std::vector <char> const &ref_seq,
			std::vector <sequence_alphabet> const &query_seq,
			std::vector <cigar_type> const &cigar_seq,
			std::vector <quality_alphabet> const &base_qualities,	// Pass empty vector to not use base qualities.
			std::int32_t const gap_opening_cost,
			std::int32_t const gap_extension_cost
		)
		{
			std::uint32_t const src_row(register_entry(src_seq_entry));
			std::uint32_t const dst_row(register_entry(dst_seq_entry));
			std::uint32_t const dst_ref(m_index->register_reference(ref_seq));
			m_index->commit();
			pv3_ctx *ctx(m_index->context());

			m_seq.resize(query_seq.size());
			for (std::size_t i(0); i < query_seq.size(); ++i) m_seq[i] = t_traits::seq_rank(query_seq[i]);
			m_qual.resize(base_qualities.size());
			for (std::size_t i(0); i < base_qualities.size(); ++i) m_qual[i] = t_traits::qual_rank(base_qualities[i]);
			m_cig_in.resize(cigar_seq.size());
			for (std::size_t i(0); i < cigar_seq.size(); ++i)
				m_cig_in[i] = (t_traits::cigar_count(cigar_seq[i]) << 4) | bam_op(t_traits::cigar_op_char(cigar_seq[i]));

			std::uint64_t const src_pos_(src_pos);
			std::uint64_t const seq_off[2]{0, m_seq.size()}, qual_off[2]{0, m_qual.size()}, cig_off[2]{0, m_cig_in.size()};
			pv3_batch_in in{};
			in.n = 1; in.src_row = &src_row; in.dst_row = &dst_row; in.dst_ref = &dst_ref; in.src_pos = &src_pos_;
			in.seq_off = seq_off; in.seq = m_seq.data(); in.qual_off = qual_off; in.qual = m_qual.data();
			in.cig_off = cig_off; in.cig = m_cig_in.data();

			if (pv3_set_params(ctx, gap_opening_cost, gap_extension_cost) < 0)
				throw std::runtime_error(pv3_last_error(ctx));

			std::uint64_t dst_pos{}, cig_o{}, rr_o{};
			std::uint32_t flags{}, cig_l{}, rr_l{};
			if (m_cig_out.size() < 64) m_cig_out.resize(64);
			if (m_rr.size() < 4 * 8) m_rr.resize(4 * 8);
			for (;;)
			{
				pv3_batch_out out{};
				out.dst_pos = &dst_pos; out.flags = &flags; out.cig_off = &cig_o; out.cig_len = &cig_l;
				out.cig = m_cig_out.data(); out.cig_cap = m_cig_out.size();
				out.rr_off = &rr_o; out.rr_len = &rr_l; out.rr = m_rr.data(); out.rr_cap = m_rr.size() / 4;
				int const st(pv3_project(ctx, &in, &out));
				if (PV3_ERR_CAPACITY == st)
				{
					m_cig_out.resize(2 * out.cig_used + 64);
					m_rr.resize(4 * (2 * out.rr_used + 8));
					continue;
				}
				if (st < 0) throw std::runtime_error(pv3_last_error(ctx));
				break;
			}
			if (flags)
				throw projection_error(flags, describe_flags(flags));

			char const ops[] = "MIDNSHP=X";
			m_cigar_realigned.clear();
			for (std::uint32_t i(0); i < cig_l; ++i)
			{
				std::uint32_t const v(m_cig_out[cig_o + i]);
				m_cigar_realigned.push_back(t_traits::make_cigar(v >> 4, ops[v & 0xf]));
			}
			for (std::uint32_t i(0); i < rr_l; ++i)                         // alignment_projector.cc:69-73
			{
				std::uint64_t const *rr(m_rr.data() + 4 * (rr_o + i));
				m_realigned_reference_ranges.push_back(t_traits::make_range(rr[0], rr[1]));
				m_realigned_query_ranges.push_back(t_traits::make_range(rr[2], rr[3]));
				if (m_delegate)
				{
					m_delegate->alignment_projector_begin_realignment(*this);
					m_delegate->alignment_projector_end_realignment(*this);
				}
			}
			return dst_pos;
		}

		cigar_vector const &alignment() const { return m_cigar_realigned; }
		range_vector const &realigned_reference_ranges() const { return m_realigned_reference_ranges; }
		range_vector const &realigned_query_ranges() const { return m_realigned_query_ranges; }
	};

	typedef basic_alignment_projector <plain_traits> plain_alignment_projector;
}


// The reference's own types. Available when the reference's include paths (include/, lib/seqan3/include,
// lib/sdsl-lite/include, ...) are given to the compiler, i.e. inside the reference's build.
#if defined(PANVC3_B200_WITH_SEQAN3) || (__has_include(<seqan3/alphabet/cigar/cigar.hpp>) && __has_include(<panvc3/msa_index.hh>))
#include <panvc3/msa_index.hh>
#include <panvc3/range.hh>
#include <seqan3/alphabet/cigar/cigar.hpp>
#include <seqan3/alphabet/nucleotide/dna5.hpp>
#include <seqan3/alphabet/quality/phred42.hpp>

namespace panvc3_b200 {

	struct seqan3_traits
	{
		typedef panvc3::msa_index::sequence_entry	sequence_entry_type;
		typedef seqan3::dna5						sequence_alphabet;
		typedef seqan3::phred42						quality_alphabet;
		typedef seqan3::cigar						cigar_type;
		typedef panvc3::range						range_type;

		static std::uint64_t n_bits(sequence_entry_type const &e) { return e.gap_positions.size(); }
		static bool gap_bit(sequence_entry_type const &e, std::uint64_t const i) { return e.gap_positions[i]; }	// rrr access
		static std::uint8_t seq_rank(sequence_alphabet const x) { return x.to_rank(); }
		static std::uint8_t qual_rank(quality_alphabet const x) { return x.to_rank(); }
		static std::uint32_t cigar_count(cigar_type const &c) { using seqan3::get; return get <0>(c); }
		static char cigar_op_char(cigar_type const &c) { using seqan3::get; return get <1>(c).to_char(); }
		static cigar_type make_cigar(std::uint32_t const count, char const op)
		{
			cigar_type cc; seqan3::cigar::operation oo; oo.assign_char(op);
			cc = oo; cc = count;
			return cc;
		}
		static range_type make_range(std::size_t const loc, std::size_t const len) { return range_type(loc, len); }
	};

	// Drop-in names: `using panvc3::alignment_projector = panvc3_b200::alignment_projector` in the reference's build.
	typedef basic_alignment_projector <seqan3_traits>			alignment_projector;
	typedef basic_alignment_projector_delegate <seqan3_traits>	alignment_projector_delegate;
}
#endif

#endif
