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
 * structs in this repository's tests (`plain_traits`).  A call is a batch of one (three stream operations and a
 * synchronisation per record: compatibility, not speed); `basic_batch_projector` below is what a pipeline should use:
 * push_back() packs records into a pinned structure-of-arrays arena, flush() runs them as one pv3_project call, results
 * are read back in submission order.
 *
 * THREADING. A device_index (one pv3_ctx) and everything hanging off it — projectors, batch projectors — belong to ONE
 * submitting thread at a time; nothing here locks. The reference's should_process_tasks_in_parallel mode runs
 * project_task::process on a concurrent queue: give every worker thread its own device_index (the index is small next
 * to HBM) or funnel the records of all workers into one batch_projector owned by one thread.
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

		// Forget a contig buffer that is about to be reused for another contig (sequence_buffer_store recycles buffers):
		// the device copy is freed and its id recycled.
		void forget_reference(std::vector <char> const &ref_seq)
		{
			auto const it(m_refs.find(&ref_seq));
			if (m_refs.end() == it) return;
			check(pv3_index_remove_ref(m_ctx, int(it->second)), "pv3_index_remove_ref");
			m_refs.erase(it);
		}

		// Uploads what was registered since the last commit (rows and contigs have their own device allocations).
		void commit() { check(pv3_index_commit(m_ctx), "pv3_index_commit"); }

		void set_params(std::int32_t const gap_opening_cost, std::int32_t const gap_extension_cost)
		{
			check(pv3_set_params(m_ctx, gap_opening_cost, gap_extension_cost), "pv3_set_params");
		}

		int project(pv3_batch_in const &in, pv3_batch_out &out) { return pv3_project(m_ctx, &in, &out); }
		char const *last_error() const { return pv3_last_error(m_ctx); }
		static constexpr bool needs_offsets = false;       // pv3_project orders the arenas over the whole batch

		// Device time of the DP + traceback kernels of the last projection call, in milliseconds. The reference's CLI times
		// each realignment between the two delegate callbacks (project_alignments.cc:932-944); on the GPU all runs of a batch
		// are realigned together, so the callbacks cannot bracket them — a caller that reports "time spent realigning" adds
		// this figure per call instead.
		double last_realignment_ms() const
		{
			pv3_timing t{};
			return PV3_OK == pv3_get_timing(m_ctx, &t) ? double(t.ms_realign) : 0.0;
		}

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

	// Several GPUs behind one submitter (pv3_multi): every row / contig is registered on every device; a batch is cut into
	// contiguous shards, one per device, and comes back in submission order. Same interface as device_index.
	class multi_device_index
	{
		pv3_multi											*m_multi{};
		std::unordered_map <void const *, std::uint32_t>	m_rows;
		std::unordered_map <void const *, std::uint32_t>	m_refs;

		void check(int const st, char const *what, pv3_ctx *ctx) const
		{
			if (st < 0)
				throw std::runtime_error(std::string(what) + ": " + pv3_last_error(ctx));
		}

	public:
		explicit multi_device_index(int const n_devices, int const *devices = nullptr)
		{
			if (PV3_OK != pv3_multi_create(&m_multi, n_devices, devices))
				throw std::runtime_error(std::string("pv3_multi_create: ") + pv3_last_error(nullptr));
		}

		~multi_device_index() { pv3_multi_destroy(m_multi); }
		multi_device_index(multi_device_index const &) = delete;
		multi_device_index &operator=(multi_device_index const &) = delete;

		int devices() const { return pv3_multi_devices(m_multi); }
		pv3_ctx *context() const { return pv3_multi_ctx(m_multi, 0); }

		std::uint32_t register_row_words(void const *key, std::uint64_t const *words, std::uint64_t const n_bits)
		{
			auto const it(m_rows.find(key));
			if (m_rows.end() != it)
				return it->second;
			int id(-1);
			for (int d(0); d < devices(); ++d)
			{
				id = pv3_index_add_row(pv3_multi_ctx(m_multi, d), words, n_bits);
				check(id, "pv3_index_add_row", pv3_multi_ctx(m_multi, d));
			}
			m_rows.emplace(key, std::uint32_t(id));
			return std::uint32_t(id);
		}

		std::uint32_t register_reference(std::vector <char> const &ref_seq)
		{
			auto const it(m_refs.find(&ref_seq));
			if (m_refs.end() != it)
				return it->second;
			int id(-1);
			for (int d(0); d < devices(); ++d)
			{
				id = pv3_index_add_ref(pv3_multi_ctx(m_multi, d), ref_seq.data(), ref_seq.size());
				check(id, "pv3_index_add_ref", pv3_multi_ctx(m_multi, d));
			}
			m_refs.emplace(&ref_seq, std::uint32_t(id));
			return std::uint32_t(id);
		}

		void forget_reference(std::vector <char> const &ref_seq)
		{
			auto const it(m_refs.find(&ref_seq));
			if (m_refs.end() == it) return;
			for (int d(0); d < devices(); ++d)
				check(pv3_index_remove_ref(pv3_multi_ctx(m_multi, d), int(it->second)), "pv3_index_remove_ref", pv3_multi_ctx(m_multi, d));
			m_refs.erase(it);
		}

		void commit() { for (int d(0); d < devices(); ++d) check(pv3_index_commit(pv3_multi_ctx(m_multi, d)), "pv3_index_commit", pv3_multi_ctx(m_multi, d)); }

		void set_params(std::int32_t const gap_opening_cost, std::int32_t const gap_extension_cost)
		{
			for (int d(0); d < devices(); ++d) check(pv3_set_params(pv3_multi_ctx(m_multi, d), gap_opening_cost, gap_extension_cost), "pv3_set_params", pv3_multi_ctx(m_multi, d));
		}

		int project(pv3_batch_in const &in, pv3_batch_out &out) { return pv3_multi_project(m_multi, &in, &out); }
		char const *last_error() const { return pv3_multi_last_error(m_multi); }
		static constexpr bool needs_offsets = true;        // arenas are ordered inside every device's shard: offsets are absolute
	};


	// Pinned, growing array (pv3_host_alloc): the arena the batch projector packs into.
	template <typename t_value>
	class pinned_vector
	{
		t_value		*m_data{};
		std::size_t	m_size{}, m_cap{};

	public:
		pinned_vector() = default;
		pinned_vector(pinned_vector const &) = delete;
		pinned_vector &operator=(pinned_vector const &) = delete;
		~pinned_vector() { pv3_host_free(m_data); }

		void reserve(std::size_t const n)
		{
			if (n <= m_cap) return;
			std::size_t const cap(n + n / 2 + 64);
			auto *p(static_cast <t_value *>(pv3_host_alloc(cap * sizeof(t_value))));
			if (!p) throw std::bad_alloc();
			for (std::size_t i(0); i < m_size; ++i) p[i] = m_data[i];
			pv3_host_free(m_data);
			m_data = p; m_cap = cap;
		}

		void resize(std::size_t const n) { reserve(n); m_size = n; }
		void push_back(t_value const v) { if (m_size == m_cap) reserve(m_size + 1); m_data[m_size++] = v; }
		t_value *append(std::size_t const n) { reserve(m_size + n); t_value *p(m_data + m_size); m_size += n; return p; }
		void clear() { m_size = 0; }
		std::size_t size() const { return m_size; }
		t_value *data() { return m_data; }
		t_value const *data() const { return m_data; }
		t_value &operator[](std::size_t const i) { return m_data[i]; }
		t_value const &operator[](std::size_t const i) const { return m_data[i]; }
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
		// 64 gap bits from column 64 w (LSB first), fewer at the end of the row
		static std::uint64_t gap_word(sequence_entry_type const &e, std::uint64_t const w)
		{
			std::uint64_t v(0);
			std::uint64_t const n(e.gap_positions.size());
			for (std::uint64_t i(64 * w); i < n && i < 64 * w + 64; ++i) if (e.gap_positions[i]) v |= std::uint64_t(1) << (i - 64 * w);
			return v;
		}
		static std::uint8_t seq_rank(sequence_alphabet const x) { return x; }
		static std::uint8_t qual_rank(quality_alphabet const x) { return x; }
		static std::uint32_t cigar_count(cigar_type const &c) { return c.count; }
		static char cigar_op_char(cigar_type const &c) { return c.op; }
		static cigar_type make_cigar(std::uint32_t const count, char const op) { return cigar_type{count, op}; }
		static range_type make_range(std::size_t const loc, std::size_t const len) { return range_type{loc, len}; }
	};


	template <typename t_traits> class basic_alignment_projector; // Fwd.

	// Registers a sequence entry's gap row word by word (64 columns per traits call).
	template <typename t_traits, typename t_index>
	std::uint32_t register_entry_words(t_index &index, typename t_traits::sequence_entry_type const &entry)
	{
		std::uint64_t const n_bits(t_traits::n_bits(entry));
		std::vector <std::uint64_t> words((n_bits + 63) / 64);
		for (std::uint64_t w(0); w < words.size(); ++w) words[w] = t_traits::gap_word(entry, w);
		return index.register_row_words(&entry, words.data(), n_bits);
	}

	template <typename t_traits>
	struct basic_alignment_projector_delegate
	{
		virtual ~basic_alignment_projector_delegate() {}
		// On the GPU all runs of a record are realigned inside one call, so both callbacks fire once per realigned run
		// after the call returns; they cannot bracket the DP (documented deviation). A timer driven by them reads
		// ~0 ns: use device_index::last_realignment_ms() for the time the DP kernels took.
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

		std::uint32_t register_entry(sequence_entry_type const &entry) { return register_entry_words <t_traits>(*m_index, entry); }

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


	// The batched entry of the host pipeline: what fills the reference's 4-record tasks
	// (project-alignments/project_alignments.cc:597-631) packs records here instead, into ONE pinned structure of arrays in
	// the compact encoding of pv3_batch_in (16-bit ids, 32-bit positions, lengths; SEQ packed to two bits per base on
	// push_back, N bases listed apart); flush() = one pv3_project (or pv3_multi_project) call; results are read in submission order. Register every
	// sequence entry and contig (register_entry / register_reference) before the first push_back that uses it.
	template <typename t_traits, typename t_index = device_index>
	class basic_batch_projector
	{
	public:
		typedef typename t_traits::sequence_entry_type	sequence_entry_type;
		typedef typename t_traits::sequence_alphabet	sequence_alphabet;
		typedef typename t_traits::quality_alphabet		quality_alphabet;
		typedef typename t_traits::cigar_type			cigar_type;
		typedef typename t_traits::range_type			range_type;

	protected:
		t_index							*m_index{};
		// input, pinned
		pinned_vector <std::uint16_t>	m_src_row, m_dst_row, m_dst_ref, m_cig_len, m_mate_row;
		pinned_vector <std::uint32_t>	m_src_pos, m_seq_len, m_cig, m_mate_pos, m_exc_rec, m_exc_off;
		pinned_vector <std::uint8_t>	m_seq, m_qual, m_has_qual;
		std::vector <std::uint64_t>		m_seq_off{0}, m_cig_off{0}, m_qual_off{0};     // host only: chunk cuts
		bool							m_any_qual{}, m_mates{};
		// output, pinned
		pinned_vector <std::uint64_t>	m_dst_pos, m_out_cig_off, m_out_rr_off, m_rr, m_mate_dst_pos;
		pinned_vector <std::uint32_t>	m_flags, m_out_cig_len, m_out_rr_len, m_out_cig;
		std::size_t						m_flushed{};

		static std::uint32_t bam_op(char const c)
		{
			char const ops[] = "MIDNSHP=X";
			for (std::uint32_t i(0); i < 9; ++i) if (ops[i] == c) return i;
			return 15;
		}

	public:
		explicit basic_batch_projector(t_index &index): m_index(&index) {}

		std::uint32_t register_entry(sequence_entry_type const &entry) { return register_entry_words <t_traits>(*m_index, entry); }
		std::uint32_t register_reference(std::vector <char> const &ref_seq) { return m_index->register_reference(ref_seq); }

		std::size_t size() const { return m_src_row.size(); }

		void clear()
		{
			for (auto *v : { &m_src_row, &m_dst_row, &m_dst_ref, &m_cig_len, &m_mate_row }) v->clear();
			for (auto *v : { &m_src_pos, &m_seq_len, &m_cig, &m_mate_pos, &m_exc_rec, &m_exc_off }) v->clear();
			for (auto *v : { &m_seq, &m_qual, &m_has_qual }) v->clear();
			m_seq_off.assign(1, 0); m_cig_off.assign(1, 0); m_qual_off.assign(1, 0);
			m_any_qual = false; m_mates = false; m_flushed = 0;
		}

		// Arguments as alignment_projector::project_alignment takes them (alignment_projector.hh:58-68), the entries and the
		// contig by the ids register_entry / register_reference returned; returns the record's index in the batch.
		std::size_t push_back(
			std::size_t const src_pos, std::uint32_t const src_row, std::uint32_t const dst_row, std::uint32_t const dst_ref,
			std::vector <sequence_alphabet> const &query_seq, std::vector <cigar_type> const &cigar_seq,
			std::vector <quality_alphabet> const &base_qualities)
		{
			if (src_row > 0xffff || dst_row > 0xffff || dst_ref > 0xffff || src_pos > 0xffffffffull || cigar_seq.size() > 0xffff)
				throw std::range_error("batch_projector: value does not fit the compact encoding");
			if (!base_qualities.empty() && base_qualities.size() != query_seq.size())
				throw std::invalid_argument("batch_projector: QUAL must be empty or as long as SEQ");
			std::size_t const i(size());
			m_src_row.push_back(std::uint16_t(src_row)); m_dst_row.push_back(std::uint16_t(dst_row)); m_dst_ref.push_back(std::uint16_t(dst_ref));
			m_src_pos.push_back(std::uint32_t(src_pos)); m_seq_len.push_back(std::uint32_t(query_seq.size())); m_cig_len.push_back(std::uint16_t(cigar_seq.size()));
			// SEQ in two bits per base (pv3_batch_in.seq_2bit): A C G T (dna5 ranks 0 1 2 4) = 0 1 2 3, four per byte, base 4 b + i in
			// bits 2 i; every other base goes to the exception list as an N
			std::size_t const n_bytes((query_seq.size() + 3) / 4);
			std::uint8_t *sp(m_seq.append(n_bytes));
			for (std::size_t b(0); b < n_bytes; ++b) sp[b] = 0;
			for (std::size_t k(0); k < query_seq.size(); ++k)
			{
				std::uint8_t const rank(t_traits::seq_rank(query_seq[k]));
				std::uint8_t const code(rank < 3 ? rank : std::uint8_t(3));
				if (3 == rank || rank > 4) { m_exc_rec.push_back(std::uint32_t(i)); m_exc_off.push_back(std::uint32_t(k)); }
				else sp[k / 4] = std::uint8_t(sp[k / 4] | (code << (2 * (k % 4))));
			}
			std::uint32_t *cp(m_cig.append(cigar_seq.size()));
			for (std::size_t k(0); k < cigar_seq.size(); ++k) cp[k] = (t_traits::cigar_count(cigar_seq[k]) << 4) | bam_op(t_traits::cigar_op_char(cigar_seq[k]));
			if (0 == i % 8) m_has_qual.push_back(0);
			if (!base_qualities.empty())
			{
				m_has_qual[i / 8] = std::uint8_t(m_has_qual[i / 8] | (1u << (i % 8)));
				std::uint8_t *qp(m_qual.append(base_qualities.size()));
				for (std::size_t k(0); k < base_qualities.size(); ++k) qp[k] = t_traits::qual_rank(base_qualities[k]);
				m_any_qual = true;
			}
			m_seq_off.push_back(m_seq.size()); m_cig_off.push_back(m_cig.size()); m_qual_off.push_back(m_qual.size());
			m_mate_row.push_back(0); m_mate_pos.push_back(0);
			return i;
		}

		// The mate of the record pushed last: its position is projected from mate_row to the record's dst_row
		// (sequence_entry::project_position, project_alignments.cc:880-906).
		void set_mate(std::uint32_t const mate_row, std::size_t const mate_pos)
		{
			m_mate_row[size() - 1] = std::uint16_t(mate_row); m_mate_pos[size() - 1] = std::uint32_t(mate_pos); m_mates = true;
		}

		// Runs every record pushed since the last clear().
		void flush(std::int32_t const gap_opening_cost, std::int32_t const gap_extension_cost)
		{
			std::size_t const n(size());
			m_flushed = 0;
			if (!n) return;
			m_index->commit();
			m_index->set_params(gap_opening_cost, gap_extension_cost);
			pv3_batch_in in{};
			in.n = n; in.compact = 1; in.seq_bam4 = 0; in.seq_2bit = 1;
			in.n_seq_exceptions = m_exc_rec.size(); in.exc_rec = m_exc_rec.data(); in.exc_off = m_exc_off.data();
			in.src_row16 = m_src_row.data(); in.dst_row16 = m_dst_row.data(); in.dst_ref16 = m_dst_ref.data(); in.src_pos32 = m_src_pos.data();
			in.seq_len = m_seq_len.data(); in.cig_len16 = m_cig_len.data(); in.has_qual = m_any_qual ? m_has_qual.data() : nullptr;
			in.seq_off = m_seq_off.data(); in.cig_off = m_cig_off.data(); in.qual_off = m_any_qual ? m_qual_off.data() : nullptr;
			in.seq = m_seq.data(); in.cig = m_cig.data(); in.qual = m_any_qual ? m_qual.data() : nullptr;
			if (m_mates) { in.mate_row16 = m_mate_row.data(); in.mate_pos32 = m_mate_pos.data(); }
			m_dst_pos.resize(n); m_flags.resize(n); m_out_cig_off.resize(n); m_out_cig_len.resize(n); m_out_rr_off.resize(n); m_out_rr_len.resize(n);
			if (m_mates) m_mate_dst_pos.resize(n);
			if (m_out_cig.size() < 2 * m_cig.size() + 8 * n) m_out_cig.resize(2 * m_cig.size() + 8 * n);
			if (m_rr.size() < 4 * (n / 8 + 64)) m_rr.resize(4 * (n / 8 + 64));
			for (;;)
			{
				pv3_batch_out out{};
				out.dst_pos = m_dst_pos.data(); out.flags = m_flags.data(); out.cig_off = m_out_cig_off.data(); out.cig_len = m_out_cig_len.data();
				out.cig = m_out_cig.data(); out.cig_cap = m_out_cig.size();
				out.rr_off = m_out_rr_off.data(); out.rr_len = m_out_rr_len.data(); out.rr = m_rr.data(); out.rr_cap = m_rr.size() / 4;
				out.mate_dst_pos = m_mates ? m_mate_dst_pos.data() : nullptr;
				int const st(m_index->project(in, out));
				if (PV3_ERR_CAPACITY == st)
				{
					m_out_cig.resize(2 * out.cig_used + 64);
					m_rr.resize(4 * (2 * out.rr_used + 8));
					continue;
				}
				if (st < 0) throw std::runtime_error(m_index->last_error());
				break;
			}
			m_flushed = n;
		}

		// -- results of the last flush(), by submission index -------------------------------------------------------------
		std::uint32_t flags(std::size_t const i) const { return m_flags[i]; }
		void throw_if_rejected(std::size_t const i) const { if (m_flags[i]) throw projection_error(m_flags[i], describe_flags(m_flags[i])); }
		std::size_t position(std::size_t const i) const { return m_dst_pos[i]; }
		std::size_t mate_position(std::size_t const i) const { return m_mate_dst_pos[i]; }
		std::size_t realigned_range_count(std::size_t const i) const { return m_out_rr_len[i]; }

		// packed ops (len << 4 | BAM op) of record i: what a BAM writer stores as they are
		std::pair <std::uint32_t const *, std::size_t> packed_alignment(std::size_t const i) const
		{
			return { m_out_cig.data() + m_out_cig_off[i], m_out_cig_len[i] };
		}

		void alignment(std::size_t const i, std::vector <cigar_type> &dst) const
		{
			char const ops[] = "MIDNSHP=X";
			auto const pa(packed_alignment(i));
			dst.clear();
			for (std::size_t k(0); k < pa.second; ++k) dst.push_back(t_traits::make_cigar(pa.first[k] >> 4, ops[pa.first[k] & 0xf]));
		}

		void realigned_ranges(std::size_t const i, std::vector <range_type> &ref_ranges, std::vector <range_type> &query_ranges) const
		{
			ref_ranges.clear(); query_ranges.clear();
			for (std::uint32_t k(0); k < m_out_rr_len[i]; ++k)                  // alignment_projector.cc:69-73
			{
				std::uint64_t const *rr(m_rr.data() + 4 * (m_out_rr_off[i] + k));
				ref_ranges.push_back(t_traits::make_range(rr[0], rr[1]));
				query_ranges.push_back(t_traits::make_range(rr[2], rr[3]));
			}
		}
	};

	typedef basic_batch_projector <plain_traits> plain_batch_projector;
	typedef basic_batch_projector <plain_traits, multi_device_index> plain_multi_batch_projector;
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
		// One rrr decode per 64 columns instead of one per column (rrr_vector<15>::get_int, rrr_vector_15.hpp:320-360):
		// registering a 2.7e8-column row costs 4.2e6 decodes, not 2.7e8 block scans.
		static std::uint64_t gap_word(sequence_entry_type const &e, std::uint64_t const w)
		{
			std::uint64_t const n(e.gap_positions.size()), first(64 * w);
			return e.gap_positions.get_int(first, std::uint8_t(n - first < 64 ? n - first : 64));
		}
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
	typedef basic_batch_projector <seqan3_traits>				batch_projector;
	typedef basic_batch_projector <seqan3_traits, multi_device_index>	multi_batch_projector;
}
#endif

#endif
