// panvc3_b200 — long records as segments (host + device code; tests/cpp/k1w_emul.cc runs it on the CPU against the oracle).
//
// rewrite_cigar() (include/panvc3/rewrite_cigar.hh:59-153) carries three things from one reference-consuming base to the
// next: src_pos, query_pos and prev_excess. The first two are running sums of the input ops; the third is, after ANY
// reference-consuming base, rank0_dst(column of that base) + (dst[column] == 0) (identity proved in pv3_rewrite.cu). So a
// record can be cut anywhere between two reference-consuming bases — inside an M / = / X / D / N op too — and every piece
// rewritten on its own, provided the piece's span starts at the column right AFTER the previous piece's last base (the
// source-gap columns up to its first base are that base's leading deletions) or, for the first piece, at the column of
// the record's first base (rewrite_cigar.hh:51-56: no leading deletion there). Pass-through ops (I S H P) stay with the
// piece in which they are met. The pieces' op lists are then concatenated through cigar_buffer::push_back's merge rule
// (libpanvc3/cigar.cc:61-86). Every piece fits the word-parallel core's quotas (pv3_k1w_core.h), so a 10-kb read becomes
// ~100 independent threads' worth of work instead of one warp walking 10^4 columns.
#pragma once

#include "pv3_k1w_core.h"

namespace pv3 {
namespace seg {

constexpr uint32_t RSEG     = 128;     // reference-consuming bases per segment ...
constexpr uint32_t QSEG     = 160;     // ... and query bases per segment (M = X bases + small I / S ops): the quotas of the core must hold them
// (256 / 288 with an 11-word core was measured slower: 15.4 instead of 14.2 ms per 5 x 10^5 10-kb reads — fewer resident tiles, two staging rounds)
constexpr uint32_t RSEG_DENSE = 64;    // the same where the record or the MSA around it is dense in events (see choose_segment_size):
constexpr uint32_t QSEG_DENSE = 96;    // more ops per base, and a segment's output has to fit SegCfg::OUTQ
constexpr uint32_t BIG_PASS = 32;      // an I / S op longer than this is a segment of its own (nothing of the query is compared in it)
constexpr uint32_t MAXOPS   = 12;      // input ops per segment
constexpr uint32_t UNPLANNABLE = 0xffffffffu;

enum : uint32_t { KIND_CORE = 0, KIND_PASS = 1 };

struct alignas(16) SegDesc {         // 48 bytes: three 16-byte stores by the planner, three loads by the segment kernel
	uint64_t rec;
	uint64_t c_start, e_start;       // its start column and rank0_dst(start column): found by the planner's walk over the two rows
	uint32_t k0;                     // input ops [k0, k0 + n_k) of the record
	uint32_t first_cnt, last_cnt;    // counts of the first / last op inside this segment when they are cut (~0: whole op)
	uint32_t ref_start, q_start;     // reference-consuming / query bases of the record before the segment
	uint16_t ref_count;              // reference-consuming bases of the segment (<= RSEG)
	uint8_t  n_k;                    // <= MAXOPS
	uint8_t  kind;
	PV3_HD uint32_t k1() const { return k0 + n_k; }
};
static_assert(sizeof(SegDesc) == 48, "SegDesc layout");

typedef k1w::CfgSegment <32> SegCfg;     // the quotas the planner cuts for (ST is irrelevant to it)

// position (0..63) of the k-th (1-based) set bit of x; popcount(x) >= k >= 1
PV3_HD uint32_t select64(uint64_t x, uint32_t k)
{
	uint32_t const lo = (uint32_t) x, c = k1w::popc32(lo);
	return k <= c ? k1w::select32(lo, k) : 32u + k1w::select32((uint32_t) (x >> 32), k - c);
}
PV3_HD uint32_t popc64(uint64_t x) { return k1w::popc32((uint32_t) x) + k1w::popc32((uint32_t) (x >> 32)); }

// The planner's cursor over the source and target rows of a record: c = next column to look at, e = rank0_dst(c).
// advance(k) moves it right behind the k-th source base at or after c (k >= 1; the caller guarantees that the row has it).
struct RowCursor {
	const uint64_t *S, *D;           // gap words of both rows (1 = gap), padded with all-ones words
	uint64_t c, e;
	PV3_HD void advance(uint32_t k)
	{
		uint64_t w = c >> 6;
		uint64_t const from = ~0ull << (c & 63u);
		uint64_t sw = ~k1w::ldg(S + w) & from, dw = ~k1w::ldg(D + w) & from;
		for (;;) {
			uint32_t const pc = popc64(sw);
			if (k <= pc) {
				uint32_t const pos = select64(sw, k);
				e += popc64(pos == 63u ? dw : (dw & ((2ull << pos) - 1ull)));
				c = w * 64 + pos + 1;
				return;
			}
			k -= pc; e += popc64(dw); ++w;
			sw = ~k1w::ldg(S + w); dw = ~k1w::ldg(D + w);
		}
	}
};

struct SegOut {
	uint64_t off;                    // first op in the segment op arena
	uint32_t n_ops;
	uint32_t flags;                  // bit 0: an I abuts a D inside the segment; bit 1: the segment could not be done here
	uint32_t first, last;            // its first and last op (what the assembly needs to know about a neighbour)
	uint64_t pad;
};

// Segment size of a record: small where events are dense — more than one input op per 24 query bases (= / X style CIGARs of
// noisy reads), or more than 12 gap / base transitions in the 256 columns of both rows at the record's start. A segment that
// still produces more ops than the core stages sends its record to the warp-per-record kernel, so this only decides speed.
PV3_HD bool dense_record(const uint64_t *S, const uint64_t *D, uint64_t c0, uint64_t nc, uint64_t nq)
{
	if (nc * 24 > nq) return true;
	uint32_t t = 0;
	for (int i = 0; i < 4; ++i) {
		uint64_t const a = k1w::ldg(S + (c0 >> 6) + i), b = k1w::ldg(D + (c0 >> 6) + i);
		t += popc64((a ^ (a >> 1)) & 0x7fffffffffffffffull) + popc64((b ^ (b >> 1)) & 0x7fffffffffffffffull);
	}
	return t > 12;
}

// Walks the ops of one record and calls emit(kind, k0, k1, first_cnt, last_cnt, ref_start, q_start, ref_count) for every segment, in
// order. Returns the number of segments, or UNPLANNABLE (an op outside MIDNSHP=X, counts beyond 32 bits of running sums).
template <typename t_emit>
PV3_HD uint32_t plan_segments(const uint32_t *cig, uint32_t nc, t_emit &&emit, uint32_t *total_ref = nullptr, uint32_t RSEG = seg::RSEG, uint32_t QSEG = seg::QSEG)
{
	uint32_t n_seg = 0;
	uint64_t ref_total = 0, q_total = 0;
	uint32_t k0 = 0, first_cnt = ~0u, ref_acc = 0, q_acc = 0, n_pass = 0, n_in = 0;
	bool open = false;                                         // the current segment holds something
	auto close = [&](uint32_t k1, uint32_t last_cnt) {
		if (open) { emit((uint32_t) KIND_CORE, k0, k1, first_cnt, last_cnt, (uint32_t) ref_total, (uint32_t) q_total, ref_acc); ++n_seg; }
		ref_total += ref_acc; q_total += q_acc;
		ref_acc = q_acc = n_pass = n_in = 0; first_cnt = ~0u; open = false;
	};
	for (uint32_t k = 0; k < nc; ++k) {
		uint32_t const v = k1w::ldg(cig + k), op = v & 0xfu, cnt = v >> 4;
		if (op > k1w::W_X) return UNPLANNABLE;
		if (ref_total + ref_acc > 0x7fffffffull || q_total + q_acc > 0x7fffffffull) return UNPLANNABLE;
		if ((0x72u >> op) & 1u) {                                  // I S H P
			bool const qcons = (op == k1w::W_I || op == k1w::W_S);
			if (qcons && cnt > BIG_PASS) {
				close(k, ~0u);
				emit((uint32_t) KIND_PASS, k, k + 1, ~0u, ~0u, (uint32_t) ref_total, (uint32_t) q_total, 0u); ++n_seg;
				q_total += cnt;
				k0 = k + 1;
				continue;
			}
			if (open && (n_in == MAXOPS || (cnt && n_pass == (uint32_t) SegCfg::NBRK) || (qcons && q_acc + cnt > QSEG))) close(k, ~0u);
			if (!open) { k0 = k; open = true; }
			n_pass += cnt ? 1u : 0u;
			q_acc += qcons ? cnt : 0u;
			++n_in;
		} else {
			bool const mlike = !(op == k1w::W_D || op == k1w::W_N);
			uint32_t remaining = cnt;
			if (!cnt) {                                               // consumes nothing; stays where it is met
				if (open && n_in == MAXOPS) close(k, ~0u);
				if (!open) { k0 = k; open = true; }
				++n_in;
				continue;
			}
			while (remaining) {
				uint32_t room = RSEG - ref_acc;
				if (mlike && QSEG - q_acc < room) room = QSEG - q_acc;
				if (open && (room == 0 || n_in == MAXOPS)) { close(k, ~0u); continue; }   // (a cut op never ends a segment here: see below)
				if (!open) { k0 = k; open = true; first_cnt = (remaining != cnt) ? remaining : ~0u; room = (mlike && QSEG < RSEG) ? QSEG : RSEG; }
				uint32_t const take = remaining < room ? remaining : room;
				ref_acc += take; q_acc += mlike ? take : 0u; remaining -= take;
				if (remaining) close(k + 1, take);                    // the op continues in the next segment
				else ++n_in;
			}
		}
	}
	close(nc, ~0u);
	if (ref_total > 0x7fffffffull || q_total > 0x7fffffffull) return UNPLANNABLE;
	if (total_ref) *total_ref = (uint32_t) ref_total;
	return n_seg;
}

// A KIND_PASS segment: its ops pushed as they are (cigar_buffer::push_back merges equal neighbours, ignores zero counts).
template <typename C>
PV3_HD uint32_t emit_pass_segment(const uint32_t *cig, uint32_t k0, uint32_t k1, k1w::Lane <C> const &L)
{
	constexpr int ST = C::ST;
	uint32_t n = 0, carry_op = k1w::W_NONE, carry_cnt = 0;
	for (uint32_t k = k0; k < k1; ++k) {
		uint32_t const v = k1w::ldg(cig + k), op = v & 0xfu, cnt = v >> 4;
		if (!cnt) continue;
		if (op == carry_op) { carry_cnt += cnt; continue; }
		if (carry_op != k1w::W_NONE) { if (n < (uint32_t) C::OUTQ) L.fin[n * ST] = (carry_cnt << 4) | carry_op; ++n; }
		carry_op = op; carry_cnt = cnt;
	}
	if (carry_op != k1w::W_NONE) { if (n < (uint32_t) C::OUTQ) L.fin[n * ST] = (carry_cnt << 4) | carry_op; ++n; }
	return n;
}

} // namespace seg
} // namespace pv3
