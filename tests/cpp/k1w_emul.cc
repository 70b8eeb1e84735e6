// TEST INFRASTRUCTURE — runs the per-record core of kernel K1W (panvc3_b200/csrc/pv3_k1w_core.h, the very code the CUDA
// kernel executes per thread) on the CPU, so that it can be checked against the oracle without a GPU
// (tests/test_k1w_core.py). Everything around the core that the kernel does cooperatively (query staging, start
// position by select0/rank0) is restated here in the plainest way.
//
//   g++ -O2 -std=c++17 -shared -fPIC -o tests/cpp/_build/libk1w_emul.so tests/cpp/k1w_emul.cc
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../panvc3_b200/csrc/pv3_k1w_core.h"
#include "../../panvc3_b200/csrc/pv3_index_build.h"
#include "../../panvc3_b200/csrc/pv3_segment_core.h"

using namespace pv3;
typedef k1w::CfgRecord <1> RC;
typedef k1w::CfgSegment <1> SC;

namespace {

struct Row {
	std::vector <uint64_t> words;     // padded with all-ones words, bits >= n_bits set
	std::vector <uint64_t> zeros_before;  // per word
	uint64_t n_bits = 0, n_zeros = 0;
};

struct Ref { std::vector <uint8_t> codes4; uint64_t n = 0; };

struct Index { std::vector <Row> rows; std::vector <Ref> refs; };

uint64_t rank0(Row const &r, uint64_t pos)
{
	if (pos >= r.n_bits) return r.n_zeros;
	uint64_t const w = pos >> 6, b = pos & 63;
	uint64_t z = r.zeros_before[w];
	if (b) z += (uint64_t) __builtin_popcountll(~r.words[w] & ((1ull << b) - 1ull));
	return z;
}

uint64_t select0(Row const &r, uint64_t k)   // position of the k-th zero (1-based); n_bits when there are fewer
{
	if (k == 0 || k > r.n_zeros) return r.n_bits;
	size_t lo = 0, hi = r.zeros_before.size() - 1;     // last word with zeros_before < k
	while (lo < hi) { size_t const mid = (lo + hi + 1) / 2; if (r.zeros_before[mid] < k) lo = mid; else hi = mid - 1; }
	uint64_t inv = ~r.words[lo];
	uint64_t need = k - r.zeros_before[lo];
	for (unsigned b = 0; b < 64; ++b) if ((inv >> b) & 1ull) { if (--need == 0) return lo * 64 + b; }
	return r.n_bits;
}

} // namespace

static int g_events_everywhere = 1;   // 1: one launch, event code for every record (what the kernel does); 0: EV = false first, EV = true for the rest

extern "C" {

void k1w_set_mode(int events_everywhere) { g_events_everywhere = events_everywhere; }

void *k1w_index_create() { return new Index; }
void k1w_index_destroy(void *p) { delete (Index *) p; }

int k1w_index_add_row(void *p, const uint64_t *words, uint64_t n_bits)
{
	Index &ix = *(Index *) p;
	Row r;
	r.n_bits = n_bits;
	size_t const nw = (n_bits + 63) / 64;
	r.words.assign(words, words + nw);
	if (n_bits % 64) r.words[nw - 1] |= ~0ull << (n_bits % 64);
	r.words.resize(nw + 72, ~0ull);
	r.zeros_before.resize(r.words.size());
	uint64_t z = 0;
	for (size_t w = 0; w < r.words.size(); ++w) { r.zeros_before[w] = z; z += (uint64_t) __builtin_popcountll(~r.words[w]); }
	r.n_zeros = z;
	ix.rows.push_back(std::move(r));
	return (int) ix.rows.size() - 1;
}

int k1w_index_add_ref(void *p, const char *bases, uint64_t n)
{
	Index &ix = *(Index *) p;
	Ref r; r.n = n;
	r.codes4.resize(codes4_bytes(n) + 64);
	pack_codes4(bases, n, r.codes4.data());
	ix.refs.push_back(std::move(r));
	return (int) ix.refs.size() - 1;
}

// Whole records. status: 0 done by the first launch (no events), 2 done by the second (events), 1 deferred (the general
// kernel would take the record). ops: [n][OUTQ].
// Returns the number of deferred records.
uint64_t k1w_emul_batch(void *p, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row, const uint32_t *dst_ref, const uint64_t *src_pos,
	const uint64_t *seq_off, const uint8_t *seq, const uint64_t *cig_off, const uint32_t *cig,
	uint8_t *status, uint64_t *dst_pos, uint32_t *n_ops, uint32_t *ops, uint8_t *adjacent)
{
	Index &ix = *(Index *) p;
	uint64_t deferred = 0;
	std::vector <uint32_t> q4;
	for (uint64_t rec = 0; rec < n; ++rec) {
		status[rec] = 1; dst_pos[rec] = 0; n_ops[rec] = 0; adjacent[rec] = 0;
		Row const &S = ix.rows[src_row[rec]], &D = ix.rows[dst_row[rec]];
		Ref const &R = ix.refs[dst_ref[rec]];
		uint64_t const spos = src_pos[rec], nq = seq_off[rec + 1] - seq_off[rec], nc = cig_off[rec + 1] - cig_off[rec];
		uint64_t const c0 = (spos < S.n_zeros) ? select0(S, spos + 1) : S.n_bits;
		uint64_t const e0 = rank0(D, c0);
		dst_pos[rec] = e0;
		bool defer = (S.n_bits != D.n_bits || e0 > R.n || nc >= 0x8000 || spos >= S.n_zeros);
		// query staging: nibbles at an arbitrary lead (as in a tile, where the record starts somewhere inside the staged bytes)
		uint32_t const qrel = (uint32_t) (rec * 7u) % 29u;
		q4.assign((qrel + nq + 7) / 8 + 4, 0xeeeeeeeeu);      // foreign nibbles around the record never equal a target code of interest
		for (uint64_t i = 0; i < nq; ++i) {
			uint8_t const rank = seq[seq_off[rec] + i];
			if (rank > 7) defer = true;
			uint32_t const a = qrel + (uint32_t) i;
			q4[a >> 3] = (q4[a >> 3] & ~(0xfu << (4 * (a & 7)))) | ((uint32_t) (rank & 0xf) << (4 * (a & 7)));
		}
		int path = 0;
		if (!defer) {
			uint32_t cmp[RC::NW], ins[RC::NW], del[RC::NW], xm[RC::NW], brk[RC::NBRK], fin[RC::OUTQ];
			k1w::Lane <RC> L; L.cmp = cmp; L.ins = ins; L.del = del; L.xm = xm; L.brk = brk; L.fin = fin;
			k1w::SpanIn I;
			I.Sp = S.words.data() + (c0 >> 6); I.Dp = D.words.data() + (c0 >> 6); I.c_lo6 = (uint32_t) c0 & 63u;
			I.src_left = S.n_zeros - spos; I.e_start = e0; I.ref_n = R.n; I.nq = (uint32_t) nq; I.q_start = 0;
			k1w::OpRange opr; opr.cig = cig + cig_off[rec]; opr.k0 = 0; opr.k1 = (uint32_t) nc; opr.first_cnt = ~0u; opr.last_cnt = ~0u; opr.first_val = nc ? opr.cig[0] : 0u;
			k1w::SpanState st;
			uint64_t const e0a = e0 & ~31ull;
			const uint32_t *rs = reinterpret_cast <const uint32_t *>(R.codes4.data() + (e0a >> 1));
			uint32_t no = 0; bool adj = false;
			// first launch: records without events; second launch (EV = true): the others
			int rc = g_events_everywhere ? 2 : k1w::build_span <RC, false>(I, opr, L, st);
			if (rc == 0) {
				k1w::compare_span <RC, false>(L, st.n_brk, opr.cig, st.w_end, q4.data(), qrel, 0u, rs, (uint32_t) (e0 - e0a));
				k1w::mark_runs <RC, false>(L, st.n_brk, st.w_end);
				no = k1w::emit_ops <RC, false>(L, st.n_brk, opr, st.w_end, adj);
			} else if (rc == 2) {
				rc = k1w::build_span <RC, true>(I, opr, L, st);
				path = st.simple ? 0 : 2;
				if (rc == 0) {
					k1w::compare_span <RC, true>(L, st.n_brk, opr.cig, st.w_end, q4.data(), qrel, 0u, rs, (uint32_t) (e0 - e0a));
					k1w::mark_runs <RC, true>(L, st.n_brk, st.w_end);
					no = k1w::emit_ops <RC, true>(L, st.n_brk, opr, st.w_end, adj);
				}
			}
			if (rc != 0 || no > (uint32_t) RC::OUTQ) defer = true;
			else {
				n_ops[rec] = no; adjacent[rec] = adj;
				memcpy(ops + rec * RC::OUTQ, fin, no * 4);
			}
		}
		status[rec] = defer ? 1 : (uint8_t) path;
		deferred += defer;
	}
	return deferred;
}

int k1w_outq() { return RC::OUTQ; }

// Records as segments (pv3_segment_core.h): plan, the word-parallel core per segment, cigar_buffer-style concatenation.
// status: 0 done, 1 the general kernel would take the record. ops: [n][cap] words, n_ops[n]. Returns the segments planned.
uint64_t k1w_emul_segments(void *p, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row, const uint32_t *dst_ref, const uint64_t *src_pos,
	const uint64_t *seq_off, const uint8_t *seq, const uint64_t *cig_off, const uint32_t *cig,
	uint8_t *status, uint64_t *dst_pos, uint32_t *n_ops, uint32_t *ops, uint64_t cap)
{
	Index &ix = *(Index *) p;
	uint64_t total_segs = 0;
	std::vector <uint32_t> q4;
	std::vector <seg::SegDesc> segs;
	for (uint64_t rec = 0; rec < n; ++rec) {
		status[rec] = 1; n_ops[rec] = 0;
		Row const &S = ix.rows[src_row[rec]], &D = ix.rows[dst_row[rec]];
		Ref const &R = ix.refs[dst_ref[rec]];
		uint64_t const spos0 = src_pos[rec], nq = seq_off[rec + 1] - seq_off[rec], nc = cig_off[rec + 1] - cig_off[rec];
		const uint32_t *rc = cig + cig_off[rec];
		uint64_t const c0 = (spos0 < S.n_zeros) ? select0(S, spos0 + 1) : S.n_bits;
		dst_pos[rec] = rank0(D, c0);
		if (S.n_bits != D.n_bits || nc >= (1u << 20)) continue;
		segs.clear();
		uint32_t total_ref = 0;
		if (spos0 >= S.n_zeros) continue;
		uint32_t rseg = seg::RSEG, qseg = seg::QSEG;
		if (seg::dense_record(S.words.data(), D.words.data(), c0, nc, nq)) { rseg = seg::RSEG_DENSE; qseg = seg::QSEG_DENSE; }
		uint32_t ns = seg::plan_segments(rc, (uint32_t) nc, [](uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t) {}, &total_ref, rseg, qseg);
		if (ns == seg::UNPLANNABLE || (uint64_t) total_ref > S.n_zeros - spos0) continue;
		seg::RowCursor cur; cur.S = S.words.data(); cur.D = D.words.data(); cur.c = c0; cur.e = dst_pos[rec];
		seg::plan_segments(rc, (uint32_t) nc, [&](uint32_t kind, uint32_t k0, uint32_t k1, uint32_t fc, uint32_t lc, uint32_t rs, uint32_t qs, uint32_t rcnt) {
			seg::SegDesc d; d.rec = rec; d.k0 = k0; d.n_k = (uint8_t) (k1 - k0); d.first_cnt = fc; d.last_cnt = lc; d.ref_start = rs; d.q_start = qs; d.kind = (uint8_t) kind; d.ref_count = (uint16_t) rcnt;
			d.c_start = cur.c; d.e_start = cur.e;
			if (rcnt) cur.advance(rcnt);
			segs.push_back(d);
		}, nullptr, rseg, qseg);
		total_segs += ns;
		bool fail = false;
		uint32_t carry_op = k1w::W_NONE, carry_cnt = 0, n_out = 0;
		uint32_t *out = ops + rec * cap;
		auto push = [&](uint32_t v) {
			uint32_t const op = v & 0xfu, cnt = v >> 4;
			if (!cnt) return;
			if (op == carry_op) { carry_cnt += cnt; return; }
			if (carry_op != k1w::W_NONE) { if (n_out < cap) out[n_out] = (carry_cnt << 4) | carry_op; ++n_out; }
			carry_op = op; carry_cnt = cnt;
		};
		for (auto const &d : segs) {
			uint32_t cmp[SC::NW], ins[SC::NW], del[SC::NW], xm[SC::NW], brk[SC::NBRK], fin[SC::OUTQ];
			k1w::Lane <SC> L; L.cmp = cmp; L.ins = ins; L.del = del; L.xm = xm; L.brk = brk; L.fin = fin;
			uint32_t no = 0;
			if (d.kind == seg::KIND_PASS) no = seg::emit_pass_segment <SC>(rc, d.k0, d.k1(), L);
			else {
				uint64_t const spos = spos0 + d.ref_start;
				uint64_t const c_start = d.c_start, e_start = d.e_start;        // the planner's walk; cross-checked against select / rank
				if (c_start != (d.ref_start == 0 ? select0(S, spos + 1) : select0(S, spos) + 1) || e_start != rank0(D, c_start)) abort();
				if (e_start > R.n || d.q_start > nq) { fail = true; break; }
				uint64_t const q_left = nq - d.q_start;
				uint32_t const nq_seg = (uint32_t) (q_left < seg::QSEG ? q_left : seg::QSEG);   // (the kernel stages up to QSEG bases whatever the segment size)
				uint32_t const qrel = (uint32_t) (rec * 5u + d.k0) % 23u;
				q4.assign((qrel + nq_seg + 7) / 8 + 4, 0xeeeeeeeeu);
				for (uint32_t i = 0; i < nq_seg; ++i) {
					uint8_t const rank = seq[seq_off[rec] + d.q_start + i];
					if (rank > 7) fail = true;
					uint32_t const a = qrel + i;
					q4[a >> 3] = (q4[a >> 3] & ~(0xfu << (4 * (a & 7)))) | ((uint32_t) (rank & 0xf) << (4 * (a & 7)));
				}
				if (fail) break;
				k1w::SpanIn I;
				I.Sp = S.words.data() + (c_start >> 6); I.Dp = D.words.data() + (c_start >> 6); I.c_lo6 = (uint32_t) c_start & 63u;
				I.src_left = S.n_zeros - spos; I.e_start = e_start; I.ref_n = R.n; I.nq = nq_seg; I.q_start = 0;
				k1w::OpRange opr; opr.cig = rc; opr.k0 = d.k0; opr.k1 = d.k1(); opr.first_cnt = d.first_cnt; opr.last_cnt = d.last_cnt; opr.first_val = rc[d.k0];
				k1w::SpanState st;
				if (k1w::build_span <SC, true>(I, opr, L, st)) { fail = true; break; }
				uint64_t const ea = e_start & ~31ull;
				k1w::compare_span <SC, true>(L, st.n_brk, opr.cig, st.w_end, q4.data(), qrel, 0u,
					reinterpret_cast <const uint32_t *>(R.codes4.data() + (ea >> 1)), (uint32_t) (e_start - ea));
				k1w::mark_runs <SC, true>(L, st.n_brk, st.w_end);
				bool adj = false;
				no = k1w::emit_ops <SC, true>(L, st.n_brk, opr, st.w_end, adj);
			}
			if (no > (uint32_t) SC::OUTQ) { fail = true; break; }
			for (uint32_t j = 0; j < no; ++j) push(fin[j]);
		}
		if (fail) continue;
		if (carry_op != k1w::W_NONE) { if (n_out < cap) out[n_out] = (carry_cnt << 4) | carry_op; ++n_out; }
		if (n_out > cap) continue;
		n_ops[rec] = n_out;
		status[rec] = 0;
	}
	return total_segs;
}

} // extern "C"
