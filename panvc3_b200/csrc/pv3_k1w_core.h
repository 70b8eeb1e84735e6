// panvc3_b200 — per-record core of kernel K1W (word-parallel projection + CIGAR rewrite), compiled for the device AND the
// host: tests/cpp/k1w_emul.cc runs exactly this code on the CPU against the oracle, the kernel in pv3_rewrite_word.cu runs
// it one thread per record (or per segment of a long record).
//
// Reference semantics: include/panvc3/rewrite_cigar.hh:50-156 with cigar_buffer::push_back/finish (libpanvc3/cigar.cc:61-95).
// Formulation (identity proved in pv3_rewrite.cu's header): every MSA column of the record's span [c0, column of its last
// reference-consuming base] yields at most one unit op decided by (source bit, target bit, class of the input op that owns
// the source base):
//     src base & dst base : M-like -> compare ('=' / 'X')   D-like -> 'D'
//     src base & dst gap  : M-like -> 'I'                   D-like -> nothing
//     src gap  & dst base : 'D' (leading deletion of the next base)        both gaps: nothing
// and the pass-through ops (I S H P, :68-77) are pushed between columns: an op met after b reference-consuming bases goes
// right after the column of base b-1 (before the leading deletions of base b).
//
// v14 (round 2). The v4-v13 kernel walked the column RUNS of a record in one flat loop (10.5 runs per 150-bp record on
// configs[1], ~110 instructions each, 13 of 32 lanes active because a warp runs as long as its record with most runs).
// Here the span is held as four bit masks over 32-column words — cmp / ins / del / mismatch — and every loop runs over
// something that is nearly the same for all records of a warp:
//   load_span    fixed: NW word pairs, span end by population counts (no per-run work)
//   apply_events one iteration per input op, only for records that have I S H P D N ops (EV = true: such records are
//                collected by the first launch and handled together by a second one, so the event code runs with full warps)
//   compare_span one iteration per <= 16 compared bases (150 bases: 10-14 steps whatever the gap structure)
//   mark_runs    per word: the columns where a new final op starts, by carry propagation through the columns that yield
//                nothing (no loop over runs)
//   emit_ops     one iteration per final op (or empty word); one cigar_buffer::push_back site
// Anything unusual (errors the reference aborts on, spans over NW words, too many events or output ops) is NOT handled
// here: the record is deferred to the warp-per-record kernel, which has the reference's sequential semantics.
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define PV3_HD __host__ __device__ __forceinline__
#define PV3_HD_COLD __host__ __device__ __noinline__
#define PV3_UNROLL _Pragma("unroll")
#define PV3_NOUNROLL _Pragma("unroll 1")
#else
#define PV3_HD inline
#define PV3_HD_COLD inline
#define PV3_UNROLL
#define PV3_NOUNROLL
#endif

namespace pv3 {
namespace k1w {

// Quotas of one instantiation of the core. ST: threads per tile (the stride of the per-thread arrays, 1 on the host);
// NW: 32-column words per span (including the offset of the start column in its word); NBRK: pass-through ops (I S H P
// with a non-zero count); OUTQ: staged output ops.
template <int t_st> struct CfgRecord  { static constexpr int ST = t_st, NW = 10, NBRK = 6, OUTQ = 24; static constexpr uint32_t QMAX = 192; };   // whole 150-bp records (K1W)
template <int t_st> struct CfgSegment { static constexpr int ST = t_st, NW = 7,  NBRK = 8, OUTQ = 32; static constexpr uint32_t QMAX = 192; };   // <= 128-base segments of long records (K1S)
constexpr uint32_t CLAMP = 1u << 20;
constexpr uint32_t NOBRK = 0xffffffffu;

enum : uint32_t { W_M = 0, W_I = 1, W_D = 2, W_N = 3, W_S = 4, W_H = 5, W_P = 6, W_EQ = 7, W_X = 8, W_NONE = 15 };

#if defined(__CUDA_ARCH__)
PV3_HD uint32_t popc32(uint32_t x) { return (uint32_t) __popc(x); }
PV3_HD uint32_t fshr(uint32_t lo, uint32_t hi, uint32_t sh) { return __funnelshift_r(lo, hi, sh); }
template <typename T> PV3_HD T ldg(const T *p) { return __ldg(p); }
#else
PV3_HD uint32_t popc32(uint32_t x) { return (uint32_t) __builtin_popcount(x); }
PV3_HD uint32_t fshr(uint32_t lo, uint32_t hi, uint32_t sh) { return (uint32_t) (((((uint64_t) hi) << 32) | lo) >> (sh & 31u)); }
template <typename T> PV3_HD T ldg(const T *p) { return *p; }
#endif
PV3_HD uint32_t ctz32(uint32_t x) { return popc32(~x & (x - 1u)); }                    // 32 for x == 0
PV3_HD uint32_t lowmask(uint32_t n) { return n >= 32u ? 0xffffffffu : ((1u << n) - 1u); }   // bits [0, n), n in 0..32
PV3_HD uint32_t umin(uint32_t a, uint32_t b) { return a < b ? a : b; }

// position (0..31) of the k-th (1-based) set bit of x; popc(x) >= k >= 1
PV3_HD uint32_t select32(uint32_t x, uint32_t k)
{
	uint32_t pos = 0, c;
	c = popc32(x & 0xffffu); if (k > c) { k -= c; pos += 16; x >>= 16; }
	c = popc32(x & 0xffu);   if (k > c) { k -= c; pos += 8;  x >>= 8; }
	c = popc32(x & 0xfu);    if (k > c) { k -= c; pos += 4;  x >>= 4; }
	c = popc32(x & 0x3u);    if (k > c) { k -= c; pos += 2;  x >>= 2; }
	c = x & 1u;              if (k > c) { pos += 1; }
	return pos;
}

// 8 nibbles of t → 8 bits (bit i set ⇔ nibble i non-zero), returned in bits 24..31
PV3_HD uint32_t nonzero_nibbles_hi(uint32_t t)
{
	uint32_t const nz = (((t & 0x77777777u) + 0x77777777u) | t) & 0x88888888u;     // bit 4i+3
	uint32_t const y = nz >> 3;                                                      // bit 4i
	uint32_t const y1 = (y | (y >> 3)) & 0x03030303u;                               // bits 8j, 8j+1
	return y1 * 0x01041040u;                                                         // byte j → bits 24+2j, 25+2j (no carries)
}

// Per-thread view of the tile's arrays: element i of an array is p[i * ST] (ST = threads per tile on the device, 1 on the host).
template <typename C> struct Lane {
	uint32_t *cmp, *ins, *del, *xm;     // [NW]   span masks, bit = column (LSB first) relative to the 32-column word of c0
	                                    //        (after mark_runs: cmp = any op, ins = code bit 1, del = code bit 0, xm = run starts)
	uint32_t *brk;                      // [NBRK] pass-through ops in input order: column position | input op index << 12
	uint32_t *fin;                      // [OUTQ] final ops (len << 4 | op); may alias memory that is dead after compare_span
};

struct OpsInfo {
	uint32_t tot_ref;   // reference-consuming bases (M = X D N), clamped
	uint32_t tot_q;     // query bases the ops consume (M = X I S), clamped
	bool     simple;    // M = X ops only: no events
	bool     bad;       // an op outside MIDNSHP=X (rewrite_cigar.hh:149-151)
};

// ops [k0, k1) of the record; the first and last op may be cut (segments of a long record): `first_cnt` / `last_cnt` replace
// their counts when not ~0u.
struct OpRange {
	const uint32_t *cig;
	uint32_t k0, k1;
	uint32_t first_cnt, last_cnt;
	uint32_t first_val;          // op k0, already loaded by the caller (its load is issued before the rank / select chain)
	PV3_HD uint32_t op_at(uint32_t k) const { return k == k0 ? first_val : ldg(cig + k); }
	PV3_HD uint32_t count(uint32_t k, uint32_t v) const
	{
		uint32_t c = v >> 4;
		if (k == k0 && first_cnt != ~0u) c = first_cnt;
		if (k + 1 == k1 && last_cnt != ~0u) c = last_cnt;
		return c;
	}
};

PV3_HD OpsInfo scan_ops(OpRange const &R)
{
	OpsInfo r; r.tot_ref = 0; r.tot_q = 0; r.simple = true; r.bad = false;
	PV3_NOUNROLL
	for (uint32_t k = R.k0; k < R.k1; ++k) {
		uint32_t const v = R.op_at(k), op = v & 0xfu, cnt = umin(R.count(k, v), CLAMP);
		bool const pass = ((0x72u >> op) & 1u) != 0, dl = (op == W_D || op == W_N), qc = (op == W_I || op == W_S || (!pass && !dl));
		r.bad |= op > W_X;
		r.simple &= !((pass || dl) && cnt);
		r.tot_ref = umin(r.tot_ref + (pass ? 0u : cnt), CLAMP);
		r.tot_q = umin(r.tot_q + (qc ? cnt : 0u), CLAMP);
	}
	return r;
}

// Loads the NW 32-column words of both rows from the word that holds the start column and builds the span masks.
// Sp / Dp: the 64-bit words holding the start column; c_lo6: start column & 63. The span ends with the column of the
// tot_ref-th source base at or after the start column. Returns false when the span does not end within NW words.
template <typename C>
PV3_HD bool load_span(const uint64_t *Sp, const uint64_t *Dp, uint32_t c_lo6, uint32_t tot_ref, Lane <C> const &L,
	uint32_t &span_begin, uint32_t &span_end, uint32_t &n_dbase)
{
	constexpr int ST = C::ST, NW = C::NW, NBRK = C::NBRK, OUTQ = C::OUTQ; (void) ST; (void) NW; (void) NBRK; (void) OUTQ;
	constexpr int NL = NW / 2 + 1;
	bool const half = (c_lo6 >> 5) != 0;
	uint32_t const off = c_lo6 & 31u, m0 = 0xffffffffu << off;
	uint64_t s[NL], d[NL];
	PV3_UNROLL
	for (int i = 0; i < NL; ++i) { s[i] = ldg(Sp + i); d[i] = ldg(Dp + i); }
	// 32-bit half i + half of the loaded words (compile-time register indices), inverted: 1 = base; the word that holds the
	// tot_ref-th source base is found on the way (one population count and three selects per word)
	uint32_t sb[NW], db[NW];
	uint32_t need = tot_ref, xf = 0, nf = 0, wf = (uint32_t) NW;
	PV3_UNROLL
	for (int i = 0; i < NW; ++i) {
		uint32_t sw, dw;
		if (i & 1) {
			sw = half ? (uint32_t) s[(i + 1) / 2] : (uint32_t) (s[i / 2] >> 32);
			dw = half ? (uint32_t) d[(i + 1) / 2] : (uint32_t) (d[i / 2] >> 32);
		} else {
			sw = half ? (uint32_t) (s[i / 2] >> 32) : (uint32_t) s[i / 2];
			dw = half ? (uint32_t) (d[i / 2] >> 32) : (uint32_t) d[i / 2];
		}
		sb[i] = ~sw & (i == 0 ? m0 : 0xffffffffu);
		db[i] = ~dw & (i == 0 ? m0 : 0xffffffffu);
		uint32_t const pc = popc32(sb[i]);
		bool const hit = wf == (uint32_t) NW && need != 0 && need <= pc;
		xf = hit ? sb[i] : xf; nf = hit ? need : nf; wf = hit ? (uint32_t) i : wf;
		need -= umin(pc, need);                                                  // stays 0 once the word is found
	}
	bool const found = tot_ref == 0 || wf != (uint32_t) NW;
	span_begin = off;
	span_end = tot_ref == 0 ? off : (wf != (uint32_t) NW ? wf * 32u + select32(xf, nf) + 1u : (uint32_t) NW * 32u);
	n_dbase = 0;
	PV3_UNROLL
	for (int i = 0; i < NW; ++i) {
		uint32_t const lo = (uint32_t) i * 32u;
		uint32_t const m = lowmask(span_end > lo ? umin(span_end - lo, 32u) : 0u);
		uint32_t const x = sb[i] & m, y = db[i] & m;
		L.cmp[i * ST] = x & y;
		L.ins[i * ST] = x & ~y;
		L.del[i * ST] = ~x & y;
		L.xm[i * ST] = 0;
		n_dbase += popc32(y);
	}
	return found;
}

// Second pass over the input ops (records with I S H P D N ops only). Pass-through ops are recorded with the column they
// go in front of; the source bases of D-like ops are collected in xm (free until compare_span) and then moved from cmp to
// del and dropped from ins (rewrite_cigar.hh:119-147). Event columns come from a cursor over the span's source bases that
// only moves forward. Returns false when the record has more than NBRK pass-through ops.
template <typename C>
PV3_HD bool apply_events(OpRange const &R, uint32_t tot_ref, uint32_t span_begin, uint32_t span_end, Lane <C> const &L, uint32_t &n_brk)
{
	constexpr int ST = C::ST, NW = C::NW, NBRK = C::NBRK, OUTQ = C::OUTQ; (void) ST; (void) NW; (void) NBRK; (void) OUTQ;
	uint32_t b = 0, wc = 0, bc = 0, xc = L.cmp[0] | L.ins[0];
	bool ok = true, any_d = false;
	n_brk = 0;
	auto col_of = [&](uint32_t j) -> uint32_t {                                  // column of source base j (j < tot_ref, non-decreasing over calls)
		PV3_NOUNROLL
		while (wc + 1u < (uint32_t) NW && bc + popc32(xc) <= j) { bc += popc32(xc); ++wc; xc = L.cmp[wc * ST] | L.ins[wc * ST]; }
		uint32_t const k = j - bc + 1u;
		return wc * 32u + (k <= popc32(xc) ? select32(xc, k) : 31u);
	};
	PV3_NOUNROLL
	for (uint32_t k = R.k0; k < R.k1; ++k) {
		uint32_t const v = ldg(R.cig + k), op = v & 0xfu, cnt = umin(R.count(k, v), CLAMP);
		if (cnt) {                                                               // cigar_buffer::push_back ignores a zero count (cigar.cc:63)
			if ((0x72u >> op) & 1u) {                                             // I S H P: right after the column of base b - 1
				uint32_t const pos = (b == 0) ? span_begin : ((b >= tot_ref) ? span_end : col_of(b - 1u) + 1u);
				if (n_brk == (uint32_t) NBRK) ok = false;
				else { L.brk[n_brk * ST] = pos | (k << 12); ++n_brk; }
			} else {
				if (op == W_D || op == W_N) {
					uint32_t const lo = col_of(b), hi = col_of(b + cnt - 1u);
					any_d = true;
					PV3_NOUNROLL
					for (uint32_t i = lo >> 5; i <= (hi >> 5); ++i) {
						uint32_t const base = i * 32u, l = lo > base ? lo - base : 0u, h = umin(hi - base, 31u);
						L.xm[i * ST] |= lowmask(h + 1u) & ~lowmask(l);
					}
				}
				b += cnt;
			}
		}
	}
	if (any_d) {
		PV3_NOUNROLL
		for (int i = 0; i < NW; ++i) {
			uint32_t const m = L.xm[i * ST], t = L.cmp[i * ST] & m;
			L.cmp[i * ST] &= ~m;
			L.del[i * ST] |= t;
			L.ins[i * ST] &= ~m;
			L.xm[i * ST] = 0;
		}
	}
	return ok;
}

// 16 nibbles at nibble offset a of a word array (reads 3 words)
PV3_HD void load16n(const uint32_t *base, uint32_t a, uint32_t &lo, uint32_t &hi, bool global)
{
	const uint32_t *p = base + (a >> 3);
	uint32_t const sh = (a & 7u) * 4u;
	uint32_t w0, w1, w2;
	if (global) { w0 = ldg(p); w1 = ldg(p + 1); w2 = ldg(p + 2); } else { w0 = p[0]; w1 = p[1]; w2 = p[2]; }
	lo = fshr(w0, w1, sh);
	hi = fshr(w1, w2, sh);
}

// Mismatch bits of every compare column (rewrite_cigar.hh:102-113), in COLUMN coordinates: one iteration per <= 16 bases of
// a maximal run of compare columns inside a 32-column word. qs: staged query nibbles (dna5 ranks), record's base 0 at
// nibble qrel. rs: packed target codes (RefDesc::codes4) as words from a 32-base boundary, base e_start at nibble ra0.
// The loop body has ONE merge point (no continue / break inside): ncu of the first version showed the lanes of a warp
// running the compare block in separate passes (12.7 of 32 lanes active) when the body had several exits.
template <typename C, bool EV>
PV3_HD void compare_span(Lane <C> const &L, uint32_t n_brk, const uint32_t *cig, uint32_t w_end,
	const uint32_t *qs, uint32_t qrel, uint32_t q_start, const uint32_t *rs, uint32_t ra0)
{
	constexpr int ST = C::ST, NW = C::NW, NBRK = C::NBRK, OUTQ = C::OUTQ; (void) ST; (void) NW; (void) NBRK; (void) OUTQ;
	uint32_t bi = 0;
	uint32_t brk = (EV && n_brk) ? L.brk[0] : NOBRK;
	uint32_t brk_pos = (EV && n_brk) ? (brk & 0xfffu) : NOBRK;
	uint32_t w = 0, cumS = 0, cumD = 0, qpass = q_start;
	uint32_t cw = w_end ? L.cmp[0] : 0u, iw = w_end ? L.ins[0] : 0u, dw = w_end ? L.del[0] : 0u;   // masks of word w
	uint32_t cur = cw;                                                            // compare columns of word w not yet handled
	while (w < w_end) {
		if (cur != 0) {
			uint32_t const s = ctz32(cur), col = w * 32u + s;
			if (EV && brk_pos <= col) {                                          // a pass-through op before this column: I and S shift the query
				uint32_t const v = ldg(cig + (brk >> 12)), op = v & 0xfu;
				if (op == W_I || op == W_S) qpass += umin(v >> 4, CLAMP);
				++bi;
				brk = bi < n_brk ? L.brk[bi * ST] : NOBRK;
				brk_pos = bi < n_brk ? (brk & 0xfffu) : NOBRK;
			} else {
				uint32_t run = ctz32(~(cur >> s));                               // <= 32 - s
				if (EV) run = umin(run, brk_pos - col);
				uint32_t const n = umin(run, 16u), below = (1u << s) - 1u;
				uint32_t const q = qpass + cumS + popc32((cw | iw) & below);
				uint32_t const r = ra0 + cumD + popc32((cw | dw) & below);
				uint32_t qlo, qhi, rlo, rhi;
				load16n(qs, qrel + q, qlo, qhi, false);
				load16n(rs, r, rlo, rhi, true);
				uint32_t bits = (nonzero_nibbles_hi(qlo ^ rlo) >> 24) | ((nonzero_nibbles_hi(qhi ^ rhi) >> 16) & 0xff00u);
				uint32_t const take = (1u << n) - 1u;                            // n <= 16
				bits &= take;
				L.xm[w * ST] |= bits << s;
				cur &= ~(take << s);
			}
		}
		// word exhausted: next word, in the same iteration as its last step; predicated, and behind an opaque barrier so that
		// the compiler does not thread the branch above into two copies of this tail (lanes then run it in separate passes)
#if defined(__CUDA_ARCH__)
		asm volatile("" : "+r"(cur));
#endif
		bool const adv = cur == 0;
		cumS += adv ? popc32(cw | iw) : 0u;
		cumD += adv ? popc32(cw | dw) : 0u;
		w += adv ? 1u : 0u;
		uint32_t const wl = umin(w, (uint32_t) NW - 1u);
		uint32_t const cw_n = L.cmp[wl * ST], iw_n = L.ins[wl * ST], dw_n = L.del[wl * ST];
		cw = adv ? cw_n : cw; iw = adv ? iw_n : iw; dw = adv ? dw_n : dw;
		cur = adv ? ((w < w_end) ? cw_n : 0u) : cur;
	}
}

// Where the final ops start. Every column has a 2-bit code (= 00, X 01, I 10, D 11) or yields nothing; a new op starts at a
// column whose code differs from the code of the previous column that yields something. The previous code reaches a column
// through the empty columns by carry propagation (adding the empty-column mask to the bits that enter its runs), one add
// per code bit and word — no loop over runs. Pass-through ops force a start at the first column at or after their position.
// In place: cmp ← columns with an op, ins ← code bit 1, del ← code bit 0, xm ← run starts.
template <typename C, bool EV>
PV3_HD void mark_runs(Lane <C> const &L, uint32_t n_brk, uint32_t w_end)
{
	constexpr int ST = C::ST, NW = C::NW, NBRK = C::NBRK, OUTQ = C::OUTQ; (void) ST; (void) NW; (void) NBRK; (void) OUTQ;
	uint32_t top0 = 0, top1 = 0, run0 = 0, run1 = 0, bj = 0;
	bool seen = false, force = false;
	PV3_NOUNROLL
	for (uint32_t w = 0; w < w_end; ++w) {
		uint32_t const cm = L.cmp[w * ST], im = L.ins[w * ST], dm = L.del[w * ST], xw = L.xm[w * ST];
		uint32_t const nn = cm | im | dm, N = ~nn;
		uint32_t const B0 = (cm & xw) | dm, B1 = im | dm;
		uint32_t const st0 = (B0 << 1) | top0 | run0, st1 = (B1 << 1) | top1 | run1;
		uint64_t const t0 = (uint64_t) (st0 & N) + N, t1 = (uint64_t) (st1 & N) + N;
		uint32_t const p0 = (st0 | (uint32_t) t0) & nn, p1 = (st1 | (uint32_t) t1) & nn;      // code of the previous op column, at op columns
		run0 = (uint32_t) (t0 >> 32); run1 = (uint32_t) (t1 >> 32);
		top0 = B0 >> 31; top1 = B1 >> 31;
		uint32_t T = nn & ((B0 ^ p0) | (B1 ^ p1));
		uint32_t const lowest = nn & (0u - nn);
		if (!seen || force) T |= lowest;                                          // the first op column of the span; a pass-through op before this word
		seen |= nn != 0;
		force &= nn == 0;
		if (EV) {
			while (bj < n_brk && ((L.brk[bj * ST] & 0xfffu) >> 5) <= w) {        // pass-through ops in front of a column of this word
				uint32_t const pos = L.brk[bj * ST] & 0xfffu;
				uint32_t const cand = ((pos >> 5) == w) ? (nn & (0xffffffffu << (pos & 31u))) : nn;
				if (cand) T |= cand & (0u - cand); else force = true;
				++bj;
			}
		}
		L.cmp[w * ST] = nn; L.ins[w * ST] = B1; L.del[w * ST] = B0; L.xm[w * ST] = T;
	}
}

// Run-length emission of the final ops after mark_runs: one iteration per final op (or word without a run start); one
// cigar_buffer::push_back site (cigar.cc:61-86) on the common path. The length of a run is the difference of the op-column
// ranks of two consecutive run starts. Returns the op count (cigar_buffer::finish included); more than OUTQ means the list
// overflowed (the caller defers the record).
template <typename C, bool EV>
PV3_HD uint32_t emit_ops(Lane <C> const &L, uint32_t n_brk, OpRange const &R, uint32_t w_end, bool &adjacent_id)
{
	constexpr int ST = C::ST, NW = C::NW, NBRK = C::NBRK, OUTQ = C::OUTQ; (void) ST; (void) NW; (void) NBRK; (void) OUTQ;
	uint32_t bi = 0;
	uint32_t brk = (EV && n_brk) ? L.brk[0] : NOBRK;
	uint32_t brk_pos = (EV && n_brk) ? (brk & 0xfffu) : NOBRK;
	uint32_t n_ops = 0, carry_op = W_NONE, carry_cnt = 0;
	bool adj = false;
	auto push = [&](uint32_t op, uint32_t len) {
		// cigar_buffer::push_back without a branch: the open op is kept in its slot (rewritten as it grows); a different op
		// closes it by advancing the count
		bool const fresh = len != 0 && carry_op != op;
		adj |= fresh && (carry_op | op) == 3u;                                    // I next to D (M and N are never produced)
		n_ops += (fresh && carry_op != W_NONE) ? 1u : 0u;
		carry_cnt = (fresh ? 0u : carry_cnt) + len;
		carry_op = fresh ? op : carry_op;
		L.fin[umin(n_ops, (uint32_t) OUTQ - 1u) * ST] = (carry_cnt << 4) | carry_op;   // clamped only when the list overflows
	};
	auto push_pass = [&]() {
		uint32_t const k = brk >> 12, v = ldg(R.cig + k);
		push(v & 0xfu, R.count(k, v));
		++bi;
		brk = bi < n_brk ? L.brk[bi * ST] : NOBRK;
		brk_pos = bi < n_brk ? (brk & 0xfffu) : NOBRK;
	};
	uint32_t w = 0, rbase = 0, prev = W_NONE, prev_r = 0;
	uint32_t tw = w_end ? L.xm[0] : 0u, nw = w_end ? L.cmp[0] : 0u;
	// Straight-line body with predicated updates (the compiler split an if / if version into two passes per iteration: 6.9
	// of 32 lanes active on the push, ncu r2_v14c). Only the pass-through ops (EV, rare) branch.
	while (w < w_end) {
		bool const has = tw != 0;
		uint32_t const p = ctz32(tw) & 31u, col = w * 32u + p;
		uint32_t const rp = rbase + popc32(nw & ((1u << p) - 1u));                 // op columns before this one
		uint32_t const len = (has && prev != W_NONE) ? rp - prev_r : 0u;
		push(prev, len);                                                           // closes the open run (no-op without one)
		bool brk_now = false;
		if (EV) {
			brk_now = has && brk_pos <= col;
			if (brk_now) {                                                         // pass-through op(s) in front of this column
				prev = W_NONE; prev_r = rp;
				push_pass();
			}
		}
		bool const take = has && !brk_now;
		uint32_t const idx = (((L.ins[w * ST] >> p) & 1u) << 1) | ((L.del[w * ST] >> p) & 1u);
		prev = take ? ((0x2187u >> (4u * idx)) & 0xfu) : prev;                       // = X I D
		prev_r = take ? rp : prev_r;
		uint32_t const tw2 = take ? (tw & (tw - 1u)) : tw;
		bool const adv = tw2 == 0;
		rbase += adv ? popc32(nw) : 0u;
		w += adv ? 1u : 0u;
		uint32_t const wl = umin(w, (uint32_t) NW - 1u);
		uint32_t const nw_n = L.cmp[wl * ST], tw_n = L.xm[wl * ST];
		nw = adv ? nw_n : nw;
		tw = adv ? ((w < w_end) ? tw_n : 0u) : tw2;
	}
	push(prev, prev != W_NONE ? rbase - prev_r : 0u);
	if (EV) while (bi < n_brk) push_pass();                                       // pass-through ops after the last column
	n_ops += (carry_op != W_NONE) ? 1u : 0u;                                      // cigar_buffer::finish (cigar.cc:89-95)
	adjacent_id = adj;
	return n_ops;
}

// Everything between "start column known" and "span masks built" for one record or segment.
struct SpanIn {
	const uint64_t *Sp, *Dp;    // 64-bit words of both rows holding the start column
	uint32_t c_lo6;             // start column & 63
	uint64_t src_left;          // source bases at or after the start column
	uint64_t e_start;           // target bases before the start column (rank0_dst)
	uint64_t ref_n;             // target contig length
	uint32_t nq;                // query bases of the record
	uint32_t q_start;           // query bases consumed before this segment (0 for a whole record)
};

struct SpanState {
	uint32_t tot_ref, span_begin, span_end, w_end, n_brk, n_dbase, n_qbase;
	bool     simple;
};

// 0: masks built; 1: the record must go to the general kernel; 2 (EV = false only): the record has events — it goes to the
// list of the second launch.
template <typename C, bool EV>
PV3_HD int build_span(SpanIn const &I, OpRange const &R, Lane <C> const &L, SpanState &st)
{
	constexpr int ST = C::ST, NW = C::NW, NBRK = C::NBRK, OUTQ = C::OUTQ; (void) ST; (void) NW; (void) NBRK; (void) OUTQ;
	OpsInfo const oi = scan_ops(R);
	st.tot_ref = oi.tot_ref; st.simple = oi.simple; st.n_brk = 0;
	st.span_begin = st.span_end = st.w_end = st.n_dbase = 0; st.n_qbase = oi.tot_q;
	if (oi.bad) return 1;
	if ((uint64_t) oi.tot_ref > I.src_left || oi.tot_ref > (uint32_t) NW * 32u) return 1;            // POS_OOB or a long span
	if ((uint64_t) I.q_start + oi.tot_q > (uint64_t) I.nq || I.nq > C::QMAX) return 1;                     // QUERY_OOB is possible
	if (!EV && !oi.simple) return 2;
	if (!load_span <C>(I.Sp, I.Dp, I.c_lo6, oi.tot_ref, L, st.span_begin, st.span_end, st.n_dbase)) return 1;
	if (I.e_start + st.n_dbase > I.ref_n) return 1;                                                    // REF_OOB is possible
	st.w_end = (st.span_end + 31u) >> 5;
	if (EV && !oi.simple && !apply_events <C>(R, oi.tot_ref, st.span_begin, st.span_end, L, st.n_brk)) return 1;
	return 0;
}

} // namespace k1w
} // namespace pv3
