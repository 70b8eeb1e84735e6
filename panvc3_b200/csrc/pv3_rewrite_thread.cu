// panvc3_b200 — kernel K1T: projection + CIGAR rewrite + run detection, ONE THREAD PER RECORD, word-parallel (sm_100a).
//
// Same semantics as the warp-per-record kernel in pv3_rewrite.cu (reference: include/panvc3/rewrite_cigar.hh:19-157,
// libpanvc3/cigar.cc:61-95, libpanvc3/indel_run_checker.cc:31-121, libpanvc3/alignment_projector.cc:61-73).
// Every MSA column of the record's span yields at most one unit op, decided by (source bit, target bit, op class):
//     src base & dst base   → compare run ('=' / 'X')          src base & dst gap → 'I'
//     src gap  & dst base   → 'D' (leading deletion)           both gaps          → nothing
//
// v4 (this file): the ncu profile of v3 (profiles/r1_v3_*) showed 451 warp instructions per record at 11 of 32 active
// lanes — nested per-record loops with four inlined copies of the run-length emitter diverged badly. The per-record
// work is now three FLAT loops, each with one body that every lane executes in lock-step and one store site:
//   C1  one iteration per column run (maximal run of equal (src bit, dst bit) inside a 64-column word) or input op:
//       → raw run list (u16: op | len << 4, or a reference to a pass-through input op) and compare-piece list
//   C2  one iteration per 8 bases of a compare piece: XOR of staged query ranks and target codes → mismatch bits in
//       query coordinates
//   C3  one iteration per output run: raw runs, '=' runs split by the mismatch bits, one cigar_buffer::push_back site
//
// v5: query ranks and target codes are staged as 4-bit nibbles (the query is packed while it is staged; the index holds
// the target codes packed, RefDesc::codes4), so that one compare step covers 16 bases and the tile needs half the shared
// memory (4 CTAs per SM instead of 3); the column walk runs on 32-bit window-relative coordinates with soft exits.
//
// Tile pipeline (32 threads = 32 consecutive records; measured: 128, 64, 32 → 0.433, 0.425, 0.421 ms per 1M records):
//   A  each thread: ids, src_pos, select0/rank0 of its record → c0, e0
//   B  cooperative, coalesced staging into shared memory: the tile's query bytes (contiguous in the batch) packed to
//      nibbles, and per record a 224-base window of packed target codes at e0
//   C1..C3 as above; run detection on the final ops (only when an I abuts a D)
//   D  block scan of op counts → one atomic on the arena cursor → ops and per-record results written
// Records that do not fit the fixed quotas (long reads, too many runs, windows beyond the staged data, malformed index,
// query ranks above 7) are appended to a deferred list and handled by the warp-per-record kernel afterwards — results
// are bit-identical.
#include "pv3_device.cuh"
#include "pv3_kernels.h"
#include "pv3_runs.cuh"
#include <cstdlib>
#include <cstdio>

namespace pv3 {

constexpr int      TT        = 32;             // threads per CTA = records per tile (one warp: the tile barriers cost nothing)
constexpr int      T_OUTQ    = 24;             // staged output ops per record
// per-variant quotas (Cfg below): raw runs per record (column runs + pass-through ops), compare pieces per record, query
// bases staged per tile
constexpr uint32_t T_RWIN    = 192;            // target bases staged per record (6 x 16 bytes of nibbles)
constexpr uint32_t T_RCH     = T_RWIN / 32;    // 16-byte chunks per staged target window
constexpr uint32_t T_MMW     = 6;              // mismatch bit words per record (query coordinates): reads up to 192 bases
constexpr uint32_t T_C1_ITERS = 192;           // bound of the column walk (long gaps / deletions go to the warp kernel)
constexpr uint32_t RAW_PASS  = 0x8000u;        // raw entry refers to input op (entry & 0x7fff)
constexpr uint32_t ERR_DEFER = 0x80000000u;

// Two layouts, chosen at launch (PV3_K1T_STAGE_REF, default 0):
//  STAGE = true  (v5-v8): the target windows are staged in shared memory; 10.3 KB per one-warp tile ⇒ 20 CTAs per SM. `fin`
//                reuses the memory of `ref` (dead after phase C2; a barrier separates them).
//  STAGE = false (v9): phase C2 reads the packed target codes through L1 (phase A prefetches the record's window), and
//                `fin` reuses the staged query and the piece list (dead after C2): 7.0 KB per tile ⇒ 28 CTAs per SM. The
//                kernel is latency-bound at 5 warps per scheduler (profiles/r1_v8_ncu_k1t_full_summary.txt), so resident
//                warps are what buys time.
template <int V> struct Cfg;
template <> struct Cfg <0> { static constexpr bool STAGE = true;  static constexpr int RAWQ = 32, PIECEQ = 12; static constexpr uint32_t QSTAGE = TT * 168; };
template <> struct Cfg <1> { static constexpr bool STAGE = false; static constexpr int RAWQ = 32, PIECEQ = 12; static constexpr uint32_t QSTAGE = TT * 168; };
template <> struct Cfg <2> { static constexpr bool STAGE = false; static constexpr int RAWQ = 24, PIECEQ = 10; static constexpr uint32_t QSTAGE = TT * 156; };   // 6.1 KB ⇒ 32 CTAs per SM
template <> struct Cfg <3> { static constexpr bool STAGE = true;  static constexpr int RAWQ = 24, PIECEQ = 10; static constexpr uint32_t QSTAGE = TT * 156; };   // 9.3 KB ⇒ 22 CTAs per SM
template <> struct Cfg <4> { static constexpr bool STAGE = false; static constexpr int RAWQ = 24, PIECEQ = 10; static constexpr uint32_t QSTAGE = TT * 156; };   // as 2, no L1 prefetch

template <int V, bool STAGE = Cfg <V>::STAGE> struct TileSmemT;
template <int V> struct TileSmemT <V, true> {
	uint16_t raw[Cfg <V>::RAWQ][TT];
	uint32_t piece[Cfg <V>::PIECEQ][TT];       // qpos | (target offset from the staged window start) << 8 | len << 24
	uint32_t mm[T_MMW + 1][TT];
	uint2    q4[Cfg <V>::QSTAGE / 16 + 2];     // packed query: 16 bases per element
	union {
		uint4    ref[TT][T_RCH];               // [record][staged packed target codes]
		uint32_t fin[T_OUTQ][TT];              // final ops
	};
	const uint8_t  *ref_ptr[TT];               // staging sources published by phase A
	uint32_t scan[TT / 32];
	unsigned long long base;
};
template <int V> struct TileSmemT <V, false> {
	uint16_t raw[Cfg <V>::RAWQ][TT];
	uint32_t mm[T_MMW + 1][TT];
	union {
		struct {
			uint2    q4[Cfg <V>::QSTAGE / 16 + 2];
			uint32_t piece[Cfg <V>::PIECEQ][TT];
		};
		uint32_t fin[T_OUTQ][TT];
	};
	uint32_t scan[TT / 32];
	unsigned long long base;
};
static_assert(sizeof(uint4) * T_RCH * TT >= sizeof(uint32_t) * T_OUTQ * TT, "fin must fit in ref");
static_assert(sizeof(uint2) * (Cfg <2>::QSTAGE / 16 + 2) + sizeof(uint32_t) * Cfg <2>::PIECEQ * TT >= sizeof(uint32_t) * T_OUTQ * TT, "fin must fit in q4 + piece");

__device__ __forceinline__ uint32_t ctz64(uint64_t x) { return (uint32_t) __ffsll((long long) x) - 1u; }   // x != 0
__device__ __forceinline__ uint32_t ctz32(uint32_t x) { return (uint32_t) __popc(~x & (x - 1u)); }             // 32 for x == 0

// 8 nibbles of t → 8 bits (bit i set ⇔ nibble i non-zero), returned in bits 24..31
__device__ __forceinline__ uint32_t nonzero_nibbles_hi(uint32_t t)
{
	uint32_t const nz = (((t & 0x77777777u) + 0x77777777u) | t) & 0x88888888u;     // bit 4i+3
	uint32_t const y = nz >> 3;                                                      // bit 4i
	uint32_t const y1 = (y | (y >> 3)) & 0x03030303u;                               // bits 8j, 8j+1
	return y1 * 0x01041040u;                                                         // byte j → bits 24+2j, 25+2j (no carries)
}

// 16 nibbles at nibble offset a of a word array in shared memory (reads 3 words)
__device__ __forceinline__ void load16n(const uint32_t *base, uint32_t a, uint32_t &lo, uint32_t &hi)
{
	const uint32_t *p = base + (a >> 3);
	uint32_t const sh = (a & 7u) * 4u;
	uint32_t const w0 = p[0], w1 = p[1], w2 = p[2];
	lo = __funnelshift_r(w0, w1, sh);
	hi = __funnelshift_r(w1, w2, sh);
}

// the same through the read-only path (word-aligned base in global memory)
__device__ __forceinline__ void load16n_ldg(const uint32_t *base, uint32_t a, uint32_t &lo, uint32_t &hi)
{
	const uint32_t *p = base + (a >> 3);
	uint32_t const sh = (a & 7u) * 4u;
	uint32_t const w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);
	lo = __funnelshift_r(w0, w1, sh);
	hi = __funnelshift_r(w1, w2, sh);
}

// the same from the packed codes in global memory (512 readable pad bytes); rare: windows beyond the staged data
__device__ __noinline__ uint2 load16n_global(const uint8_t *codes4, uint64_t nib)
{
	const uint8_t *p = codes4 + (nib >> 1);
	uint64_t v = 0;
	#pragma unroll
	for (int i = 0; i < 8; ++i) v |= (uint64_t) __ldg(p + i) << (8 * i);
	if (nib & 1u) v = (v >> 4) | ((uint64_t) __ldg(p + 8) << 60);
	return make_uint2((uint32_t) v, (uint32_t) (v >> 32));
}

// 16 query ranks (one per byte) → 16 nibbles; `bad` collects bits 3..7 of every byte (ranks the nibble compare cannot hold)
__device__ __forceinline__ uint2 pack16(uint4 v, uint32_t &bad)
{
	bad |= (v.x | v.y | v.z | v.w) & 0xf8f8f8f8u;
	uint32_t const a = (v.x | (v.x >> 4)) & 0x00ff00ffu, b = (v.y | (v.y >> 4)) & 0x00ff00ffu;
	uint32_t const c = (v.z | (v.z >> 4)) & 0x00ff00ffu, d = (v.w | (v.w >> 4)) & 0x00ff00ffu;
	return make_uint2(__byte_perm(a, b, 0x6420), __byte_perm(c, d, 0x6420));
}

// 8 BAM 4-bit codes (first base in the HIGH nibble of each byte) → 8 dna5 ranks, first base in the lowest nibble.
// One-hot codes 1 2 4 8 (A C G T) become code >> 1 = 0 1 2 4; every other code becomes 3 (N), see bam4_to_rank.
__device__ __forceinline__ uint32_t bam4_word_to_ranks(uint32_t w)
{
	uint32_t const x = ((w & 0x0f0f0f0fu) << 4) | ((w >> 4) & 0x0f0f0f0fu);
	uint32_t const cnt = x - ((x >> 1) & 0x77777777u) - ((x >> 2) & 0x33333333u) - ((x >> 3) & 0x11111111u);   // bits set per nibble
	uint32_t const t = cnt ^ 0x11111111u;
	uint32_t const one = (~(((t & 0x77777777u) + 0x77777777u) | t) & 0x88888888u) >> 3;                           // 1 where exactly one bit
	uint32_t const m = one * 15u;
	return ((x >> 1) & 0x77777777u & m) | (0x33333333u & ~m);
}

template <int V>
__global__ void __launch_bounds__(TT)
k1_thread_kernel(K1Params P)
{
	constexpr bool STAGE = Cfg <V>::STAGE;
	constexpr int T_RAWQ = Cfg <V>::RAWQ, T_PIECEQ = Cfg <V>::PIECEQ;
	constexpr uint32_t T_QSTAGE = Cfg <V>::QSTAGE;
	extern __shared__ __align__(16) uint8_t smem_raw[];
	TileSmemT <V> &sm = *reinterpret_cast <TileSmemT <V> *>(smem_raw);
	unsigned const tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	uint64_t const n_tiles = (P.n + TT - 1) / TT;

	for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
		uint64_t const t0 = tile * TT;
		uint32_t const n_here = (uint32_t) ((P.n - t0 < (uint64_t) TT) ? (P.n - t0) : TT);
		uint64_t const rec = t0 + tid;
		bool const active = tid < n_here;

		// The tile's query bytes are contiguous in the batch: [seq_off[t0], seq_off[t0 + n_here])
		uint64_t const tq0 = __ldg(P.seq_off + t0), tq1 = __ldg(P.seq_off + t0 + n_here);
		uint64_t const q_addr = (uint64_t) (uintptr_t) P.seq + tq0;
		uint64_t const q_addr16 = q_addr & ~15ull;
		uint32_t const q_lead = (uint32_t) (q_addr - q_addr16);
		bool const packed = P.seq_len != nullptr;                           // BAM 4-bit input: offsets count bytes, two bases each
		bool const tile_fits = ((tq1 - tq0 + q_lead) << (packed ? 1 : 0)) <= (uint64_t) T_QSTAGE;

		// ---- phase B1: the tile's query, packed to nibbles while it is staged (issued first: the loads overlap phase A) ----
		uint32_t bad_rank = 0;
		if (tile_fits) {
			uint32_t const n16 = (uint32_t) ((tq1 - tq0 + q_lead + 15) / 16);
			const uint4 *g = reinterpret_cast <const uint4 *>((uintptr_t) q_addr16);
			if (!packed) for (uint32_t i = tid; i < n16; i += TT) sm.q4[i] = pack16(__ldg(g + i), bad_rank);
			else for (uint32_t i = tid; i < n16; i += TT) {
				uint4 const v = __ldg(g + i);
				sm.q4[2 * i] = make_uint2(bam4_word_to_ranks(v.x), bam4_word_to_ranks(v.y));
				sm.q4[2 * i + 1] = make_uint2(bam4_word_to_ranks(v.z), bam4_word_to_ranks(v.w));
			}
		}

		// ---- phase A: per-record projection of the start position ------------------------------------------------
		uint32_t flags = 0;
		bool defer = !tile_fits;
		RowDesc S = {}, D = {};
		RefDesc R = {};
		uint64_t c0 = 0, e0 = 0, spos = 0, qoff = 0, nq = 0, cbase = 0;
		uint32_t nc = 0, first_op = 0;
		if constexpr (STAGE) sm.ref_ptr[tid] = nullptr;
		if (active) {
			uint32_t const sr = __ldg(P.src_row + rec), dr = __ldg(P.dst_row + rec), rf = __ldg(P.dst_ref + rec);
			if (sr >= P.n_rows || dr >= P.n_rows || rf >= P.n_refs) flags = PV3_FLAG_BAD_ID;
			else {
				S = P.rows[sr]; D = P.rows[dr]; R = P.refs[rf];
				qoff = __ldg(P.seq_off + rec);
				nq = packed ? (uint64_t) __ldg(P.seq_len + rec) : __ldg(P.seq_off + rec + 1) - qoff;
				if (P.qual_off) {
					uint64_t const ql = __ldg(P.qual_off + rec + 1) - __ldg(P.qual_off + rec);
					if (ql && ql != nq) flags |= PV3_FLAG_BAD_QUAL;
				}
				cbase = __ldg(P.cig_off + rec);
				nc = (uint32_t) (__ldg(P.cig_off + rec + 1) - cbase);
				if (nc) first_op = __ldg(P.cig + cbase);                             // in flight while select0/rank0 run
				spos = __ldg(P.src_pos + rec);
				c0 = (spos < S.n_zeros) ? select0_thread(S, spos + 1) : S.n_bits;     // rewrite_cigar.hh:51
				e0 = rank0_thread(D, c0);                                              // :52
				if (S.n_bits != D.n_bits || e0 > R.n || nq > 32u * T_MMW || nc >= RAW_PASS) defer = true;   // unusual: the general kernel guards it
				if (!defer) {
					const uint8_t *rp = R.codes4 + ((e0 & ~31ull) >> 1);
					if constexpr (STAGE) sm.ref_ptr[tid] = rp;
					else if (V != 4) {                                                   // the window C2 will read: 96 bytes from a 16-byte boundary
						asm volatile("prefetch.global.L1 [%0];" :: "l"(rp));
						asm volatile("prefetch.global.L1 [%0];" :: "l"(rp + 95));
					}
				}
			}
		}
		if (P.mate_row && active) {                                              // sequence_entry::project_position (msa_index.hh:253-258), as positions_kernel
			uint32_t const mr = __ldg(P.mate_row + rec), dr = __ldg(P.dst_row + rec);
			uint64_t mo = ~0ull;
			if (mr < P.n_rows && dr < P.n_rows) mo = rank0_thread(P.rows[dr], select0_thread(P.rows[mr], __ldg(P.mate_pos + rec) + 1));
			P.mate_dst_pos[rec] = mo;
		}
		// a query rank above 7 anywhere in the tile: the nibble compare cannot represent it, the general kernel takes the tile
		if (__syncthreads_or((int) bad_rank)) defer = true;

		// ---- phase B2: per record a window of packed target codes (8 lanes per record, 16 bytes per lane) --------------
		if constexpr (STAGE) for (uint32_t r = warp * 32 + (lane >> 3); r < warp * 32 + 32; r += 4) {
			const uint8_t *rp = sm.ref_ptr[r];
			uint32_t const l8 = lane & 7u;
			if (rp && l8 < T_RCH) sm.ref[r][l8] = __ldg(reinterpret_cast <const uint4 *>(rp) + l8);
		}

		// ---- phase C1: one flat walk over the column runs of the record's span -------------------------------------
		// Output: raw runs ('=' stands for a compare run of unknown content) and the compare pieces. Coordinates are
		// 32-bit: columns relative to the word of c0, target offsets relative to the staged window. Errors do not leave
		// the loop through a branch: they end the op list (`k = nc, rem = 0`), so every lane runs the same straight body.
		uint32_t n_raw = 0, n_piece = 0, n_ops = 0, n_runs = 0;
		unsigned long long job_base = 0;
		uint64_t const e0a = e0 & ~31ull;
		bool const work = active && !flags && !defer;
		if (work) {
			const uint64_t *Sp = S.words + (c0 >> 6), *Dp = D.words + (c0 >> 6);
			uint64_t const r_room = R.n - e0a;                                   // e0 <= R.n
			uint32_t const r_lim = r_room > 0x10000ull ? 0x10000u : (uint32_t) r_room, nq32 = (uint32_t) nq;
			uint64_t const left64 = (spos < S.n_zeros) ? (S.n_zeros - spos) : 0;
			uint32_t src_left = left64 > 0xffffffffull ? 0xffffffffu : (uint32_t) left64;   // op counts are < 2^28
			uint32_t crel = (uint32_t) c0 & 63u, ra = (uint32_t) (e0 - e0a), qpos = 0, k = 0, rem = 0, err = 0, wcur = ~0u;
			uint64_t Sinv = 0, Dinv = 0;
			uint64_t Snext = __ldg(Sp), Dnext = __ldg(Dp);                      // the walk starts in word 0 of the window; one word ahead
			bool mlike = false;
			// The open run lives in registers AND in the first free slot of the lists (rewritten whenever it grows), so that
			// closing it is a counter increment: no flush branch that a few lanes take in every iteration.
			uint32_t pend_op = OP_NONE, pend_len = 0, pend_q = 0, pend_ra = 0;
			auto close_run = [&]() {
				n_raw += (pend_op != OP_NONE) ? 1u : 0u;
				n_piece += (pend_op == OP_EQ) ? 1u : 0u;
			};
			for (uint32_t it = 0; ; ++it) {
				if (it >= T_C1_ITERS) { err = ERR_DEFER; break; }
				bool walk = true;
				if (rem == 0) {                                                    // next input op
					if (k == nc) break;
					uint32_t const packed = k ? __ldg(P.cig + cbase + k) : first_op;
					uint32_t const cnt = packed >> 4, op = packed & 0xfu;
					walk = false;
					if (op > OP_X) { err = PV3_FLAG_BAD_OP; k = nc; }                 // rewrite_cigar.hh:149-151
					else if ((0x72u >> op) & 1u) {                                     // I S H P pass through (:68-77)
						close_run(); pend_op = OP_NONE;
						sm.raw[min(n_raw, (uint32_t) T_RAWQ - 1u)][tid] = (uint16_t) (RAW_PASS | k);   // a clamped slot is only hit when the list overflows (⇒ deferred)
						++n_raw;
						if (op == OP_I || op == OP_S) qpos = min(qpos + cnt, 0x10000u);
						++k;
					} else {
						mlike = (op != OP_D && op != OP_N);
						if (cnt > src_left) { err = PV3_FLAG_POS_OOB; k = nc; }
						else { src_left -= cnt; rem = cnt; ++k; walk = cnt != 0; }
					}
				}
				if (walk) {
					uint32_t const wi = crel >> 6, b = crel & 63u;
					if (wi != wcur) {                                               // next word pair is already in flight
						Sinv = ~Snext; Dinv = ~Dnext; wcur = wi;
						Snext = __ldg(Sp + wi + 1); Dnext = __ldg(Dp + wi + 1);
					}
					uint64_t const s = Sinv >> b, d = Dinv >> b;
					uint32_t const sb = (uint32_t) s & 1u, db = (uint32_t) d & 1u;
					// columns until either row changes: trailing zeros of (s ^ -sb) | (d ^ -db), in 32-bit halves and without a
					// branch (ctz32(0) = 32, so an all-zero word gives 64)
					uint32_t const ms = 0u - sb, md = 0u - db;
					uint32_t const nlo = ((uint32_t) s ^ ms) | ((uint32_t) d ^ md), nhi = ((uint32_t) (s >> 32) ^ ms) | ((uint32_t) (d >> 32) ^ md);
					uint32_t len = ctz32(nlo ? nlo : nhi) + (nlo ? 0u : 32u);         // one population count: the half that holds the change
					len = min(len, 64u - b);
					if (sb) { len = min(len, rem); rem -= len; }
					crel += len;
					bool const isq = sb && mlike;                                   // the run consumes query bases
					bool const cmp = isq && db;                                     // compare run (:102-113)
					uint32_t e1 = 0;
					if (cmp) {
						if (qpos + len > nq32) e1 = PV3_FLAG_QUERY_OOB;
						else if (ra + len > r_lim) e1 = (r_room > 0x10000ull) ? ERR_DEFER : PV3_FLAG_REF_OOB;
					}
					if (e1) { err = e1; k = nc; rem = 0; }                         // ends the op list; nothing is emitted for this run
					uint32_t const op = db ? (isq ? OP_EQ : OP_D) : OP_I;         // :92-100, :127-137
					bool const emit = (db || isq) && !e1;
					// a run that continues the pending one (across a word boundary or over columns that are gaps in both rows)
					// extends it: fewer raw runs and compare pieces for the next two loops
					bool const ext = op == pend_op && pend_len + len <= 250u;
					bool const fresh = emit && !ext;
					if (fresh) close_run();
					pend_op = fresh ? op : pend_op; pend_q = fresh ? qpos : pend_q; pend_ra = fresh ? ra : pend_ra;
					pend_len = (fresh ? 0u : pend_len) + (emit ? len : 0u);
					// (a clamped slot is only written while a run is open, and closing that run overflows the list ⇒ deferred)
					if (pend_op != OP_NONE) sm.raw[min(n_raw, (uint32_t) T_RAWQ - 1u)][tid] = (uint16_t) ((pend_len << 4) | pend_op);
					if (pend_op == OP_EQ) sm.piece[min(n_piece, (uint32_t) T_PIECEQ - 1u)][tid] = pend_q | (pend_ra << 8) | (pend_len << 24);
					if (isq) qpos = min(qpos + len, 0x10000u);
					if (db) ra = min(ra + len, 0x20000u);
				}
			}
			close_run();
			if (err == ERR_DEFER || n_raw > (uint32_t) T_RAWQ || n_piece > (uint32_t) T_PIECEQ) defer = true;
			else flags |= err;
		}
		__syncthreads();        // staged query and target windows are complete

		// ---- phase C2: one flat loop over the compare pieces, 16 bases per step: mismatch bits in query coordinates ------
		bool const work2 = work && !flags && !defer;
		if (work2) {
			const uint32_t *qs = reinterpret_cast <const uint32_t *>(sm.q4);
			const uint32_t *rs;                 // the record's target window: staged, or the packed codes themselves (16-byte aligned, 512 readable pad bytes)
			if constexpr (STAGE) rs = reinterpret_cast <const uint32_t *>(sm.ref[tid]);
			else rs = reinterpret_cast <const uint32_t *>(R.codes4 + (e0a >> 1));
			uint32_t const qrel = (q_lead + (uint32_t) (qoff - tq0)) << (packed ? 1 : 0);
			#pragma unroll
			for (uint32_t w = 0; w <= T_MMW; ++w) sm.mm[w][tid] = 0;
			if constexpr (STAGE) {
				uint32_t j = 0, remaining = 0, qpos = 0, ra = 0;
				for (;;) {
					if (!remaining) {
						if (j == n_piece) break;
						uint32_t const v = sm.piece[j][tid]; ++j;
						qpos = v & 0xffu; ra = (v >> 8) & 0xffffu; remaining = v >> 24;
					}
					uint32_t const n = min(remaining, 16u);
					uint32_t qlo, qhi, rlo, rhi;
					load16n(qs, qrel + qpos, qlo, qhi);
					if (ra + 24u <= T_RWIN) load16n(rs, ra, rlo, rhi);
					else { uint2 const g = load16n_global(R.codes4, e0a + ra); rlo = g.x; rhi = g.y; }   // beyond the staged window
					uint32_t bits = (nonzero_nibbles_hi(qlo ^ rlo) >> 24) | ((nonzero_nibbles_hi(qhi ^ rhi) >> 16) & 0xff00u);
					bits &= (1u << n) - 1u;
					uint32_t const w = qpos >> 5, b = qpos & 31u;
					uint64_t const placed = (uint64_t) bits << b;
					sm.mm[w][tid] |= (uint32_t) placed;
					sm.mm[w + 1][tid] |= (uint32_t) (placed >> 32);
					qpos += n; ra += n; remaining -= n;
				}
			} else {
				// The target words come from global memory (L1 when the target is small, L2 or HBM at chromosome scale), so
				// the loop is rotated: the three words of step i+1 are requested before step i is compared.
				uint32_t j = 0, rem_n = 0, qpos_n = 0, ra_n = 0, w0 = 0, w1 = 0, w2 = 0;
				bool have = false;
				auto request = [&]() {                                              // next step: position and loads
					if (!rem_n) {
						have = j != n_piece;
						if (have) { uint32_t const v = sm.piece[j][tid]; ++j; qpos_n = v & 0xffu; ra_n = (v >> 8) & 0xffffu; rem_n = v >> 24; }
					}
					if (have) { const uint32_t *p = rs + (ra_n >> 3); w0 = __ldg(p); w1 = __ldg(p + 1); w2 = __ldg(p + 2); }
				};
				request();
				while (have) {
					uint32_t const n = min(rem_n, 16u), qpos = qpos_n, sh = (ra_n & 7u) * 4u;
					uint32_t const rlo = __funnelshift_r(w0, w1, sh), rhi = __funnelshift_r(w1, w2, sh);
					qpos_n += n; ra_n += n; rem_n -= n;
					request();
					uint32_t qlo, qhi;
					load16n(qs, qrel + qpos, qlo, qhi);
					uint32_t bits = (nonzero_nibbles_hi(qlo ^ rlo) >> 24) | ((nonzero_nibbles_hi(qhi ^ rhi) >> 16) & 0xff00u);
					bits &= (1u << n) - 1u;
					uint32_t const w = qpos >> 5, b = qpos & 31u;
					uint64_t const placed = (uint64_t) bits << b;
					sm.mm[w][tid] |= (uint32_t) placed;
					sm.mm[w + 1][tid] |= (uint32_t) (placed >> 32);
				}
			}
		}

		__syncthreads();        // the staged target windows are dead from here on: fin overwrites them

		// ---- phase C3: final ops — one flat loop, one cigar_buffer::push_back (cigar.cc:61-86) site ----------------------
		unsigned long long cells = 0, path = 0, pool = 0, max_pool = 0;
		uint32_t n_big = 0;
		bool affine = false;
		if (work2) {
			uint32_t j = 0, left = 0, qpos = 0, carry_op = OP_NONE, carry_cnt = 0;
			bool adjacent_id = false;
			for (;;) {
				uint32_t op = OP_NONE, len = 0;
				if (!left) {
					if (j == n_raw) break;
					uint32_t const v = sm.raw[j][tid]; ++j;
					if (v & RAW_PASS) {
						uint32_t const packed = __ldg(P.cig + cbase + (v & 0x7fffu));
						op = packed & 0xfu; len = packed >> 4;
						if (op == OP_I || op == OP_S) qpos = min(qpos + len, 32u * T_MMW);   // beyond |SEQ|: no compare can follow
					} else {
						op = v & 0xfu; len = v >> 4;
						if (op == OP_EQ) left = len;
						else if (op == OP_I) qpos += len;
					}
				}
				if (left) {                                                         // split the compare run at the next change
					uint32_t const w = qpos >> 5, b = qpos & 31u;
					uint32_t const m0 = sm.mm[w][tid], m1 = sm.mm[w + 1][tid], m2 = (w + 2 <= T_MMW) ? sm.mm[w + 2][tid] : 0u;
					uint32_t const lo = __funnelshift_r(m0, m1, b), hi = __funnelshift_r(m1, m2, b);
					uint32_t const first = lo & 1u;
					uint32_t const xlo = lo ^ (0u - first), xhi = hi ^ (0u - first);
					uint32_t l = ctz32(xlo ? xlo : xhi) + (xlo ? 0u : 32u);             // 64 when neither half changes
					l = min(l, left);
					op = first ? OP_X : OP_EQ; len = l;
					left -= l; qpos += l;
				}
				// cigar_buffer::push_back without a branch: the open op is kept in its slot (rewritten as it grows); a different
				// op closes it by advancing the count
				bool const fresh = len != 0 && carry_op != op;
				adjacent_id |= fresh && (carry_op | op) == 3u;                         // I next to D (N is never produced)
				n_ops += (fresh && carry_op != OP_NONE) ? 1u : 0u;
				carry_cnt = (fresh ? 0u : carry_cnt) + len;
				carry_op = fresh ? op : carry_op;
				sm.fin[min(n_ops, (uint32_t) T_OUTQ - 1u)][tid] = (carry_cnt << 4) | carry_op;   // clamped only when the list overflows (⇒ deferred)
			}
			n_ops += (carry_op != OP_NONE) ? 1u : 0u;                                // cigar_buffer::finish (cigar.cc:89-95)
			if (n_ops > (uint32_t) T_OUTQ) defer = true;
			if (!defer && P.stage == 0) {
				if (n_ops == 0) flags |= PV3_FLAG_EMPTY_CIGAR;                                   // cigar.cc:16-18 UB
				else if (adjacent_id) {
					bool has_qual = false;
					if (P.qual_off) has_qual = __ldg(P.qual_off + rec + 1) > __ldg(P.qual_off + rec);
					affine = has_qual || P.gap_open != 0;
					n_runs = scan_runs <false>(&sm.fin[0][tid], TT, n_ops, e0, rec, nq, R.n, nullptr, nullptr, flags, cells, path, pool, max_pool, affine, false, &n_big);
					if (flags) n_runs = 0;
				}
			}
		}
		// Runs found: one set of atomics per WARP (per-record atomics on six shared counters serialise in L2 when most
		// records realign), then each record writes its jobs and reported ranges behind the warp's base.
		if (__any_sync(FULL, n_runs != 0)) {
			uint32_t incl = n_runs;
			unsigned long long w_cells = n_runs ? cells : 0, w_path = n_runs ? path : 0, w_pool = n_runs ? pool : 0, w_max = n_runs ? max_pool : 0;
			uint32_t const w_recs = __popc(__ballot_sync(FULL, n_runs != 0));
			uint32_t w_bigjobs = n_runs ? n_big : 0u;                  // runs the thread-per-run DP kernel leaves to the warp kernel
			#pragma unroll
			for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
			#pragma unroll
			for (int d = 16; d >= 1; d >>= 1) {
				w_cells += shfl_u64(w_cells, (int) (lane ^ d)); w_path += shfl_u64(w_path, (int) (lane ^ d)); w_pool += shfl_u64(w_pool, (int) (lane ^ d));
				unsigned long long const o = shfl_u64(w_max, (int) (lane ^ d)); if (o > w_max) w_max = o;
				w_bigjobs += __shfl_xor_sync(FULL, w_bigjobs, d);
			}
			uint32_t const w_runs = __shfl_sync(FULL, incl, 31);
			unsigned long long wbase = 0;
			if (lane == 0) {
				if (w_bigjobs) atomicAdd(&P.counters->n_big_jobs, (unsigned long long) w_bigjobs);
				wbase = atomicAdd(&P.counters->job_cursor, (unsigned long long) w_runs);
				atomicAdd(&P.counters->n_realigned, (unsigned long long) w_recs);
				atomicAdd(&P.counters->dp_cells, w_cells);
				atomicAdd(&P.counters->dp_path, w_path);
				atomicAdd(&P.counters->dp_pool, w_pool);
				atomicMax(&P.counters->dp_max_pool, w_max);
				if (wbase + w_runs > P.rr_cap) atomicOr(&P.counters->overflow, 2u);
			}
			wbase = shfl_u64(wbase, 0);
			if (n_runs) {
				job_base = wbase + (incl - n_runs);
				if (wbase + w_runs <= P.rr_cap) {
					uint32_t f = 0; unsigned long long a = 0, b = 0, c = 0, d = 0;
					scan_runs <true>(&sm.fin[0][tid], TT, n_ops, e0, rec, nq, R.n, P.jobs + job_base, P.rr + 4 * job_base, f, a, b, c, d, affine, true);
				}
			}
		}
		if (active && defer && !flags) {
			unsigned long long const slot = atomicAdd(&P.counters->n_deferred, 1ull);
			P.deferred[slot] = rec;
		}

		// ---- phase D: block scan, one arena allocation per tile, write-out ----------------------------------------------
		uint32_t const cnt = (active && !defer && !flags) ? n_ops : 0u;
		uint32_t incl = cnt;
		#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
		if (lane == 31) sm.scan[warp] = incl;
		__syncthreads();
		uint32_t warp_off = 0, total = 0;
		#pragma unroll
		for (int w = 0; w < TT / 32; ++w) { uint32_t const v = sm.scan[w]; if (w < (int) warp) warp_off += v; total += v; }
		if (tid == 0) {
			unsigned long long base = 0;
			if (total) {
				base = atomicAdd(&P.counters->arena_cursor, (unsigned long long) total);
				if (base + total > P.arena_cap) { atomicOr(&P.counters->overflow, 1u); base = ~0ull; }
			}
			sm.base = base;
		}
		__syncthreads();
		unsigned long long const base = sm.base;
		uint32_t const my_off = warp_off + incl - cnt;
		if (active && (!defer || flags)) {
			if (flags) { n_runs = 0; atomicAdd(&P.counters->n_flagged, 1ull); }
			P.dst_pos[rec] = e0;
			P.flags[rec] = flags;
			P.rr_off[rec] = job_base + P.rr_base;
			P.rr_len[rec] = n_runs;
			P.cig_out_off[rec] = (base == ~0ull || flags) ? 0 : base + my_off + P.arena_off;
			P.cig_out_len[rec] = (base == ~0ull) ? 0 : cnt;
			if (base != ~0ull) for (uint32_t j = 0; j < cnt; ++j) P.arena[base + my_off + j] = sm.fin[j][tid];
		}
		__syncthreads();
	}
}

template <int V>
static void launch_variant(const K1Params &P, int n_sms, cudaStream_t stream)
{
	static int per_sm = 0;
	size_t const bytes = sizeof(TileSmemT <V>);
	if (!per_sm) {
		cudaFuncSetAttribute(k1_thread_kernel <V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes);
		if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_thread_kernel <V>, TT, bytes) != cudaSuccess || per_sm < 1) per_sm = 1;
		if (getenv("PV3_TRACE")) fprintf(stderr, "[pv3 trace] k1_thread_kernel variant %d: %zu bytes of shared memory per tile, %d CTAs per SM\n", V, bytes, per_sm);
	}
	uint64_t const n_tiles = (P.n + TT - 1) / TT;
	if (!n_tiles) return;
	uint64_t blocks = (uint64_t) n_sms * per_sm;  // persistent: every resident CTA strides over the tiles (shared memory bounds the count)
	if (blocks > n_tiles) blocks = n_tiles;
	k1_thread_kernel <V><<<(unsigned) blocks, TT, bytes, stream>>>(P);
}

// Layout choice. Reading the target codes through L1 (no staging, 32 CTAs per SM) wins while the packed target stays close
// to the SMs (L1 / the near part of L2); with a large target every tile's windows come from L2 / HBM and the staged layout
// wins, because its coalesced 16-byte loads are issued before the column walk and used after it. Measured on B200, 1 M
// 150-bp records in random order, staged (variant 0) against through-L1 (variant 4), ms per launch (profiles/r1_v12_*):
//   configs[1] (25 KB of codes) 0.335 / 0.277    10^5 columns (0.2 MB) 0.274 / 0.246    10^6 (2 MB) 0.292 / 0.302
//   10^7 (20 MB) 0.298 / 0.313                   10^8 (200 MB) 0.409 / 0.477 (0.504 without the L1 prefetch)
// PV3_K1T_VARIANT overrides: 0 = staged windows (v5-v8 layout, 20 CTAs per SM), 1 = through L1 with the old quotas (28),
// 2 = through L1 with smaller quotas (32) and an L1 prefetch of the window, 3 = staged with the smaller quotas (21),
// 4 = as 2 without the prefetch.
constexpr uint64_t L1_RESIDENT_TARGET_BYTES = 512 * 1024;   // between the 0.2 MB and 2 MB measurements

bool k1_thread_layout_is_tunable(uint64_t target_code_bytes) { return target_code_bytes > L1_RESIDENT_TARGET_BYTES && !getenv("PV3_K1T_VARIANT"); }

// layout: -1 = by target size; 0 = staged windows, 1 = through L1 — the caller's measured choice for a large target (with a
// large target the better layout depends on the ORDER of the records: staged wins when they are scattered over the genome,
// through-L1 when neighbours share windows, as in a coordinate-sorted BAM; configs[2], ms per 10^6 records, staged /
// through-L1: scattered 0.566 / 0.672, sorted window 0.317 / 0.226 — profiles/r2_v25_chr_layout_variants.json)
void launch_k1_thread(const K1Params &P, int n_sms, uint64_t target_code_bytes, cudaStream_t stream, int layout)
{
	const char *e = getenv("PV3_K1T_VARIANT");                          // read per launch: the parity tests switch layouts in-process
	int const forced = (e && *e >= '0' && *e <= '4') ? (*e - '0') : -1;
	int const by_size = target_code_bytes <= L1_RESIDENT_TARGET_BYTES ? 4 : 0;
	int const variant = forced >= 0 ? forced : (by_size == 4 || layout < 0) ? by_size : (layout == 0 ? 0 : 2);
	switch (variant) {
		case 0: launch_variant <0>(P, n_sms, stream); break;
		case 1: launch_variant <1>(P, n_sms, stream); break;
		case 2: launch_variant <2>(P, n_sms, stream); break;
		case 3: launch_variant <3>(P, n_sms, stream); break;
		default: launch_variant <4>(P, n_sms, stream); break;
	}
}

} // namespace pv3
