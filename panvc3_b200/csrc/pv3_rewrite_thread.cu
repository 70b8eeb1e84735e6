// panvc3_b200 — kernel K1T: projection + CIGAR rewrite + run detection, ONE THREAD PER RECORD, word-parallel (sm_100a).
//
// Same semantics as the warp-per-record kernel in pv3_rewrite.cu (reference: include/panvc3/rewrite_cigar.hh:19-157,
// libpanvc3/cigar.cc:61-95, libpanvc3/indel_run_checker.cc:31-121, libpanvc3/alignment_projector.cc:61-73) but the
// per-column work is done 64 columns at a time with bit operations on the source/target gap words:
//     src base & dst base   → compare run ('=' / 'X')          src base & dst gap → 'I'
//     src gap  & dst base   → 'D' (leading deletion)           both gaps          → nothing
// Maximal runs of one class inside a word are found with ctz; a compare run XORs 4 query ranks against 4 target codes per
// step (RefDesc::codes: equal byte ⇔ dna5::to_char(query) == raw target byte) and extracts the mismatch bits with SWAR
// arithmetic.  The ncu profile of the lane-per-column kernel (profiles/r1_v1_*) showed ≈1600 warp instructions per
// 150-bp record, i.e. instruction-bound at 3 % of the HBM roofline; here a record costs O(10²..10³) *thread*
// instructions, and 128 records are processed per CTA tile.
//
// Tile pipeline (128 threads = 128 consecutive records):
//   A  each thread: ids, src_pos, select0/rank0 of its record → c0, e0
//   B  cooperative, coalesced staging into shared memory: the tile's query bytes (contiguous in the batch), and per record
//      a 208-byte window of target codes at e0 (gap words are read through L1: a record touches 3-4 words per row)
//   C  each thread rewrites its record out of shared memory into a 24-op staging column; detects indel runs
//   D  block scan of op counts → one atomic on the arena cursor → ops and per-record results written (coalesced
//      per-record arrays)
// Records that do not fit the fixed quotas (long reads, > 24 output ops, windows beyond the staged data, malformed index)
// are appended to a deferred list and handled by the warp-per-record kernel afterwards — results are bit-identical.
#include "pv3_device.cuh"
#include "pv3_kernels.h"
#include "pv3_runs.cuh"

namespace pv3 {

constexpr int      TT        = 128;            // threads per CTA = records per tile
constexpr int      T_OUTQ    = 24;             // staged output ops per record
constexpr uint32_t T_QSTAGE  = TT * 168;       // bytes of query staged per tile
constexpr uint32_t T_RWIN    = 208;            // bytes of target codes staged per record (13 x 16)
constexpr uint32_t T_MMW     = 6;              // mismatch bit words per record (query coordinates): reads up to 192 bases

// 64.6 KB ⇒ 3 CTAs per SM. `fin` reuses the memory of `ref` (dead after phase C2; a barrier separates them).
struct TileSmem {
	uint32_t prov[T_OUTQ][TT];                 // provisional ops: '=' means "compare run", decided in C2/C3
	uint32_t mm[T_MMW][TT];
	uint4    q[T_QSTAGE / 16 + 2];
	union {
		uint4    ref[TT][T_RWIN / 16];         // [record][staged target codes]
		uint32_t fin[T_OUTQ][TT];              // final ops
	};
	const uint8_t  *ref_ptr[TT];               // staging sources published by phase A
	uint32_t scan[TT / 32];
	unsigned long long base;
};
static_assert(sizeof(uint4) * (T_RWIN / 16) * TT >= sizeof(uint32_t) * T_OUTQ * TT, "fin must fit in ref");

struct TEmitter {
	uint32_t *out;        // &smem.out[0][tid]
	uint32_t  n_out, carry_op, carry_cnt;
	bool      overflow, adjacent_id;

	__device__ __forceinline__ void init(uint32_t *o) { out = o; n_out = 0; carry_op = OP_NONE; carry_cnt = 0; overflow = false; adjacent_id = false; }
	__device__ __forceinline__ void flush()
	{
		if (carry_op != OP_NONE) {
			if (n_out < (uint32_t) T_OUTQ) out[n_out * TT] = (carry_cnt << 4) | carry_op; else overflow = true;
			++n_out;
		}
	}
	// cigar_buffer::push_back (cigar.cc:61-86)
	__device__ __forceinline__ void push(uint32_t op, uint32_t cnt)
	{
		if (!cnt) return;
		if (carry_op == op) { carry_cnt += cnt; return; }
		if ((carry_op == OP_I && op == OP_D) || (carry_op == OP_D && op == OP_I)) adjacent_id = true;
		flush();
		carry_op = op; carry_cnt = cnt;
	}
};

__device__ __forceinline__ uint32_t ctz64(uint64_t x) { return (uint32_t) __ffsll((long long) x) - 1u; }   // x != 0

// 4 bytes at byte offset a of a word array in shared memory (may read the following word)
__device__ __forceinline__ uint32_t load4(const uint32_t *base, uint32_t a)
{
	const uint32_t *p = base + (a >> 2);
	return __funnelshift_r(p[0], p[1], (a & 3u) * 8u);
}

__device__ __forceinline__ uint32_t load4_global(const uint8_t *p)   // codes have 512 readable pad bytes
{
	return (uint32_t) __ldg(p) | ((uint32_t) __ldg(p + 1) << 8) | ((uint32_t) __ldg(p + 2) << 16) | ((uint32_t) __ldg(p + 3) << 24);
}

// bit i set ⇔ byte i of t is non-zero (SWAR; no per-byte branches)
__device__ __forceinline__ uint32_t nonzero_bytes(uint32_t t)
{
	uint32_t const nz = (((t & 0x7f7f7f7fu) + 0x7f7f7f7fu) | t) & 0x80808080u;
	return (nz * 0x00204081u) >> 28;
}

__global__ void __launch_bounds__(TT)
k1_thread_kernel(K1Params P)
{
	extern __shared__ __align__(16) uint8_t smem_raw[];
	TileSmem &sm = *reinterpret_cast <TileSmem *>(smem_raw);
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
		bool const tile_staged = (tq1 - tq0 + q_lead) <= (uint64_t) T_QSTAGE;

		// ---- phase A: per-record projection of the start position ------------------------------------------------
		uint32_t flags = 0;
		bool defer = !tile_staged;
		RowDesc S = {}, D = {};
		RefDesc R = {};
		uint64_t c0 = 0, e0 = 0, spos = 0, qoff = 0, nq = 0, cbase = 0;
		uint32_t nc = 0;
		sm.ref_ptr[tid] = nullptr;
		if (active) {
			uint32_t const sr = __ldg(P.src_row + rec), dr = __ldg(P.dst_row + rec), rf = __ldg(P.dst_ref + rec);
			if (sr >= P.n_rows || dr >= P.n_rows || rf >= P.n_refs) flags = PV3_FLAG_BAD_ID;
			else {
				S = P.rows[sr]; D = P.rows[dr]; R = P.refs[rf];
				qoff = __ldg(P.seq_off + rec);
				nq = __ldg(P.seq_off + rec + 1) - qoff;
				if (P.qual_off) {
					uint64_t const ql = __ldg(P.qual_off + rec + 1) - __ldg(P.qual_off + rec);
					if (ql && ql != nq) flags |= PV3_FLAG_BAD_QUAL;
				}
				cbase = __ldg(P.cig_off + rec);
				nc = (uint32_t) (__ldg(P.cig_off + rec + 1) - cbase);
				spos = __ldg(P.src_pos + rec);
				c0 = (spos < S.n_zeros) ? select0_thread(S, spos + 1) : S.n_bits;     // rewrite_cigar.hh:51
				e0 = rank0_thread(D, c0);                                              // :52
				if (S.n_bits != D.n_bits || e0 > R.n || nq > 32u * T_MMW) defer = true;                    // unusual index: the general kernel guards it
				if (!defer) {
					sm.ref_ptr[tid] = R.codes + (e0 & ~15ull);
				}
			}
		}
		__syncthreads();

		// ---- phase B: cooperative staging -------------------------------------------------------------------------
		if (tile_staged) {
			uint32_t const n16 = (uint32_t) ((tq1 - tq0 + q_lead + 15) / 16);
			const uint4 *g = reinterpret_cast <const uint4 *>((uintptr_t) q_addr16);
			for (uint32_t i = tid; i < n16; i += TT) sm.q[i] = __ldg(g + i);
		}
		for (uint32_t r = warp * 32 + (lane >> 4); r < warp * 32 + 32; r += 2) {       // half-warp per record
			const uint8_t *rp = sm.ref_ptr[r];
			uint32_t const l16 = lane & 15u;
			if (rp && l16 < T_RWIN / 16) sm.ref[r][l16] = __ldg(reinterpret_cast <const uint4 *>(rp) + l16);
		}
		__syncthreads();

		// ---- phase C1: gap words → provisional ops (compare runs become '=' of unknown content) --------------------
		uint32_t n_prov = 0, n_ops = 0, n_runs = 0;
		unsigned long long job_base = 0;
		bool adjacent_id = false;
		uint64_t const e0a = e0 & ~15ull;
		bool const work = active && !flags && !defer;
		if (work) {
			TEmitter em; em.init(&sm.prov[0][tid]);
			uint64_t col = c0, e = e0, qpos = 0;
			uint64_t src_left = (spos < S.n_zeros) ? (S.n_zeros - spos) : 0;

			for (uint32_t k = 0; k < nc && !flags; ++k) {
				uint32_t const packed = __ldg(P.cig + cbase + k);
				uint32_t const cnt = packed >> 4, op = packed & 0xfu;
				if (op == OP_I || op == OP_S) { em.push(op, cnt); qpos += cnt; continue; }     // rewrite_cigar.hh:68-72
				if (op == OP_H || op == OP_P) { em.push(op, cnt); continue; }                  // :74-77
				if (op > OP_X) { flags |= PV3_FLAG_BAD_OP; break; }                             // :149-151
				bool const mlike = (op == OP_M || op == OP_EQ || op == OP_X);
				if (cnt > src_left) { flags |= PV3_FLAG_POS_OOB; break; }
				src_left -= cnt;
				uint32_t rem = cnt;
				while (rem && !flags) {
					uint64_t const wi = col >> 6;
					uint64_t const Sw = __ldg(S.words + wi), Dw = __ldg(D.words + wi);
					uint64_t valid = ~0ull << (col & 63);
					uint64_t const z = ~Sw & valid;
					uint32_t const zc = __popcll(z);
					uint32_t take;
					if (zc >= rem) {
						unsigned const last = select_in_word(z, rem);
						if (last < 63) valid &= (2ull << last) - 1ull;
						take = rem;
						col = wi * 64 + last + 1;
					} else { take = zc; col = (wi + 1) * 64; }
					uint64_t const sz = ~Sw & valid, dz = ~Dw & valid;
					if (!mlike) {                                        // D, N: rewrite_cigar.hh:119-147
						uint32_t const nd = __popcll(dz);
						em.push(OP_D, nd); e += nd;
					} else {                                             // M, =, X: :79-117
						uint64_t x = sz | dz;
						while (x) {
							uint32_t const p = ctz64(x);
							uint64_t const sb = (sz >> p) & 1ull, db = (dz >> p) & 1ull;
							uint64_t const same = ~((sz ^ (0ull - sb)) | (dz ^ (0ull - db)));
							uint64_t const nr = ~(same >> p);
							uint32_t const len = nr ? ctz64(nr) : 64u;
							if (sb && db) {
								if (qpos + len > nq) { flags |= PV3_FLAG_QUERY_OOB; break; }      // :102
								if (e + len > R.n) { flags |= PV3_FLAG_REF_OOB; break; }          // :103
								em.push(OP_EQ, len);
								qpos += len; e += len;
							} else if (sb) { em.push(OP_I, len); qpos += len; }
							else { em.push(OP_D, len); e += len; }
							x = (p + len >= 64u) ? 0ull : (x & (~0ull << (p + len)));
						}
					}
					rem -= take;
				}
			}
			em.flush();
			n_prov = em.n_out;
			adjacent_id = em.adjacent_id;
			if (em.overflow) defer = true;
		}

		// ---- phase C2: one flat loop over the provisional ops, 4 bases per step: mismatch bits in query coordinates -----
		bool const work2 = work && !flags && !defer;
		if (work2) {
			const uint32_t *qs = reinterpret_cast <const uint32_t *>(sm.q);
			const uint32_t *rs = reinterpret_cast <const uint32_t *>(sm.ref[tid]);
			uint32_t const qrel = q_lead + (uint32_t) (qoff - tq0);
			#pragma unroll
			for (uint32_t w = 0; w < T_MMW; ++w) sm.mm[w][tid] = 0;
			uint32_t j = 0, remaining = 0, qpos = 0, ra = (uint32_t) (e0 - e0a), acc = 0, wcur = 0;
			while (j < n_prov || remaining) {
				if (!remaining) {
					uint32_t const v = sm.prov[j][tid]; ++j;
					uint32_t const op = v & 0xfu, len = v >> 4;
					if (op == OP_EQ) remaining = len;
					else if (op == OP_I || op == OP_S) {
						qpos += len;
						if (qpos > 32u * T_MMW) qpos = 32u * T_MMW;           // I/S beyond |SEQ| (no compare can follow)
						uint32_t const wn = qpos >> 5;
						if (wn != wcur) { if (wcur < T_MMW) sm.mm[wcur][tid] = acc; acc = 0; wcur = wn; }
					} else if (op == OP_D) ra += len;
				} else {
					uint32_t const n = remaining < 4u ? remaining : 4u;
					uint32_t const rv = (ra + 8u <= T_RWIN) ? load4(rs, ra) : load4_global(R.codes + e0a + ra);   // beyond the staged window: L1
					uint32_t bits = nonzero_bytes(load4(qs, qrel + qpos) ^ rv) & ((1u << n) - 1u);
					uint32_t const bit = qpos & 31u;
					acc |= bits << bit;
					qpos += n; ra += n; remaining -= n;
					if (bit + n >= 32u) {
						sm.mm[wcur][tid] = acc;
						acc = (bit + n > 32u) ? (bits >> (32u - bit)) : 0u;
						++wcur;
					}
				}
			}
			if (wcur < T_MMW) sm.mm[wcur][tid] = acc;
		}

		__syncthreads();        // the staged target windows are dead from here on: fin overwrites them

		// ---- phase C3: final ops: provisional '=' runs are split where the mismatch bits say so ---------------------------
		if (work2) {
			TEmitter em; em.init(&sm.fin[0][tid]);
			uint32_t qpos = 0;
			for (uint32_t j = 0; j < n_prov; ++j) {
				uint32_t const v = sm.prov[j][tid];
				uint32_t const op = v & 0xfu, len = v >> 4;
				if (op != OP_EQ) { em.push(op, len); if (op == OP_I || op == OP_S) qpos += len; continue; }
				uint32_t left = len;
				while (left) {
					uint32_t const bit = qpos & 31u;
					uint32_t const n = (32u - bit < left) ? (32u - bit) : left;
					uint32_t m = (sm.mm[qpos >> 5][tid] >> bit) & ((n == 32u) ? ~0u : ((1u << n) - 1u));
					if (!m) em.push(OP_EQ, n);
					else {
						uint32_t pos = 0;
						while (pos < n) {
							uint32_t const rest = m >> pos;
							uint32_t l;
							if (rest & 1u) { uint32_t const inv = ~rest; l = inv ? (uint32_t) (__ffs((int) inv) - 1) : 32u; if (l > n - pos) l = n - pos; em.push(OP_X, l); }
							else { l = rest ? (uint32_t) (__ffs((int) rest) - 1) : 32u; if (l > n - pos) l = n - pos; em.push(OP_EQ, l); }
							pos += l;
						}
					}
					qpos += n; left -= n;
				}
			}
			em.flush();
			n_ops = em.n_out;
			if (em.overflow) defer = true;
			if (!defer && P.stage == 0) {
				if (n_ops == 0) flags |= PV3_FLAG_EMPTY_CIGAR;                                   // cigar.cc:16-18 UB
				else if (adjacent_id) {
					bool has_qual = false;
					if (P.qual_off) has_qual = __ldg(P.qual_off + rec + 1) > __ldg(P.qual_off + rec);
					bool const affine = has_qual || P.gap_open != 0;
					unsigned long long cells = 0, path = 0, pool = 0, max_pool = 0;
					n_runs = scan_runs <false>(&sm.fin[0][tid], TT, n_ops, e0, rec, nq, R.n, nullptr, nullptr, flags, cells, path, pool, max_pool, affine, false);
					if (!flags && n_runs) {
						job_base = atomicAdd(&P.counters->job_cursor, (unsigned long long) n_runs);
						atomicAdd(&P.counters->n_realigned, 1ull);
						atomicAdd(&P.counters->dp_cells, cells);
						atomicAdd(&P.counters->dp_path, path);
						atomicAdd(&P.counters->dp_pool, pool);
						atomicMax(&P.counters->dp_max_pool, max_pool);
						if (job_base + n_runs > P.rr_cap) atomicOr(&P.counters->overflow, 2u);
						else {
							uint32_t f = 0; unsigned long long a = 0, b = 0, c = 0, d = 0;
							scan_runs <true>(&sm.fin[0][tid], TT, n_ops, e0, rec, nq, R.n, P.jobs + job_base, P.rr + 4 * job_base, f, a, b, c, d, affine, true);
						}
					} else n_runs = 0;
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

void launch_k1_thread(const K1Params &P, int n_sms, cudaStream_t stream)
{
	static bool configured = false;
	if (!configured) {
		cudaFuncSetAttribute(k1_thread_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sizeof(TileSmem));
		configured = true;
	}
	uint64_t const n_tiles = (P.n + TT - 1) / TT;
	if (!n_tiles) return;
	uint64_t blocks = (uint64_t) n_sms * 3;       // 3 CTAs of 75 KB shared memory per SM
	if (blocks > n_tiles) blocks = n_tiles;
	k1_thread_kernel<<<(unsigned) blocks, TT, sizeof(TileSmem), stream>>>(P);
}

} // namespace pv3
