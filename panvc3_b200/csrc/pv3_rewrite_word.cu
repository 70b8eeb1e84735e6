// panvc3_b200 — kernel K1W: projection + CIGAR rewrite + run detection, one thread per record, WORD-parallel (sm_100a).
//
// Reference: include/panvc3/rewrite_cigar.hh:19-157, libpanvc3/cigar.cc:61-95, libpanvc3/indel_run_checker.cc:31-121,
// libpanvc3/alignment_projector.cc:61-73. The per-record algorithm is in pv3_k1w_core.h (compiled for the host too and
// checked there against the oracle, tests/test_k1w_core.py); this file is the tile pipeline around it:
//   B1  the tile's query bytes (contiguous in the batch) → 4-bit nibbles in shared memory, coalesced 16-byte loads
//   A   per thread: ids, select0 / rank0 of the start position, op scan, span masks (NW word pairs per row), events
//   C2  compare_span: mismatch bits in column coordinates, <= 16 bases per step
//   C3  emit_ops: final ops into shared memory; run detection (pv3_runs.cuh) where an I abuts a D
//   D   warp scan of op counts → one atomic on the arena cursor per tile → coalesced per-record results
// Records the core does not take (errors, long spans, many events / ops, query ranks above 7, long reads) go to the
// deferred list; the warp-per-record kernel handles them afterwards with the reference's sequential semantics.
#include "pv3_device.cuh"
#include "pv3_kernels.h"
#include "pv3_runs.cuh"
#include "pv3_k1w_core.h"
#include <cstdlib>
#include <cstdio>

namespace pv3 {

namespace {

constexpr int      WT       = 32;                    // threads per CTA = records per tile (one warp)
typedef k1w::CfgRecord <WT> WC;
constexpr uint32_t W_QSTAGE = WT * 156;              // query bases staged per tile

constexpr uint32_t W_QSLOT  = 29;                    // LIST mode: words of staged query per record (odd: conflict-free lanes), 232 nibbles

template <bool LIST> struct TileW {
	uint32_t cmp[WC::NW][WT], ins[WC::NW][WT], del[WC::NW][WT], xm[WC::NW][WT];
	uint32_t brk[WC::NBRK][WT];
	union {
		uint2    q4[W_QSTAGE / 16 + 2];              // !LIST: the tile's packed query, 16 bases per element (dead after C2)
		uint32_t qs[LIST ? WT * W_QSLOT : 1];        //  LIST: one slot per record
		uint32_t fin[WC::OUTQ][WT];                 // final ops
	};
};

// 16-byte streaming load: the query bytes are read once; keeping them out of L1 leaves it to the index words and the
// packed target windows (24 tiles x 4.8 KB of query per SM would otherwise flush the ~30 KB of L1 left beside shared memory)
__device__ __forceinline__ uint4 ld_stream16(const uint4 *p)
{
	uint4 v;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
	return v;
}

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }

// 16 query ranks (one per byte) → 16 nibbles; `bad` collects bits 3..7 of every byte (ranks the nibble compare cannot hold)
__device__ __forceinline__ uint2 pack16w(uint4 v, uint32_t &bad)
{
	bad |= (v.x | v.y | v.z | v.w) & 0xf8f8f8f8u;
	uint32_t const a = (v.x | (v.x >> 4)) & 0x00ff00ffu, b = (v.y | (v.y >> 4)) & 0x00ff00ffu;
	uint32_t const c = (v.z | (v.z >> 4)) & 0x00ff00ffu, d = (v.w | (v.w >> 4)) & 0x00ff00ffu;
	return make_uint2(__byte_perm(a, b, 0x6420), __byte_perm(c, d, 0x6420));
}

// 8 BAM 4-bit codes (first base in the HIGH nibble of each byte) → 8 dna5 ranks, first base in the lowest nibble.
__device__ __forceinline__ uint32_t bam4_word_to_ranks_w(uint32_t w)
{
	uint32_t const x = ((w & 0x0f0f0f0fu) << 4) | ((w >> 4) & 0x0f0f0f0fu);
	uint32_t const cnt = x - ((x >> 1) & 0x77777777u) - ((x >> 2) & 0x33333333u) - ((x >> 3) & 0x11111111u);
	uint32_t const t = cnt ^ 0x11111111u;
	uint32_t const one = (~(((t & 0x77777777u) + 0x77777777u) | t) & 0x88888888u) >> 3;
	uint32_t const m = one * 15u;
	return ((x >> 1) & 0x77777777u & m) | (0x33333333u & ~m);
}

// indel_run_checker over the staged final ops (rare: only records where an I abuts a D): out of line, the hot path stays small
__device__ __noinline__ uint32_t count_runs(const uint32_t *fin, uint32_t n_ops, uint64_t e0, uint64_t rec, uint64_t nq, uint64_t ref_n, uint32_t &flags,
	unsigned long long &cells, unsigned long long &path, unsigned long long &pool, unsigned long long &max_pool, bool affine, uint32_t &n_big)
{
	return scan_runs <false>(fin, WT, n_ops, e0, rec, nq, ref_n, nullptr, nullptr, flags, cells, path, pool, max_pool, affine, false, &n_big);
}

__device__ __noinline__ void write_runs(const uint32_t *fin, uint32_t n_ops, uint64_t e0, uint64_t rec, uint64_t nq, uint64_t ref_n, Job *jobs, uint64_t *rr, bool affine)
{
	uint32_t f = 0; unsigned long long a = 0, b = 0, c = 0, d = 0;
	scan_runs <true>(fin, WT, n_ops, e0, rec, nq, ref_n, jobs, rr, f, a, b, c, d, affine, true);
}

} // namespace

// LIST = false: records 0..n in tiles of 32 consecutive records, records without events only (the others are appended to
// P.complex_list). LIST = true: the records of P.rec_list (length read on the device), events enabled, per-record staging.
template <bool LIST>
__global__ void __launch_bounds__(WT)
k1_word_kernel(K1Params P)
{
	constexpr bool EV = true;        // events are handled in place: a second launch for the records with I S H P D N ops cost more
	                                 // (0.12 ms per 157 k records, latency-bound) than running the event code in every warp
	extern __shared__ __align__(16) uint8_t smem_raw_w[];
	TileW <LIST> &sm = *reinterpret_cast <TileW <LIST> *>(smem_raw_w);
	unsigned const tid = threadIdx.x, lane = tid;
	uint64_t const n_items = LIST ? (uint64_t) *P.n_list_dev : P.n;
	uint64_t const n_tiles = (n_items + WT - 1) / WT;
	k1w::Lane <WC> L;
	L.cmp = &sm.cmp[0][tid]; L.ins = &sm.ins[0][tid]; L.del = &sm.del[0][tid]; L.xm = &sm.xm[0][tid];
	L.brk = &sm.brk[0][tid]; L.fin = &sm.fin[0][tid];
	bool const packed = P.seq_len != nullptr;                               // BAM 4-bit input: offsets count bytes, two bases each

	// Every CTA takes a contiguous range of tiles: the next tile's input follows this tile's in memory, so it can be pulled
	// towards L2 one tile ahead without knowing any offset.
	uint64_t const per_cta = (n_tiles + gridDim.x - 1) / gridDim.x;
	uint64_t const tile_begin = (uint64_t) blockIdx.x * per_cta, tile_end = (tile_begin + per_cta < n_tiles) ? tile_begin + per_cta : n_tiles;
	for (uint64_t tile = tile_begin; tile < tile_end; ++tile) {
		uint64_t const t0 = tile * WT;
		uint32_t const n_here = (uint32_t) ((n_items - t0 < (uint64_t) WT) ? (n_items - t0) : WT);
		bool const active = tid < n_here;
		uint64_t const rec = LIST ? (active ? P.rec_list[t0 + tid] : 0ull) : t0 + tid;

		uint32_t bad_rank = 0, q_lead = 0;
		uint64_t tq0 = 0;
		bool tile_fits = true;
		if (!LIST) {
			// The tile's query bytes are contiguous in the batch: [seq_off[t0], seq_off[t0 + n_here])
			tq0 = __ldg(P.seq_off + t0);
			uint64_t const tq1 = __ldg(P.seq_off + t0 + n_here);
			uint64_t const q_addr = (uint64_t) (uintptr_t) P.seq + tq0;
			uint64_t const q_addr16 = q_addr & ~15ull;
			q_lead = (uint32_t) (q_addr - q_addr16);
			tile_fits = ((tq1 - tq0 + q_lead) << (packed ? 1 : 0)) <= (uint64_t) W_QSTAGE;
			if (tile + 1 < tile_end) {
				// the next tile's streams towards L2 (its query bytes start where this tile's end; about the same length)
				const uint8_t *nq_ = P.seq + tq1;
				uint64_t const nbytes = tq1 - tq0 + 256;
				if ((uint64_t) tid * 128 < nbytes) prefetch_l2(nq_ + (uint64_t) tid * 128);
				if ((uint64_t) (tid + 32) * 128 < nbytes) prefetch_l2(nq_ + (uint64_t) (tid + 32) * 128);
				uint64_t const nt = t0 + WT;
				const void *sp = nullptr;
				switch (tid) {
					case 0: sp = P.src_row + nt; break;       case 1: sp = P.dst_row + nt; break;       case 2: sp = P.dst_ref + nt; break;
					case 3: sp = P.src_pos + nt; break;       case 4: sp = P.src_pos + nt + 16; break;  case 5: sp = P.seq_off + nt; break;
					case 6: sp = P.seq_off + nt + 16; break;  case 7: sp = P.cig_off + nt; break;       case 8: sp = P.cig_off + nt + 16; break;
					case 9: sp = P.qual_off ? P.qual_off + nt : nullptr; break;
					case 10: sp = P.qual_off ? P.qual_off + nt + 16 : nullptr; break;
					case 11: sp = P.seq_len ? P.seq_len + nt : nullptr; break;
					default: break;
				}
				if (sp) prefetch_l2(sp);
			}
			// ---- phase B1: the tile's query, packed to nibbles while it is staged. Five 16-byte loads per lane are in flight
			// at a time (the first version waited for every load: long_scoreboard on the pack was the top stall) ------------------
			if (tile_fits) {
				uint32_t const n16 = (uint32_t) ((tq1 - tq0 + q_lead + 15) / 16);
				const uint4 *g = reinterpret_cast <const uint4 *>((uintptr_t) q_addr16);
				for (uint32_t i0 = 0; i0 < n16; i0 += 5 * WT) {
					uint4 v[5];
					#pragma unroll
					for (int j = 0; j < 5; ++j) { uint32_t const i = i0 + j * WT + tid; if (i < n16) v[j] = ld_stream16(g + i); }
					#pragma unroll
					for (int j = 0; j < 5; ++j) {
						uint32_t const i = i0 + j * WT + tid;
						if (i < n16) {
							if (!packed) sm.q4[i] = pack16w(v[j], bad_rank);
							else {
								sm.q4[2 * i] = make_uint2(bam4_word_to_ranks_w(v[j].x), bam4_word_to_ranks_w(v[j].y));
								sm.q4[2 * i + 1] = make_uint2(bam4_word_to_ranks_w(v[j].z), bam4_word_to_ranks_w(v[j].w));
							}
						}
					}
				}
			}
		}

		// ---- phase A: start position, op scan, span masks -----------------------------------------------------------
		uint32_t flags = 0;
		bool defer = !tile_fits, complex_rec = false;
		RefDesc R = {};
		uint64_t e0 = 0, qoff = 0, nq = 0, cbase = 0, c0_keep = 0;
		uint32_t nc = 0;
		k1w::SpanState st = {};
		k1w::OpRange opr = {};
		if (active) {
			uint32_t const sr = __ldg(P.src_row + rec), dr = __ldg(P.dst_row + rec), rf = __ldg(P.dst_ref + rec);
			if (sr >= P.n_rows || dr >= P.n_rows || rf >= P.n_refs) flags = PV3_FLAG_BAD_ID;
			else {
				RowDesc const S = P.rows[sr], D = P.rows[dr];
				R = P.refs[rf];
				qoff = __ldg(P.seq_off + rec);
				nq = packed ? (uint64_t) __ldg(P.seq_len + rec) : __ldg(P.seq_off + rec + 1) - qoff;
				if (P.qual_off) {
					uint64_t const ql = __ldg(P.qual_off + rec + 1) - __ldg(P.qual_off + rec);
					if (ql && ql != nq) flags |= PV3_FLAG_BAD_QUAL;
				}
				cbase = __ldg(P.cig_off + rec);
				uint64_t const nc64 = __ldg(P.cig_off + rec + 1) - cbase;
				nc = (uint32_t) nc64;
				uint32_t const first_op = nc ? __ldg(P.cig + cbase) : 0u;                            // in flight while select0 / rank0 run
				uint64_t const spos = __ldg(P.src_pos + rec);
				uint64_t c0;
				if (LIST) { c0 = P.complex_ce[2 * (t0 + tid)]; e0 = P.complex_ce[2 * (t0 + tid) + 1]; }   // found by the first launch
				else {
					c0 = (spos < S.n_zeros) ? select0_thread(S, spos + 1) : S.n_bits;             // rewrite_cigar.hh:51
					e0 = rank0_thread(D, c0);                                                      // :52
				}
				c0_keep = c0;
				if (S.n_bits != D.n_bits || e0 > R.n || nc64 >= 0x8000ull || spos >= S.n_zeros || nq > WC::QMAX) defer = true;   // unusual: the general kernel guards it
				if (!defer && !flags) {
					// the window of packed target codes C2 will read: towards L1 while the masks are built
					const uint8_t *rp = R.codes4 + ((e0 & ~31ull) >> 1);
					asm volatile("prefetch.global.L1 [%0];" :: "l"(rp));
					asm volatile("prefetch.global.L1 [%0];" :: "l"(rp + 96));
					k1w::SpanIn I;
					I.Sp = S.words + (c0 >> 6); I.Dp = D.words + (c0 >> 6); I.c_lo6 = (uint32_t) c0 & 63u;
					I.src_left = S.n_zeros - spos; I.e_start = e0; I.ref_n = R.n;
					I.nq = (uint32_t) nq; I.q_start = 0;
					opr.cig = P.cig + cbase; opr.k0 = 0; opr.k1 = nc; opr.first_cnt = ~0u; opr.last_cnt = ~0u; opr.first_val = first_op;
					int const rc = k1w::build_span <WC, EV>(I, opr, L, st);
					defer = rc == 1;
					complex_rec = rc == 2;
				}
			}
		}
		if (LIST) {
			// ---- phase B1 (LIST): every lane stages its own record's query into its slot ----------------------------------
			if (active && !defer && !flags) {
				uint64_t const q_addr = (uint64_t) (uintptr_t) P.seq + qoff, q_addr16 = q_addr & ~15ull;
				q_lead = (uint32_t) (q_addr - q_addr16);
				uint64_t const nbytes = packed ? (nq + 1) / 2 : nq;
				uint32_t const n16 = (uint32_t) ((nbytes + q_lead + 15) / 16);              // <= 13 (7 when packed)
				const uint4 *g = reinterpret_cast <const uint4 *>((uintptr_t) q_addr16);
				uint32_t *slot = &sm.qs[tid * W_QSLOT];
				for (uint32_t i0 = 0; i0 < n16; i0 += 7) {                                   // two rounds of seven loads in flight
					uint4 v[7];
					#pragma unroll
					for (int j = 0; j < 7; ++j) if (i0 + j < n16) v[j] = ld_stream16(g + i0 + j);
					#pragma unroll
					for (int j = 0; j < 7; ++j) {
						uint32_t const i = i0 + j;
						if (i < n16) {
							if (!packed) { uint2 const t = pack16w(v[j], bad_rank); slot[2 * i] = t.x; slot[2 * i + 1] = t.y; }
							else {
								slot[4 * i] = bam4_word_to_ranks_w(v[j].x); slot[4 * i + 1] = bam4_word_to_ranks_w(v[j].y);
								slot[4 * i + 2] = bam4_word_to_ranks_w(v[j].z); slot[4 * i + 3] = bam4_word_to_ranks_w(v[j].w);
							}
						}
					}
				}
				if (bad_rank) defer = true;                  // a query rank above 7: the nibble compare cannot represent it
			}
			__syncwarp();
		} else {
			// a query rank above 7 anywhere in the tile: the nibble compare cannot represent it, the general kernel takes the tile
			if (__syncthreads_or((int) bad_rank)) defer = true;      // also: the staged query is complete
		}

		// ---- phase C2: mismatch bits ------------------------------------------------------------------------------------
		bool const work = active && !flags && !defer && !complex_rec;
		uint64_t const e0a = e0 & ~31ull;
		if (work) {
			const uint32_t *qs = LIST ? &sm.qs[tid * W_QSLOT] : reinterpret_cast <const uint32_t *>(sm.q4);
			uint32_t const qrel = LIST ? (q_lead << (packed ? 1 : 0)) : ((q_lead + (uint32_t) (qoff - tq0)) << (packed ? 1 : 0));
			k1w::compare_span <WC, EV>(L, st.n_brk, opr.cig, st.w_end, qs, qrel, 0u,
				reinterpret_cast <const uint32_t *>(R.codes4 + (e0a >> 1)), (uint32_t) (e0 - e0a));
			k1w::mark_runs <WC, EV>(L, st.n_brk, st.w_end);
		}
		__syncthreads();        // the staged query is dead from here on: fin overwrites it

		// ---- phase C3: final ops, run detection -------------------------------------------------------------------------
		uint32_t n_ops = 0, n_runs = 0, n_big = 0;
		unsigned long long cells = 0, path = 0, pool = 0, max_pool = 0, job_base = 0;
		bool affine = false;
		if (work) {
			bool adjacent_id = false;
			n_ops = k1w::emit_ops <WC, EV>(L, st.n_brk, opr, st.w_end, adjacent_id);
			if (n_ops > (uint32_t) WC::OUTQ) defer = true;
			if (!defer && P.stage == 0) {
				if (n_ops == 0) flags |= PV3_FLAG_EMPTY_CIGAR;                               // cigar.cc:16-18 UB
				else if (adjacent_id) {
					bool has_qual = false;
					if (P.qual_off) has_qual = __ldg(P.qual_off + rec + 1) > __ldg(P.qual_off + rec);
					affine = has_qual || P.gap_open != 0;
					n_runs = count_runs(&sm.fin[0][tid], n_ops, e0, rec, nq, R.n, flags, cells, path, pool, max_pool, affine, n_big);
					if (flags) n_runs = 0;
				}
			}
		}
		// Runs found: one set of atomics per WARP, then each record writes its jobs and reported ranges behind the warp's base.
		if (__any_sync(FULL, n_runs != 0)) {
			uint32_t incl = n_runs;
			unsigned long long w_cells = n_runs ? cells : 0, w_path = n_runs ? path : 0, w_pool = n_runs ? pool : 0, w_max = n_runs ? max_pool : 0;
			uint32_t const w_recs = __popc(__ballot_sync(FULL, n_runs != 0));
			uint32_t w_bigjobs = n_runs ? n_big : 0u;
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
				if (wbase + w_runs <= P.rr_cap) write_runs(&sm.fin[0][tid], n_ops, e0, rec, nq, R.n, P.jobs + job_base, P.rr + 4 * job_base, affine);
			}
		}
		// records for the second launch / for the general kernel: one atomic per warp and list
		{
			bool const dfr = active && defer && !flags;
			uint32_t const m = __ballot_sync(FULL, dfr);
			if (m) {
				unsigned long long slot = 0;
				if (lane == 0) slot = atomicAdd(&P.counters->n_deferred, (unsigned long long) __popc(m));
				slot = shfl_u64(slot, 0);
				if (dfr) P.deferred[slot + __popc(m & lanemask_lt())] = rec;
			}
			if (!LIST) {
				bool const cpx = active && complex_rec && !defer && !flags;
				uint32_t const mc = __ballot_sync(FULL, cpx);
				if (mc) {
					unsigned long long slot = 0;
					if (lane == 0) slot = atomicAdd(&P.counters->n_complex, (unsigned long long) __popc(mc));
					slot = shfl_u64(slot, 0);
					if (cpx) {
						unsigned long long const at = slot + __popc(mc & lanemask_lt());
						P.complex_list[at] = rec;
						P.complex_ce[2 * at] = c0_keep; P.complex_ce[2 * at + 1] = e0;
					}
				}
			}
		}

		// ---- phase D: warp scan, one arena allocation per tile, write-out ---------------------------------------------------
		bool const done = active && !defer && !complex_rec;                      // this launch owns the record's outputs (flagged ones included)
		uint32_t const cnt = (done && !flags) ? n_ops : 0u;
		uint32_t incl = cnt;
		#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
		uint32_t const total = __shfl_sync(FULL, incl, 31);
		unsigned long long base = 0;
		if (lane == 0 && total) {
			base = atomicAdd(&P.counters->arena_cursor, (unsigned long long) total);
			if (base + total > P.arena_cap) { atomicOr(&P.counters->overflow, 1u); base = ~0ull; }
		}
		base = shfl_u64(base, 0);
		uint32_t const my_off = incl - cnt;
		uint32_t const w_flagged = __popc(__ballot_sync(FULL, active && flags != 0));
		if (lane == 0 && w_flagged) atomicAdd(&P.counters->n_flagged, (unsigned long long) w_flagged);
		if (done || (active && flags)) {
			if (flags) n_runs = 0;
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

template <bool LIST>
static void launch_word_variant(const K1Params &P, int n_sms, cudaStream_t stream)
{
	static int per_sm = 0;
	size_t const bytes = sizeof(TileW <LIST>);
	if (!per_sm) {
		cudaFuncSetAttribute(k1_word_kernel <LIST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes);
		if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_word_kernel <LIST>, WT, bytes) != cudaSuccess || per_sm < 1) per_sm = 1;
		if (getenv("PV3_TRACE")) fprintf(stderr, "[pv3 trace] k1_word_kernel<%d>: %zu bytes of shared memory per tile, %d CTAs per SM\n", (int) LIST, bytes, per_sm);
	}
	uint64_t blocks = (uint64_t) n_sms * per_sm;     // persistent: every resident CTA strides over the tiles
	if (!LIST) {
		uint64_t const n_tiles = (P.n + WT - 1) / WT;
		if (!n_tiles) return;
		if (blocks > n_tiles) blocks = n_tiles;
	} else {
		uint64_t const max_tiles = (P.n + WT - 1) / WT;    // the list cannot be longer than the batch
		if (!max_tiles) return;
		if (blocks > max_tiles) blocks = max_tiles;
	}
	k1_word_kernel <LIST><<<(unsigned) blocks, WT, bytes, stream>>>(P);
}

// First launch: all records, those without events are finished; second launch: the records with I S H P D N ops, collected
// by the first (the list length is read on the device: no host round trip).
void launch_k1_word(const K1Params &P, int n_sms, cudaStream_t stream)
{
	launch_word_variant <false>(P, n_sms, stream);
}

} // namespace pv3
