// panvc3_b200 — kernels K1S: long records cut into segments, one THREAD per segment (sm_100a).
//
// The records a thread-per-record rewrite kernel cannot take (10-kb reads above all: BASELINE configs[4]) used to go to the
// warp-per-record kernel, which walks 32 MSA columns per step (144 ms per 10^6 x 10-kb reads, 2.5 % of the HBM roofline).
// Here they are cut into segments of <= 64 reference-consuming bases (pv3_segment_core.h: why that is exact) and every
// segment is rewritten by the word-parallel core (pv3_k1w_core.h) like a short record of its own:
//   plan kernel      one thread per deferred record: walks its ops twice (count, then write) → segment descriptors
//   segment kernel   one thread per segment: select0 / rank0 of its start, query staging, span masks, compare, final ops
//                    → segment op arena (one atomic per warp)
//   assembly kernel  (pv3_rewrite.cu) one warp per record: concatenates its segments' ops with cigar_buffer's merge rule,
//                    run detection, per-record outputs
// A record with a segment the core does not take (span beyond its window, more ops than its quota, an error the reference
// aborts on) is left to the warp-per-record kernel, as before. Reference: include/panvc3/rewrite_cigar.hh:19-157,
// libpanvc3/cigar.cc:61-95.
#include "pv3_device.cuh"
#include "pv3_kernels.h"
#include "pv3_segment_core.h"
#include <cstdlib>
#include <cstdio>

namespace pv3 {

namespace {

constexpr int      ST       = 32;      // threads per CTA = segments per tile (one warp)
typedef k1w::CfgSegment <ST> SC;
constexpr uint32_t S_QSLOT  = 27;      // words of staged query per segment (odd: conflict-free lanes): 15 lead + 160 bases + read-ahead
constexpr int      S_LOADS  = 11;      // 16-byte loads that cover 15 + 160 bytes

struct TileS {
	uint32_t cmp[SC::NW][ST], ins[SC::NW][ST], del[SC::NW][ST], xm[SC::NW][ST];
	uint32_t brk[SC::NBRK][ST];
	union {
		uint32_t qs[ST * S_QSLOT];
		uint32_t fin[SC::OUTQ][ST];
	};
};

__device__ __forceinline__ uint4 ld_stream16s(const uint4 *p)
{
	uint4 v;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
	return v;
}

__device__ __forceinline__ uint2 pack16s(uint4 v, uint32_t &bad)
{
	bad |= (v.x | v.y | v.z | v.w) & 0xf8f8f8f8u;
	uint32_t const a = (v.x | (v.x >> 4)) & 0x00ff00ffu, b = (v.y | (v.y >> 4)) & 0x00ff00ffu;
	uint32_t const c = (v.z | (v.z >> 4)) & 0x00ff00ffu, d = (v.w | (v.w >> 4)) & 0x00ff00ffu;
	return make_uint2(__byte_perm(a, b, 0x6420), __byte_perm(c, d, 0x6420));
}

__device__ __forceinline__ uint32_t bam4_word_to_ranks_s(uint32_t w)
{
	uint32_t const x = ((w & 0x0f0f0f0fu) << 4) | ((w >> 4) & 0x0f0f0f0fu);
	uint32_t const cnt = x - ((x >> 1) & 0x77777777u) - ((x >> 2) & 0x33333333u) - ((x >> 3) & 0x11111111u);
	uint32_t const t = cnt ^ 0x11111111u;
	uint32_t const one = (~(((t & 0x77777777u) + 0x77777777u) | t) & 0x88888888u) >> 3;
	uint32_t const m = one * 15u;
	return ((x >> 1) & 0x77777777u & m) | (0x33333333u & ~m);
}

} // namespace

// (A cursor that kept the next 512 columns of both rows in shared memory — eight words per row fetched together instead of one
// dependent load per word — changed nothing: 14.02 against 14.04 ms per 5 x 10^5 reads. The plan kernel is bound by divergent
// issue, 9 of 32 lanes active, not by the loads; the plain seg::RowCursor, which the CPU emulation checks, stays.)
constexpr int PLAN_THREADS = 128;

// One thread per deferred record. Records that cannot be planned go straight to deferred2.
__global__ void __launch_bounds__(PLAN_THREADS)
k1_plan_kernel(K1Params P)
{
	uint64_t const n_items = (uint64_t) *((volatile const unsigned long long *) &P.counters->n_deferred);
	uint64_t const stride = (uint64_t) gridDim.x * blockDim.x;
	seg::SegDesc *segs = (seg::SegDesc *) P.segs;
	for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n_items; i += stride) {
		uint64_t const rec = P.deferred[i];
		uint64_t const cbase = __ldg(P.cig_off + rec), nc64 = __ldg(P.cig_off + rec + 1) - cbase;
		const uint32_t *cig = P.cig + cbase;
		uint32_t ns = seg::UNPLANNABLE, total_ref = 0, rseg = seg::RSEG, qseg = seg::QSEG;
		auto nothing = [](uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t) {};
		// the walk over the two rows needs a record the reference would not abort on (ids, start position, enough source bases)
		uint32_t const sr = __ldg(P.src_row + rec), dr = __ldg(P.dst_row + rec);
		uint64_t const spos = __ldg(P.src_pos + rec);
		seg::RowCursor cur; cur.S = nullptr; cur.D = nullptr; cur.c = 0; cur.e = 0;
		if (nc64 < (1ull << 20) && sr < P.n_rows && dr < P.n_rows) {
			RowDesc const S = P.rows[sr], D = P.rows[dr];
			if (S.n_bits == D.n_bits && spos < S.n_zeros) {
				cur.S = S.words; cur.D = D.words;
				cur.c = select0_thread(S, spos + 1);          // rewrite_cigar.hh:51
				cur.e = rank0_thread(D, cur.c);               // :52
				if (seg::dense_record(cur.S, cur.D, cur.c, nc64, query_length(P, rec))) { rseg = seg::RSEG_DENSE; qseg = seg::QSEG_DENSE; }
				ns = seg::plan_segments(cig, (uint32_t) nc64, nothing, &total_ref, rseg, qseg);
				if (ns != seg::UNPLANNABLE && (uint64_t) total_ref > S.n_zeros - spos) ns = seg::UNPLANNABLE;
			}
		}
		if (ns == seg::UNPLANNABLE) {
			unsigned long long const slot = atomicAdd(&P.counters->n_deferred2, 1ull);
			P.deferred2[slot] = rec;
			continue;
		}
		unsigned long long const base = atomicAdd(&P.counters->seg_cursor, (unsigned long long) ns);
		unsigned long long const slot = atomicAdd(&P.counters->n_planned, 1ull);
		P.plan_rec[slot] = rec; P.plan_seg0[slot] = base; P.plan_nseg[slot] = ns;
		if (base + ns > P.seg_cap) { atomicOr(&P.counters->overflow, 128u); continue; }
		seg::SegDesc *out = segs + base;
		uint32_t w = 0;
		seg::plan_segments(cig, (uint32_t) nc64, [&](uint32_t kind, uint32_t k0, uint32_t k1, uint32_t fc, uint32_t lc, uint32_t rs, uint32_t qs, uint32_t rc) {
			seg::SegDesc d;
			d.rec = rec; d.k0 = k0; d.n_k = (uint8_t) (k1 - k0); d.first_cnt = fc; d.last_cnt = lc; d.ref_start = rs; d.q_start = qs; d.kind = (uint8_t) kind; d.ref_count = (uint16_t) rc;
			d.c_start = cur.c; d.e_start = cur.e;                 // the first base's column, or right behind the previous segment's last base
			if (rc) cur.advance(rc);
			out[w++] = d;
		}, nullptr, rseg, qseg);
	}
}

__global__ void __launch_bounds__(ST)      // (94 registers, 20 tiles per SM; capped at 80 for 24 tiles it spills and is 20 % slower)
k1_segment_kernel(K1Params P)
{
	extern __shared__ __align__(16) uint8_t smem_raw_s[];
	TileS &sm = *reinterpret_cast <TileS *>(smem_raw_s);
	if (*((volatile const unsigned int *) &P.counters->overflow) & 128u) return;       // the segment list is incomplete: the host re-runs
	unsigned const tid = threadIdx.x, lane = tid;
	uint64_t const n_items = (uint64_t) *((volatile const unsigned long long *) &P.counters->seg_cursor);
	uint64_t const n_tiles = (n_items + ST - 1) / ST;
	const seg::SegDesc *segs = (const seg::SegDesc *) P.segs;
	seg::SegOut *seg_out = (seg::SegOut *) P.seg_out;
	k1w::Lane <SC> L;
	L.cmp = &sm.cmp[0][tid]; L.ins = &sm.ins[0][tid]; L.del = &sm.del[0][tid]; L.xm = &sm.xm[0][tid];
	L.brk = &sm.brk[0][tid]; L.fin = &sm.fin[0][tid];
	bool const packed = P.seq_len != nullptr;
	uint64_t const per_cta = (n_tiles + gridDim.x - 1) / gridDim.x;
	uint64_t const tile_begin = (uint64_t) blockIdx.x * per_cta, tile_end = (tile_begin + per_cta < n_tiles) ? tile_begin + per_cta : n_tiles;

	for (uint64_t tile = tile_begin; tile < tile_end; ++tile) {
		uint64_t const s = tile * ST + tid;
		bool const active = s < n_items;
		bool fail = false, core = false;
		uint32_t n_ops = 0, qrel = 0;
		bool adj = false;
		seg::SegDesc d = {};
		k1w::SpanState st = {};
		k1w::OpRange opr = {};
		RefDesc R = {};
		uint64_t e_start = 0;
		if (active) {
			d = segs[s];
			uint64_t const rec = d.rec;
			uint64_t const cbase = __ldg(P.cig_off + rec);
			opr.cig = P.cig + cbase; opr.k0 = d.k0; opr.k1 = d.k1(); opr.first_cnt = d.first_cnt; opr.last_cnt = d.last_cnt;
			if (d.kind == seg::KIND_CORE) {
				opr.first_val = d.n_k != 0 ? __ldg(opr.cig + d.k0) : 0u;
				uint32_t const sr = __ldg(P.src_row + rec), dr = __ldg(P.dst_row + rec), rf = __ldg(P.dst_ref + rec);
				if (sr >= P.n_rows || dr >= P.n_rows || rf >= P.n_refs) fail = true;      // the general kernel flags the record
				else {
					RowDesc const S = P.rows[sr], D = P.rows[dr];
					R = P.refs[rf];
					uint64_t const qoff = __ldg(P.seq_off + rec);
					uint64_t const nq = packed ? (uint64_t) __ldg(P.seq_len + rec) : __ldg(P.seq_off + rec + 1) - qoff;
					uint64_t const spos = __ldg(P.src_pos + rec) + d.ref_start;
					if (d.q_start > nq) fail = true;
					else {
						uint64_t const c_start = d.c_start;                  // found by the planner (no select / rank here)
						e_start = d.e_start;
						if (e_start > R.n) fail = true;
						else {
							uint64_t const q_left = nq - d.q_start;
							uint32_t const nq_seg = (uint32_t) (q_left < seg::QSEG ? q_left : seg::QSEG);
							// query staging: this segment's bases into its slot
							uint64_t const q_byte = packed ? (d.q_start >> 1) : (uint64_t) d.q_start;
							uint64_t const q_addr = (uint64_t) (uintptr_t) P.seq + qoff + q_byte, q_addr16 = q_addr & ~15ull;
							uint32_t const q_lead = (uint32_t) (q_addr - q_addr16);
							uint32_t const nbytes = packed ? (nq_seg + (d.q_start & 1u) + 1u) / 2u : nq_seg;
							uint32_t const n16 = (nbytes + q_lead + 15u) / 16u;               // <= S_LOADS
							const uint4 *g = reinterpret_cast <const uint4 *>((uintptr_t) q_addr16);
							uint32_t *slot = &sm.qs[tid * S_QSLOT];
							uint32_t bad_rank = 0;
							for (uint32_t j0 = 0; j0 < n16; j0 += S_LOADS) {                    // all loads of a segment in flight at once (two rounds of six cost 20 %: latency)
								uint4 v[S_LOADS];
								#pragma unroll
								for (int j = 0; j < S_LOADS; ++j) if (j0 + j < n16) v[j] = ld_stream16s(g + j0 + j);
								#pragma unroll
								for (int j = 0; j < S_LOADS; ++j) {
									uint32_t const i = j0 + j;
									if (i < n16) {
										if (!packed) { uint2 const t = pack16s(v[j], bad_rank); slot[2 * i] = t.x; slot[2 * i + 1] = t.y; }
										else {
											slot[4 * i] = bam4_word_to_ranks_s(v[j].x); slot[4 * i + 1] = bam4_word_to_ranks_s(v[j].y);
											slot[4 * i + 2] = bam4_word_to_ranks_s(v[j].z); slot[4 * i + 3] = bam4_word_to_ranks_s(v[j].w);
										}
									}
								}
							}
							qrel = packed ? ((q_lead << 1) + (d.q_start & 1u)) : q_lead;
							if (bad_rank) fail = true;
							else {
								const uint8_t *rp = R.codes4 + ((e_start & ~31ull) >> 1);
								asm volatile("prefetch.global.L1 [%0];" :: "l"(rp));
								k1w::SpanIn I;
								I.Sp = S.words + (c_start >> 6); I.Dp = D.words + (c_start >> 6); I.c_lo6 = (uint32_t) c_start & 63u;
								I.src_left = S.n_zeros - spos; I.e_start = e_start; I.ref_n = R.n; I.nq = nq_seg; I.q_start = 0;
								if (k1w::build_span <SC, true>(I, opr, L, st)) fail = true;
								else core = true;
							}
						}
					}
				}
			}
		}
		__syncwarp();
		if (core) {
			uint64_t const ea = e_start & ~31ull;
			k1w::compare_span <SC, true>(L, st.n_brk, opr.cig, st.w_end, &sm.qs[tid * S_QSLOT], qrel, 0u,
				reinterpret_cast <const uint32_t *>(R.codes4 + (ea >> 1)), (uint32_t) (e_start - ea));
			k1w::mark_runs <SC, true>(L, st.n_brk, st.w_end);
		}
		__syncthreads();        // the staged query is dead from here on: fin overwrites it
		if (core) n_ops = k1w::emit_ops <SC, true>(L, st.n_brk, opr, st.w_end, adj);
		else if (active && !fail) n_ops = seg::emit_pass_segment <SC>(opr.cig, d.k0, d.k1(), L);      // KIND_PASS
		if (n_ops > (uint32_t) SC::OUTQ) { fail = true; }
		if (fail) n_ops = 0;
		// segment op arena: one atomic per warp
		uint32_t incl = n_ops;
		#pragma unroll
		for (int dd = 1; dd < 32; dd <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, dd); if ((int) lane >= dd) incl += t; }
		uint32_t const total = __shfl_sync(FULL, incl, 31);
		unsigned long long base = 0;
		if (lane == 0 && total) {
			base = atomicAdd(&P.counters->seg_ops_cursor, (unsigned long long) total);
			if (base + total > P.seg_ops_cap) { atomicOr(&P.counters->overflow, 256u); base = ~0ull; }
		}
		base = shfl_u64(base, 0);
		if (active) {
			seg::SegOut o;
			o.off = base == ~0ull ? 0 : base + (incl - n_ops);
			o.n_ops = base == ~0ull ? 0 : n_ops;
			o.flags = (adj ? 1u : 0u) | (fail ? 2u : 0u);
			o.first = n_ops ? sm.fin[0][tid] : 0u; o.last = n_ops ? sm.fin[n_ops - 1][tid] : 0u; o.pad = 0;
			seg_out[s] = o;
			if (base != ~0ull) for (uint32_t j = 0; j < n_ops; ++j) P.seg_ops[o.off + j] = sm.fin[j][tid];
		}
		__syncthreads();
	}
}

void launch_k1_segments(const K1Params &P, int n_sms, uint64_t expected, cudaStream_t stream)
{
	static int per_sm = 0;
	size_t const bytes = sizeof(TileS);
	if (!per_sm) {
		cudaFuncSetAttribute(k1_segment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes);
		if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k1_segment_kernel, ST, bytes) != cudaSuccess || per_sm < 1) per_sm = 1;
		if (getenv("PV3_TRACE")) fprintf(stderr, "[pv3 trace] k1_segment_kernel: %zu bytes of shared memory per tile, %d CTAs per SM\n", bytes, per_sm);
	}
	// both grids are resident and read their item counts on the device (no host round trip after the thread kernel)
	// (a handful of expected records: small grids — filling 148 x 24 CTAs that find nothing to do costs 15 us per launch)
	uint64_t plan_ctas = (expected + 127) / 128 + 4, seg_ctas = expected * 4 + 32;
	if (plan_ctas > (uint64_t) n_sms * 8) plan_ctas = (uint64_t) n_sms * 8;
	if (seg_ctas > (uint64_t) n_sms * per_sm) seg_ctas = (uint64_t) n_sms * per_sm;
	k1_plan_kernel<<<(unsigned) plan_ctas, PLAN_THREADS, 0, stream>>>(P);
	k1_segment_kernel<<<(unsigned) seg_ctas, ST, bytes, stream>>>(P);
}

} // namespace pv3
