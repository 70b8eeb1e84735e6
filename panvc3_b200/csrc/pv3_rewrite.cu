// panvc3_b200 — kernel K1: start-position projection + CIGAR rewrite + indel-run detection (sm_100a).
//
// Restates, as a warp-cooperative column scan, what the reference does one base at a time:
//   rewrite_cigar()            include/panvc3/rewrite_cigar.hh:19-157   (3 rrr queries per reference-consuming base)
//   cigar_buffer::push_back    libpanvc3/cigar.cc:61-95                 (run-length merge on push)
//   indel_run_checker          libpanvc3/indel_run_checker.cc:31-121
//   range bookkeeping          libpanvc3/alignment_projector.cc:61-73   (incl. the shifted reported location)
//
// Formulation. Let S and D be the source and target gap rows. The k-th reference-consuming base of the read sits at the
// column of the k-th zero of S counted from c0 = select0_S(src_pos + 1).  For that base the reference computes
// e = rank0_D(column), emits D x (e - prev_excess) and then one op; prev_excess afterwards is always
// e + (D[column] == 0), whatever the op kind, so "leading D" = number of target non-gaps in the columns strictly between
// two consecutive source non-gaps.  Hence every MSA column in [c0, c_last] yields at most one unit op:
//      S=0,D=0 : M-like -> '=' / 'X' (raw byte compare)     D-like -> 'D'
//      S=0,D=1 : M-like -> 'I'                                         D-like -> nothing
//      S=1,D=0 : 'D' (leading deletion of the next base)               S=1,D=1 : nothing
// One lane handles one column of a 32-column chunk; popcounts of the chunk's S/D bits below the lane give the query and
// target offsets; run-length merging is done with ballots.  Pass-through ops (I S H P) are pushed between column
// segments exactly where the reference pushes them.  No per-base rank/select is ever issued.
#include "pv3_device.cuh"
#include "pv3_kernels.h"
#include "pv3_runs.cuh"
#include "pv3_segment_core.h"

namespace pv3 {

// ---------------------------------------------------------------------------------------------- run-length emitter
// Warp-uniform state; out may point to shared (staging) or global (direct) memory.
struct Emitter {
	uint32_t *out;
	uint32_t  cap;
	uint32_t  n_out;
	uint32_t  carry_op;    // OP_NONE when empty
	uint32_t  carry_cnt;
	bool      overflow;
	bool      adjacent_id; // an I op directly followed by a D op or vice versa ⇒ an indel run with both kinds exists

	__device__ __forceinline__ void init(uint32_t *o, uint32_t c)
	{
		out = o; cap = c; n_out = 0; carry_op = OP_NONE; carry_cnt = 0; overflow = false; adjacent_id = false;
	}

	__device__ __forceinline__ void write(uint32_t idx, uint32_t op, uint32_t cnt)
	{
		if (idx < cap) out[idx] = (cnt << 4) | op;
	}

	__device__ __forceinline__ void note_boundary(uint32_t prev_op, uint32_t op)
	{
		if ((prev_op == OP_I && op == OP_D) || (prev_op == OP_D && op == OP_I)) adjacent_id = true;
	}

	__device__ __forceinline__ void flush_carry()
	{
		if (carry_op != OP_NONE) {
			if (n_out >= cap) overflow = true;
			if (lane_id() == 0) write(n_out, carry_op, carry_cnt);
			++n_out;
		}
	}

	// cigar_buffer::push_back for a warp-uniform (op, count)
	__device__ __forceinline__ void push_uniform(uint32_t op, uint32_t cnt)
	{
		if (!cnt) return;
		if (carry_op == op) { carry_cnt += cnt; return; }
		note_boundary(carry_op, op);
		flush_carry();
		carry_op = op; carry_cnt = cnt;
	}

	// One unit op per lane (OP_NONE = nothing), pushed in lane order.
	__device__ __forceinline__ void push_lanes(uint32_t code)
	{
		unsigned const lane = lane_id(), lt = lanemask_lt();
		unsigned const nn = __ballot_sync(FULL, code != OP_NONE);
		if (!nn) return;
		unsigned const below = nn & lt;
		int const prev_lane = 31 - __clz(below);                       // -1 when none
		uint32_t prev_code = __shfl_sync(FULL, code, prev_lane & 31);
		if (!below) prev_code = carry_op;
		bool const is_head = (code != OP_NONE) && (code != prev_code);
		unsigned const heads = __ballot_sync(FULL, is_head);
		if (!heads) { carry_cnt += __popc(nn); return; }
		if (__any_sync(FULL, is_head && ((prev_code == OP_I && code == OP_D) || (prev_code == OP_D && code == OP_I))))
			adjacent_id = true;
		int const first_head = __ffs(heads) - 1;
		carry_cnt += __popc(nn & ((1u << first_head) - 1u));
		flush_carry();
		// run length of the run starting at this lane
		unsigned const above = heads & ~lt & ~(1u << lane);            // heads strictly above this lane
		unsigned const upto = above ? ((1u << (__ffs(above) - 1)) - 1u) : FULL;
		uint32_t const my_cnt = __popc(nn & upto & ~lt);
		int const last_head = 31 - __clz(heads);
		uint32_t const n_heads = __popc(heads);
		if (is_head && (int) lane != last_head) {
			uint32_t const idx = n_out + __popc(heads & lt);
			write(idx, code, my_cnt);
		}
		if (n_out + n_heads - 1 > cap) overflow = true;
		n_out += n_heads - 1;
		carry_op = __shfl_sync(FULL, code, last_head);
		carry_cnt = __shfl_sync(FULL, my_cnt, last_head);
	}

	__device__ __forceinline__ void finish() { flush_carry(); carry_op = OP_NONE; carry_cnt = 0; }
};

__device__ __forceinline__ char dna5_to_char(uint8_t rank)
{
	// seqan3 dna5 rank_to_char: A C G N T
	return rank < 5 ? (char) ((0x544e474341ull >> (8 * rank)) & 0xff) : 'N';
}

struct RecordResult {
	uint64_t dst_pos;
	uint32_t n_ops;
	uint32_t flags;
	bool     overflow;
	bool     adjacent_id;
};

// rewrite_cigar() for one record by one warp. All lanes return the same result.
__device__ RecordResult rewrite_record(const K1Params &P, uint64_t rec, uint32_t *out, uint32_t cap)
{
	unsigned const lane = lane_id(), lt = lanemask_lt();
	RecordResult res; res.dst_pos = 0; res.n_ops = 0; res.flags = 0; res.overflow = false; res.adjacent_id = false;

	uint32_t const sr = __ldg(P.src_row + rec), dr = __ldg(P.dst_row + rec), rf = __ldg(P.dst_ref + rec);
	if (sr >= P.n_rows || dr >= P.n_rows || rf >= P.n_refs) { res.flags = PV3_FLAG_BAD_ID; return res; }
	RowDesc const S = P.rows[sr];
	RowDesc const D = P.rows[dr];
	RefDesc const R = P.refs[rf];
	const uint8_t *qptr = P.seq + __ldg(P.seq_off + rec);
	bool const seq_packed = P.seq_len != nullptr;                      // SEQ in BAM's 4-bit packing
	uint64_t const nq = query_length(P, rec);
	if (P.qual_off) {
		uint64_t const ql = __ldg(P.qual_off + rec + 1) - __ldg(P.qual_off + rec);
		if (ql && ql != nq) res.flags |= PV3_FLAG_BAD_QUAL;
	}
	uint64_t const cbase = __ldg(P.cig_off + rec);
	uint32_t const nc = (uint32_t) (__ldg(P.cig_off + rec + 1) - cbase);
	uint64_t const spos = __ldg(P.src_pos + rec);

	// rewrite_cigar.hh:51-56
	uint64_t col = (spos < S.n_zeros) ? select0_warp(S, spos + 1) : S.n_bits;
	uint64_t ecur = rank0_warp(D, col);           // zeros of D in [0, col)
	res.dst_pos = ecur;
	uint64_t src_left = (spos < S.n_zeros) ? (S.n_zeros - spos) : 0;   // source non-gaps at or after col
	uint64_t qpos = 0;
	uint32_t lane_err = 0;

	Emitter em; em.init(out, cap);

	uint64_t wbase = 1ull << 62; // first word of the cached 32-word window (initially far away)
	uint64_t sw = 0, dw = 0;   // lane l caches words wbase + l of both rows

	for (uint32_t k = 0; k < nc; ++k) {
		uint32_t const packed = __ldg(P.cig + cbase + k);
		uint32_t const cnt = packed >> 4, op = packed & 0xfu;
		if (op == OP_I || op == OP_S) { em.push_uniform(op, cnt); qpos += cnt; continue; }   // :68-72
		if (op == OP_H || op == OP_P) { em.push_uniform(op, cnt); continue; }                // :74-77
		if (op > OP_X) { res.flags |= PV3_FLAG_BAD_OP; break; }                               // :149-151
		bool const mlike = (op == OP_M || op == OP_EQ || op == OP_X);                         // :79-117 vs :119-147
		if (cnt > src_left) { res.flags |= PV3_FLAG_POS_OOB; break; }
		src_left -= cnt;
		uint32_t rem = cnt;
		while (rem) {
			uint64_t const wi = col >> 6;
			if (wi - wbase >= 32ull) {           // unsigned wrap: also true for the initial wbase
				wbase = wi;
				sw = __ldg(S.words + wbase + lane);
				dw = (wbase + lane < D.n_words_alloc) ? __ldg(D.words + wbase + lane) : ~0ull;
			}
			int const wsel = (int) (wi - wbase);
			uint64_t const sword = shfl_u64(sw, wsel), dword = shfl_u64(dw, wsel);
			bool const hi_half = (col >> 5) & 1;
			uint32_t const s32 = hi_half ? (uint32_t) (sword >> 32) : (uint32_t) sword;
			uint32_t const d32 = hi_half ? (uint32_t) (dword >> 32) : (uint32_t) dword;
			uint32_t valid = FULL << (col & 31);
			uint32_t z = ~s32 & valid;             // source non-gap columns: one reference-consuming base each
			uint32_t const zc = __popc(z);
			uint32_t take;
			uint64_t const chunk_base = col & ~31ull;
			if (zc >= rem) {
				unsigned const last = select_in_u32(z, rem);
				valid &= (last == 31) ? FULL : ((2u << last) - 1u);
				z &= valid;
				take = rem;
				col = chunk_base + last + 1;
			} else {
				take = zc;
				col = chunk_base + 32;
			}
			uint32_t const dz = ~d32 & valid;      // target non-gap columns
			if (!z) {
				em.push_uniform(OP_D, __popc(dz)); // only leading deletions in this chunk
			} else {
				uint32_t const bit = 1u << lane;
				uint32_t code = OP_NONE;
				if (valid & bit) {
					if (z & bit) {
						if (dz & bit) {
							if (mlike) {
								uint64_t const qi = qpos + __popc(z & lt);
								uint64_t const ei = ecur + __popc(dz & lt);
								if (qi >= nq) lane_err |= PV3_FLAG_QUERY_OOB;          // :102
								else if (ei >= R.n) lane_err |= PV3_FLAG_REF_OOB;       // :103
								else code = (dna5_to_char(seq_packed ? (uint8_t) bam4_to_rank((qi & 1u) ? __ldg(qptr + (qi >> 1)) : (uint32_t) (__ldg(qptr + (qi >> 1)) >> 4)) : __ldg(qptr + qi)) == __ldg(R.bases + ei)) ? OP_EQ : OP_X;
							} else code = OP_D;
						} else if (mlike) code = OP_I;
					} else if (dz & bit) code = OP_D;
				}
				em.push_lanes(code);
			}
			if (mlike) qpos += take;
			ecur += __popc(dz);
			rem -= take;
		}
	}
	em.finish();
	res.flags |= __reduce_or_sync(FULL, lane_err);
	res.n_ops = em.n_out;
	res.overflow = em.overflow;
	res.adjacent_id = em.adjacent_id;
	return res;
}

} // namespace pv3

namespace pv3 {

// Records are handled in groups of GROUP consecutive records per warp: the rewritten CIGARs of a group are staged in
// shared memory and leave with one arena allocation (one atomic per group) and one coalesced copy.
constexpr int GROUP       = 8;
constexpr int STAGE_OPS   = 256;    // per warp, uint32 each
constexpr int K1_WARPS    = 8;

struct GroupState {
	uint32_t  stage_n;     // ops staged so far
	uint32_t  my_off;      // lane r: offset of the group's r-th record in the stage
	uint32_t  my_len;
	uint64_t  my_rec;      // lane r: record index, ~0 when not staged
};

__device__ __forceinline__ uint64_t arena_alloc(const K1Params &P, uint32_t n_words)
{
	unsigned long long base = 0;
	if (lane_id() == 0) base = atomicAdd(&P.counters->arena_cursor, (unsigned long long) n_words);
	base = shfl_u64(base, 0);
	if (base + n_words > P.arena_cap) {
		if (lane_id() == 0) atomicOr(&P.counters->overflow, 1u);
		return ~0ull;
	}
	return base;
}

__device__ void flush_group(const K1Params &P, GroupState &g, const uint32_t *stage)
{
	unsigned const lane = lane_id();
	if (g.stage_n) {
		uint64_t const base = arena_alloc(P, g.stage_n);
		if (g.my_rec != ~0ull) {
			P.cig_out_off[g.my_rec] = (base == ~0ull) ? 0 : base + g.my_off + P.arena_off;
			P.cig_out_len[g.my_rec] = (base == ~0ull) ? 0 : g.my_len;
		}
		if (base != ~0ull) {
			__syncwarp();
			for (uint32_t i = lane; i < g.stage_n; i += 32) P.arena[base + i] = stage[i];
		}
	} else if (g.my_rec != ~0ull) {
		P.cig_out_off[g.my_rec] = 0;
		P.cig_out_len[g.my_rec] = 0;
	}
	__syncwarp();
	g.stage_n = 0; g.my_off = 0; g.my_len = 0; g.my_rec = ~0ull;
}

// Post-rewrite bookkeeping of one record whose ops are at `ops` (shared or global): flags, run detection, per-record outputs.
__device__ void finish_record(const K1Params &P, uint64_t rec, RecordResult &res, const uint32_t *ops)
{
	unsigned const lane = lane_id();
	uint32_t n_runs = 0;
	unsigned long long job_base = 0;
	if (!res.flags && P.stage == 0) {
		if (res.n_ops == 0) res.flags |= PV3_FLAG_EMPTY_CIGAR;            // cigar.cc:16-18 UB
		else if (res.adjacent_id) {
			__syncwarp();
			uint64_t const nq = query_length(P, rec);
			uint64_t const ref_n = P.refs[__ldg(P.dst_ref + rec)].n;
			bool has_qual = false;
			if (P.qual_off) has_qual = __ldg(P.qual_off + rec + 1) > __ldg(P.qual_off + rec);
			bool const affine = has_qual || P.gap_open != 0;
			unsigned long long cells = 0, path = 0, pool = 0, max_pool = 0;
			uint32_t n_big = 0;
			n_runs = scan_runs_warp<false>(ops, res.n_ops, res.dst_pos, rec, nq, ref_n, nullptr, nullptr, res.flags, cells, path, pool, max_pool, affine, &n_big);
			if (!res.flags && n_runs) {
				if (lane == 0) {
					job_base = atomicAdd(&P.counters->job_cursor, (unsigned long long) n_runs);
					atomicAdd(&P.counters->n_realigned, 1ull);
					if (n_big) atomicAdd(&P.counters->n_big_jobs, (unsigned long long) n_big);
					atomicAdd(&P.counters->dp_cells, cells);
					atomicAdd(&P.counters->dp_path, path);
					atomicAdd(&P.counters->dp_pool, pool);
					atomicMax(&P.counters->dp_max_pool, max_pool);
				}
				job_base = shfl_u64(job_base, 0);
				if (job_base + n_runs > P.rr_cap) {
					if (lane == 0) atomicOr(&P.counters->overflow, 2u);
				} else {
					uint32_t f = 0; unsigned long long a = 0, b = 0, c = 0, d = 0;
					scan_runs_warp<true>(ops, res.n_ops, res.dst_pos, rec, nq, ref_n, P.jobs + job_base, P.rr + 4 * job_base, f, a, b, c, d, affine);
				}
			} else n_runs = 0;
		}
	}
	if (res.flags) { n_runs = 0; res.n_ops = 0; }
	if (lane == 0) {
		P.dst_pos[rec] = res.dst_pos;
		P.flags[rec] = res.flags;
		P.rr_off[rec] = job_base + P.rr_base;
		P.rr_len[rec] = n_runs;
		if (res.flags) atomicAdd(&P.counters->n_flagged, 1ull);
	}
}

__global__ void __launch_bounds__(K1_WARPS * 32)
k1_rewrite_kernel(K1Params P, int group)
{
	__shared__ uint32_t stage_all[K1_WARPS][STAGE_OPS];
	unsigned const lane = lane_id();
	unsigned const warp_in_block = threadIdx.x >> 5;
	uint32_t *stage = stage_all[warp_in_block];
	uint64_t const n_items = P.n_list_dev ? (uint64_t) *P.n_list_dev : (P.rec_list ? P.n_list : P.n);
	uint64_t const n_groups = (n_items + group - 1) / group;
	uint64_t const warp_global = (uint64_t) blockIdx.x * K1_WARPS + warp_in_block;
	uint64_t const n_warps = (uint64_t) gridDim.x * K1_WARPS;

	GroupState g; g.stage_n = 0; g.my_off = 0; g.my_len = 0; g.my_rec = ~0ull;

	for (uint64_t grp = warp_global; grp < n_groups; grp += n_warps) {
		uint64_t const rec0 = grp * group;
		int const n_in_group = (int) ((n_items - rec0 < (uint64_t) group) ? (n_items - rec0) : group);
		for (int r = 0; r < n_in_group; ++r) {
			uint64_t const rec = P.rec_list ? P.rec_list[rec0 + r] : rec0 + r;
			RecordResult res = rewrite_record(P, rec, stage + g.stage_n, STAGE_OPS - g.stage_n);
			if (res.overflow && g.stage_n) {
				// does not fit behind the records already staged: send those off and redo this one on an empty stage
				flush_group(P, g, stage);
				res = rewrite_record(P, rec, stage, STAGE_OPS);
			}
			if (res.overflow) {
				// larger than the whole stage: the first pass counted the ops; allocate and write straight to the arena
				uint32_t const n_ops = res.n_ops;
				uint64_t const base = res.flags ? ~0ull : arena_alloc(P, n_ops);
				if (base != ~0ull) {
					res = rewrite_record(P, rec, P.arena + base, n_ops);
					__syncwarp();
					finish_record(P, rec, res, P.arena + base);
				} else {
					res.n_ops = 0;
					if (!res.flags) res.flags = 0; // arena overflow: the host re-runs the batch with a larger arena
					finish_record(P, rec, res, stage);
				}
				if (lane == 0) {
					P.cig_out_off[rec] = (base == ~0ull) ? 0 : base + P.arena_off;
					P.cig_out_len[rec] = (base == ~0ull || res.flags) ? 0 : res.n_ops;
				}
				continue;
			}
			__syncwarp();
			finish_record(P, rec, res, stage + g.stage_n);
			if ((int) lane == r) { g.my_rec = rec; g.my_off = g.stage_n; g.my_len = res.n_ops; }
			g.stage_n += res.n_ops;
		}
		flush_group(P, g, stage);
	}
}

// Assembly of the records cut into segments (pv3_rewrite_seg.cu): one warp per record concatenates its segments' op lists
// through cigar_buffer::push_back's merge rule (cigar.cc:61-86) into the CIGAR arena, then the usual post-rewrite
// bookkeeping (run detection, per-record outputs). A record with a failed segment goes to deferred2.
__global__ void __launch_bounds__(K1_WARPS * 32)
k1_assemble_kernel(K1Params P)
{
	if (*((volatile const unsigned int *) &P.counters->overflow) & (128u | 256u)) return;       // incomplete segment data: the host re-runs
	unsigned const lane = lane_id();
	uint64_t const n_items = (uint64_t) *((volatile const unsigned long long *) &P.counters->n_planned);
	uint64_t const warp_global = (uint64_t) blockIdx.x * K1_WARPS + (threadIdx.x >> 5), n_warps = (uint64_t) gridDim.x * K1_WARPS;
	const seg::SegOut *seg_out = (const seg::SegOut *) P.seg_out;
	for (uint64_t i = warp_global; i < n_items; i += n_warps) {
		uint64_t const rec = P.plan_rec[i], s0 = P.plan_seg0[i];
		uint32_t const ns = P.plan_nseg[i];
		// any failed segment? upper bound of the op count
		bool bad = false;
		uint32_t upper = 0;
		for (uint32_t b = 0; b < ns; b += 32) {
			uint32_t n = 0, fl = 0;
			if (b + lane < ns) { n = seg_out[s0 + b + lane].n_ops; fl = seg_out[s0 + b + lane].flags; }
			bad |= __any_sync(FULL, (fl & 2u) != 0);
			upper += __reduce_add_sync(FULL, n);
		}
		if (bad) {
			if (lane == 0) { unsigned long long const slot = atomicAdd(&P.counters->n_deferred2, 1ull); P.deferred2[slot] = rec; }
			continue;
		}
		RecordResult res; res.flags = 0; res.overflow = false; res.adjacent_id = false; res.n_ops = 0; res.dst_pos = 0;
		uint32_t const sr = __ldg(P.src_row + rec), dr = __ldg(P.dst_row + rec), rf = __ldg(P.dst_ref + rec);
		if (sr >= P.n_rows || dr >= P.n_rows || rf >= P.n_refs) res.flags = PV3_FLAG_BAD_ID;        // (a record without segments only)
		else {
			RowDesc const S = P.rows[sr], D = P.rows[dr];
			uint64_t const spos = __ldg(P.src_pos + rec);
			uint64_t const col = (spos < S.n_zeros) ? select0_warp(S, spos + 1) : S.n_bits;        // rewrite_cigar.hh:51-56
			res.dst_pos = rank0_warp(D, col);
			if (P.qual_off) {
				uint64_t const ql = __ldg(P.qual_off + rec + 1) - __ldg(P.qual_off + rec);
				if (ql && ql != query_length(P, rec)) res.flags |= PV3_FLAG_BAD_QUAL;
			}
		}
		uint64_t const base = arena_alloc(P, upper);
		if (base != ~0ull) {
			// 32 segments per step, lane = segment. A segment's first op merges with the open op iff it equals the last op of
			// the previous non-empty segment (an absorbed one-op segment has that same op); every other op gets a slot of its
			// own, and a merged first op adds its count to the slot written last before it — the head of its run.
			uint32_t *out = P.arena + base;
			uint32_t n_out = 0, carry_op = OP_NONE;
			bool adj = false;
			for (uint32_t b = 0; b < ns; b += 32) {
				seg::SegOut o; o.off = 0; o.n_ops = 0; o.flags = 0; o.first = 0; o.last = 0;
				if (b + lane < ns) o = seg_out[s0 + b + lane];
				unsigned const nonempty = __ballot_sync(FULL, o.n_ops != 0);
				unsigned const below = nonempty & lanemask_lt();
				uint32_t prev_last = __shfl_sync(FULL, o.last & 0xfu, below ? 31 - __clz(below) : 0);
				if (!below) prev_last = carry_op;
				bool const merges = o.n_ops != 0 && (o.first & 0xfu) == prev_last;
				uint32_t const w = o.n_ops - (merges ? 1u : 0u);
				uint32_t incl = w;
				#pragma unroll
				for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
				uint32_t const at = n_out + incl - w;                                        // this segment's first slot
				adj |= __any_sync(FULL, (o.flags & 1u) != 0 || (o.n_ops != 0 && !merges && (prev_last | (o.first & 0xfu)) == 3u));
				for (uint32_t j = merges ? 1u : 0u; j < o.n_ops; ++j) out[at + j - (merges ? 1u : 0u)] = P.seg_ops[o.off + j];
				__syncwarp();
				if (merges) atomicAdd(out + at - 1, (o.first >> 4) << 4);                    // at >= 1: an open op exists
				__syncwarp();
				n_out += __shfl_sync(FULL, incl, 31);
				if (nonempty) carry_op = __shfl_sync(FULL, o.last & 0xfu, 31 - __clz(nonempty));
			}
			res.n_ops = n_out; res.adjacent_id = adj;
			__threadfence_block();
			__syncwarp();
			finish_record(P, rec, res, out);
		} else {
			res.n_ops = 0;
			finish_record(P, rec, res, P.arena);
		}
		if (lane == 0) {
			P.cig_out_off[rec] = (base == ~0ull) ? 0 : base + P.arena_off;
			P.cig_out_len[rec] = (base == ~0ull || res.flags) ? 0 : res.n_ops;
		}
	}
}

void launch_k1_assemble(const K1Params &P, int n_sms, uint64_t expected, cudaStream_t stream)
{
	uint64_t ctas = expected / K1_WARPS + 4;
	if (ctas > (uint64_t) n_sms * 8) ctas = (uint64_t) n_sms * 8;
	k1_assemble_kernel<<<(unsigned) ctas, K1_WARPS * 32, 0, stream>>>(P);
}

void launch_k1_list(const K1Params &P, int n_sms, uint64_t expected, cudaStream_t stream)
{
	// list length known only on the device: a resident grid strides over it, one record per warp
	uint64_t ctas = expected / K1_WARPS + 8;
	if (ctas > (uint64_t) n_sms * 4) ctas = (uint64_t) n_sms * 4;
	k1_rewrite_kernel<<<(unsigned) ctas, K1_WARPS * 32, 0, stream>>>(P, 1);
}

void launch_k1(const K1Params &P, int n_sms, cudaStream_t stream)
{
	if (P.n_list_dev) {
		// deferred list of the thread kernel, length known only on the device: a resident grid strides over it, one record per warp
		k1_rewrite_kernel<<<(unsigned) n_sms * 4, K1_WARPS * 32, 0, stream>>>(P, 1);
		return;
	}
	uint64_t const n_items = P.rec_list ? P.n_list : P.n;
	// few records (the deferred list of the thread kernel): one record per warp so that they spread over the SMs
	int const group = (n_items < (uint64_t) n_sms * 8 * K1_WARPS * GROUP) ? 1 : GROUP;
	uint64_t const n_groups = (n_items + group - 1) / group;
	uint64_t blocks = (n_groups + K1_WARPS - 1) / K1_WARPS;
	uint64_t const max_blocks = (uint64_t) n_sms * 8;   // 8 resident CTAs of 256 threads per SM
	if (blocks > max_blocks) blocks = max_blocks;
	if (blocks == 0) return;
	k1_rewrite_kernel<<<(unsigned) blocks, K1_WARPS * 32, 0, stream>>>(P, group);
}

} // namespace pv3
