// panvc3_b200 — kernel K3: realignment of indel runs (global pairwise alignment + traceback), one warp per run (sm_100a).
//
// Restates the three algorithms panvc3::align_global (reference include/panvc3/align.hh:66-113) reaches through
// seqan3::align_pairwise, bit-exactly including tie-breaks:
//   EDIT    qual empty && gap_open == 0  → seqan3 edit distance (alignment_configurator.hpp:323-349); unit costs,
//           gap_extension ignored; traceback priority Left > Up > Diagonal (edit_distance_trace_matrix_full.hpp:273-286)
//   AFFINE  otherwise → alignment_algorithm + affine_gap_policy (affine_gap_policy.hpp:87-131), int32 scores, 5-bit
//           trace, stateful traceback (trace_iterator_base.hpp:150-179,254-274). Scores: 0/-1 on dna5 rank equality, or
//           the int32-truncated Malde LUT indexed by the query base's phred (align.hh:28-37; reference at phred 41)
//
// Mapping: DP rows = query bases, DP columns = reference window (seqan3 sequence1 = reference).  The reference fills
// column by column; the recurrence only needs (i-1,j-1), (i-1,j), (i,j-1), so a skewed wavefront gives the same values:
// lane l owns row i0+1+l of a 32-row strip and visits column j at step t = j + l; the vertical gap state and the diagonal
// score arrive from lane l-1 by shuffle, the horizontal gap state stays in the lane.  Strip boundaries (row i0) live in
// three arrays; the byte trace matrix lives in shared memory when it fits (6 KB per warp) and in an HBM slot otherwise.
// Integer ALU work only — not a dense contraction, no tensor cores.
#include "pv3_device.cuh"
#include "pv3_kernels.h"

namespace pv3 {

constexpr int      K3_WARPS  = 8;
constexpr uint32_t SMEM_POOL = 6144;   // bytes per warp

enum : uint32_t { T_D = 1, T_UO = 2, T_U = 4, T_LO = 8, T_L = 16 };
enum : int { DIR_NONE = -1, DIR_DIAG = 0, DIR_UP = 1, DIR_LEFT = 2 };

// seqan3 dna5::assign_char: case-insensitive, U→T, everything else → N (dna5.hpp:99-110)
__device__ __forceinline__ uint8_t dna5_rank_of_char(char c)
{
	switch (c) {
		case 'A': case 'a': return 0;
		case 'C': case 'c': return 1;
		case 'G': case 'g': return 2;
		case 'T': case 't': case 'U': case 'u': return 4;
		default: return 3;
	}
}

__device__ __forceinline__ int trace_pick(uint32_t t)
{
	return (t & T_D) ? DIR_DIAG : (t & (T_U | T_UO)) ? DIR_UP : (t & (T_L | T_LO)) ? DIR_LEFT : DIR_NONE;
}

uint32_t k3_smem_pool_bytes() { return SMEM_POOL; }
uint64_t k3_default_grid_warps(int n_sms) { return (uint64_t) n_sms * 4 * K3_WARPS; }

template <bool PACKED>   // PACKED: SEQ in BAM's 4-bit packing (K1Params::seq_len)
__global__ void __launch_bounds__(K3_WARPS * 32)
k3_realign_kernel(K3Params P)
{
	__shared__ __align__(16) uint8_t smem_pool[K3_WARPS][SMEM_POOL];
	K1Params const &B = P.k1;
	unsigned const lane = lane_id();
	unsigned const warp_in_block = threadIdx.x >> 5;
	uint64_t const warp_global = (uint64_t) blockIdx.x * K3_WARPS + warp_in_block;
	int32_t const gext = B.gap_ext, gopen = B.gap_open + B.gap_ext;   // affine_gap_policy.hpp:183-190
	// run counts from the device (no host round trip after the rewrite); small runs were realigned by the thread-per-run
	// kernel iff it decided to run (same counters, same rule): then only its list of large runs is left
	uint64_t const n_jobs = dp_n_jobs(B.counters, P.n_jobs);
	bool const only_big = P.big_list && dp_use_thread_kernel(B.counters, n_jobs, P.dp_thread_min);
	uint64_t const n_mine = only_big ? (uint64_t) B.counters->big_cursor : n_jobs;

	for (;;) {
		unsigned long long jid = 0;
		if (lane == 0) jid = atomicAdd(&B.counters->job_ticket, 1ull);
		jid = shfl_u64(jid, 0);
		if (jid >= n_mine) break;
		if (only_big) jid = P.big_list[jid];

		Job const job = B.jobs[jid];
		uint64_t const rec = job.rec;
		uint32_t const R = job.ref_len, Q = job.qry_len;
		uint32_t const cols = R + 1;
		const char *rbase = B.refs[__ldg(B.dst_ref + rec)].bases + job.ref_loc;
		uint32_t const qloc = job.qry_loc;
		const uint8_t *qptr = B.seq + __ldg(B.seq_off + rec) + (PACKED ? 0u : qloc);
		auto qrank = [&](uint32_t w) -> uint8_t {                       // rank of base w of the run's query window
			if (!PACKED) return __ldg(qptr + w);
			uint32_t const g = qloc + w, b = __ldg(qptr + (g >> 1));
			return (uint8_t) bam4_to_rank((g & 1u) ? b : (b >> 4));
		};
		bool has_qual = false;
		const uint8_t *qlptr = nullptr;
		if (B.qual_off) {
			uint64_t const q0 = __ldg(B.qual_off + rec);
			has_qual = __ldg(B.qual_off + rec + 1) > q0;
			qlptr = P.job_qual_off ? (P.packed_qual + P.job_qual_off[jid]) : (B.qual + q0 + job.qry_loc);
		}
		bool const edit = !has_qual && B.gap_open == 0;

		unsigned long long const need = dp_pool_bytes(R, Q, !edit);
		if (need > SMEM_POOL && (!P.pool || need > P.pool_slot_bytes || warp_global >= P.pool_slots)) {
			if (lane == 0) atomicOr(&B.counters->overflow, 8u);      // HBM trace slots missing or too small: the host grows them and re-runs
			continue;
		}
		uint8_t *pool = (need <= SMEM_POOL) ? smem_pool[warp_in_block] : (P.pool + warp_global * P.pool_slot_bytes);
		uint32_t const trace_bytes = (uint32_t) ((((unsigned long long) (Q + 1) * cols) + 3ull) & ~3ull);
		uint8_t *T = pool;
		int32_t *Hb = (int32_t *) (pool + trace_bytes);
		int32_t *Ub = Hb + cols;
		uint8_t *Tb = (uint8_t *) (Ub + cols);
		uint8_t *rs = Tb + ((cols + 3u) & ~3u);

		__syncwarp();
		for (uint32_t j = lane; j < R; j += 32) rs[j] = dna5_rank_of_char(__ldg(rbase + j));   // alignment_projector.cc:78-82
		for (uint32_t j = lane; j <= R; j += 32) {
			if (edit) { Hb[j] = (int32_t) j; Ub[j] = (int32_t) j; Tb[j] = 0; }
			else {
				int32_t const h = j ? gopen + (int32_t) (j - 1) * gext : 0;      // row 0 of the score matrix
				Hb[j] = h; Ub[j] = h + gopen; Tb[j] = T_UO;
				T[j] = j ? (j == 1 ? T_LO : T_L) : 0;
			}
		}
		__syncwarp();

		for (uint32_t i0 = 0; i0 < Q; i0 += 32) {
			uint32_t const i = i0 + 1 + lane;
			bool const row_ok = i <= Q;
			uint32_t const nl = (Q - i0 < 32u) ? (Q - i0) : 32u;
			uint8_t const qi = row_ok ? qrank(i - 1) : 0;
			int32_t s_match, s_mis;
			if (edit) { s_match = 0; s_mis = 1; }
			else if (has_qual) {
				uint32_t ql = row_ok ? __ldg(qlptr + i - 1) : 0; if (ql > 41) ql = 41;
				s_match = P.lut[2 * ql]; s_mis = P.lut[2 * ql + 1];
			} else { s_match = 0; s_mis = -1; }
			int32_t Hcur, Hprev = 0, left = 0, up_out = 0;
			uint32_t tleft = T_LO, tup_out = 0;
			if (edit) { Hcur = (int32_t) i; up_out = Hcur; }
			else {
				Hcur = gopen + (int32_t) (i - 1) * gext;          // column 0 (affine_gap_init_policy.hpp)
				left = Hcur + gopen;
				if (row_ok) T[(uint64_t) i * cols] = (i == 1) ? T_UO : T_U;
			}
			bool const last_lane_feeds = (lane == 31) && (i0 + 32 < Q);
			uint32_t const n_steps = R + nl - 1;
			for (uint32_t t = 1; t <= n_steps && R; ++t) {
				int32_t dH = __shfl_up_sync(FULL, Hprev, 1);
				int32_t uI = __shfl_up_sync(FULL, up_out, 1);
				uint32_t tI = __shfl_up_sync(FULL, tup_out, 1);
				int32_t const jc = (int32_t) t - (int32_t) lane;
				bool const act = row_ok && jc >= 1 && jc <= (int32_t) R;
				if (lane == 0 && act) { dH = Hb[jc - 1]; uI = Ub[jc]; tI = Tb[jc]; }
				if (act) {
					int32_t const s = (rs[jc - 1] == qi) ? s_match : s_mis;
					uint32_t tr;
					if (edit) {
						int32_t const lft = Hcur + 1, upp = uI + 1;
						int32_t best = dH + s;
						if (upp < best) best = upp;
						if (lft < best) best = lft;
						tr = (best == lft ? 1u : 0u) | (best == upp ? 2u : 0u);
						Hprev = Hcur; Hcur = best; up_out = best;
					} else {
						int32_t tmp = dH + s;
						if (tmp < uI) { tr = tI; tmp = uI; } else tr = T_D | tI;
						if (tmp < left) { tr = tleft; tmp = left; } else tr |= tleft;
						Hprev = Hcur; Hcur = tmp;
						tmp += gopen;
						int32_t const upe = uI + gext, wl = left + gext;
						if (upe < tmp) { tup_out = T_UO; up_out = tmp; } else { tup_out = T_U; up_out = upe; }
						if (wl < tmp) { tleft = T_LO; left = tmp; } else { tleft = T_L; left = wl; }
					}
					T[(uint64_t) i * cols + jc] = (uint8_t) tr;
					if (last_lane_feeds) { Hb[jc] = Hcur; Ub[jc] = up_out; Tb[jc] = (uint8_t) tup_out; }
				}
			}
			__syncwarp();
			if (lane == 0 && i0 + 32 < Q) {
				uint32_t const ib = i0 + 32;                       // H[ib][0] feeds the diagonal of (ib+1, 1)
				Hb[0] = edit ? (int32_t) ib : gopen + (int32_t) (ib - 1) * gext;
			}
			__syncwarp();
		}

		// ---- traceback (warp-uniform) + backward run-length emission into the reserved R+Q words -------------------
		unsigned long long obase = 0;
		uint32_t const reserve = R + Q;
		if (lane == 0) obase = atomicAdd(&B.counters->dp_arena_cursor, (unsigned long long) reserve);
		obase = shfl_u64(obase, 0);
		if (obase + reserve > P.dp_arena_cap) {
			if (lane == 0) atomicOr(&B.counters->overflow, 4u);
			continue;
		}
		uint32_t *out = P.dp_arena + obase;
		uint32_t wp = reserve, cur_op = OP_NONE, cur_cnt = 0;
		uint32_t i = Q, j = R;
		bool ub = false;
		auto emit = [&](uint32_t op) {
			if (op == cur_op) { ++cur_cnt; return; }
			if (cur_op != OP_NONE) { --wp; if (lane == 0) out[wp] = (cur_cnt << 4) | cur_op; }
			cur_op = op; cur_cnt = 1;
		};
		auto diag_op = [&]() -> uint32_t { return (rs[j - 1] == qrank(i - 1)) ? OP_EQ : OP_X; };   // align.hh:104-107

		if (edit) {
			while (i > 0 || j > 0) {
				uint32_t const tr = (i > 0 && j > 0) ? T[(uint64_t) i * cols + j] : 0u;
				if (j > 0 && (i == 0 || (tr & 1u))) { emit(OP_D); --j; }          // Left  ⇒ gap in query ⇒ D
				else if (i > 0 && (j == 0 || (tr & 2u))) { emit(OP_I); --i; }     // Up    ⇒ gap in reference ⇒ I
				else { emit(diag_op()); --i; --j; }
			}
		} else {
			uint32_t tr = T[(uint64_t) i * cols + j];
			int dir = trace_pick(tr);
			uint32_t steps = 0;
			while (tr != 0) {
				if ((dir == DIR_UP && i == 0) || (dir == DIR_LEFT && j == 0) || (dir == DIR_DIAG && (i == 0 || j == 0))
					|| dir == DIR_NONE || steps > reserve) { ub = true; break; }
				uint32_t const old = tr;
				if (dir == DIR_UP) { emit(OP_I); --i; tr = T[(uint64_t) i * cols + j]; if (old & T_UO) dir = trace_pick(tr); }
				else if (dir == DIR_LEFT) { emit(OP_D); --j; tr = T[(uint64_t) i * cols + j]; if (old & T_LO) dir = trace_pick(tr); }
				else { emit(diag_op()); --i; --j; tr = T[(uint64_t) i * cols + j]; dir = trace_pick(tr); }
				++steps;
			}
			if (!ub && (i || j)) ub = true;
		}
		if (cur_op != OP_NONE) { --wp; if (lane == 0) out[wp] = (cur_cnt << 4) | cur_op; }
		if (lane == 0) {
			B.jobs[jid].out_off = obase + wp;
			B.jobs[jid].out_len = reserve - wp;
			if (ub) {   // reference UB: flag the record, do not emulate
				uint32_t const old = atomicOr(&B.flags[rec], PV3_FLAG_TRACE_UB);
				if (!old) atomicAdd(&B.counters->n_flagged, 1ull);
			}
		}
	}
}

void launch_k3(const K3Params &P, int n_sms, uint64_t grid_warps, cudaStream_t stream)
{
	uint64_t const n = P.n_jobs;          // ~0: read on the device
	if (!n) return;
	uint64_t warps = grid_warps ? grid_warps : k3_default_grid_warps(n_sms);
	if (warps > n) warps = n;
	unsigned const blocks = (unsigned) ((warps + K3_WARPS - 1) / K3_WARPS);
	if (P.k1.seq_len) k3_realign_kernel <true><<<blocks, K3_WARPS * 32, 0, stream>>>(P);
	else k3_realign_kernel <false><<<blocks, K3_WARPS * 32, 0, stream>>>(P);
}

} // namespace pv3
