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
// three arrays, read and written through 32-column windows in shared memory; the trace is stored anti-diagonal-major (the 32
// cells of a step side by side, four steps per word: one coalesced store per four steps), a byte per cell for the affine
// modes and two bits for edit distance, in shared memory when it fits (8 KB per warp) and in an HBM slot otherwise.
// Integer ALU work only — not a dense contraction, no tensor cores.
#include "pv3_device.cuh"
#include "pv3_kernels.h"
#include <type_traits>

namespace pv3 {

constexpr int      K3_WARPS  = 8;
constexpr uint32_t SMEM_POOL = 8192;   // bytes per warp: a 64 x 64 affine trace with its boundary arrays fits (3 CTAs per SM either way: 80 registers)

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

// What lane 0 of a strip needs per step — row i0's score, vertical gap state and its trace bit at the step's column, and the
// reference base there — and what lane 31 leaves for the next strip: both go through these 32-column windows, so that the
// boundary arrays (HBM for wide windows) are touched by coalesced loads / stores once per 32 steps, one window ahead, and the
// per-step dependency chain never waits for global memory.
struct StripStage {
	int32_t  in_h[32], in_u[32];
	uint16_t in_tr[32];               // trace bit of the vertical state | reference rank << 8
	int32_t  out_h[64], out_u[64];    // rings over two blocks: a block's columns are flushed while the next block's arrive
	uint8_t  out_t[64];
};

uint32_t k3_smem_pool_bytes() { return SMEM_POOL; }
uint64_t k3_default_grid_warps(int n_sms) { return (uint64_t) n_sms * 3 * K3_WARPS; }     // 3 CTAs per SM are resident (80 registers, 72 KB of shared memory)

template <bool PACKED>   // PACKED: SEQ in BAM's 4-bit packing (K1Params::seq_len)
__global__ void __launch_bounds__(K3_WARPS * 32)
k3_realign_kernel(K3Params P)
{
	extern __shared__ __align__(16) uint8_t k3_smem[];
	K1Params const &B = P.k1;
	unsigned const lane = lane_id();
	unsigned const warp_in_block = threadIdx.x >> 5;
	uint8_t *const my_pool = k3_smem + (size_t) warp_in_block * SMEM_POOL;
	StripStage &S = *(StripStage *) (k3_smem + (size_t) K3_WARPS * SMEM_POOL + (size_t) warp_in_block * sizeof(StripStage));
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
		uint8_t *pool = (need <= SMEM_POOL) ? my_pool : (P.pool + warp_global * P.pool_slot_bytes);
		// Trace, anti-diagonal-major: strip s (rows 32 s + 1 ..), step t (lane l is at column t - l), lane — so the 32 cells of a
		// step are adjacent; four steps share a word per lane (affine: a byte per cell) or a byte (edit distance: two bits).
		uint32_t const groups = dp_trace_groups(R);
		uint8_t *T = pool;
		int32_t *Hb = (int32_t *) (pool + dp_trace_bytes(R, Q, !edit));
		int32_t *Ub = Hb + cols;
		uint32_t *Tb = (uint32_t *) (Ub + cols);      // (only when Q > 32)
		auto trace_at = [&](uint32_t i, uint32_t j) -> uint32_t {     // i, j >= 1
			uint32_t const s = (i - 1) >> 5, l = (i - 1) & 31u, u = j + l - 1;
			size_t const w = ((size_t) s * groups + (u >> 2)) * 32 + l;
			return edit ? ((uint32_t) T[w] >> (2 * (u & 3u))) & 3u : (uint32_t) T[w * 4 + (u & 3u)];
		};
		// row 0 of the matrices (strip 0's lane 0 reads these instead of the boundary arrays)
		auto row0_h = [&](uint32_t j) -> int32_t { return edit ? (int32_t) j : (j ? gopen + (int32_t) (j - 1) * gext : 0); };

		// One strip of 32 query rows. Steps u = 0 .. R + nl - 2 (lane l is at column u + 1 - l), four per group (one trace word),
		// 32 per block (one window of lane-0 inputs, one flush of lane-31 outputs); the tail of the last block runs past the
		// matrix with every lane inactive. Specialised per scoring mode: the step is ~40 instructions, a mode test per step shows.
		auto fill_strip = [&](auto edit_tag, uint32_t i0) {
			constexpr bool EDIT = decltype(edit_tag)::value;
			uint32_t const i = i0 + 1 + lane;
			bool const row_ok = i <= Q;
			uint32_t const nl = (Q - i0 < 32u) ? (Q - i0) : 32u;
			uint32_t const my_cols = row_ok ? R : 0u;              // lane is active while u - lane < my_cols (unsigned)
			uint32_t const qi = row_ok ? qrank(i - 1) : 0xffu;
			int32_t s_match, s_mis;
			if (EDIT) { s_match = 0; s_mis = 1; }
			else if (has_qual) {
				uint32_t ql = row_ok ? __ldg(qlptr + i - 1) : 0; if (ql > 41) ql = 41;
				s_match = P.lut[2 * ql]; s_mis = P.lut[2 * ql + 1];
			} else { s_match = 0; s_mis = -1; }
			int32_t Hcur, Hprev = 0, left = 0, up_out = 0;
			uint32_t tleft = T_LO, tup_out = 0, my_ref = 0;
			if (EDIT) { Hcur = (int32_t) i; up_out = Hcur; }
			else {
				Hcur = gopen + (int32_t) (i - 1) * gext;          // column 0 (affine_gap_init_policy.hpp)
				left = Hcur + gopen;
			}
			bool const first = i0 == 0, feeds = i0 + 32 < Q;
			uint32_t const n_steps = R ? R + nl - 1 : 0;
			size_t const strip_base = (size_t) (i0 >> 5) * groups * 32;
			// the window of columns c = c0 + lane that lane 0 walks through one block from now: H[i0][c - 1], U[i0][c], its trace bit, ref[c - 1]
			int32_t pre_h = 0, pre_u = 0; uint32_t pre_tr = 0;
			auto load_window = [&](uint32_t c0) {
				uint32_t const c = c0 + lane;
				if (c > R) return;
				uint32_t const rr = dna5_rank_of_char(__ldg(rbase + c - 1));          // alignment_projector.cc:78-82
				if (first) {
					pre_h = row0_h(c - 1);
					pre_u = EDIT ? (int32_t) c : row0_h(c) + gopen;
					pre_tr = (EDIT ? 0u : (uint32_t) T_UO) | (rr << 8);
				} else {
					pre_h = (c == 1) ? (EDIT ? (int32_t) i0 : gopen + (int32_t) (i0 - 1) * gext) : Hb[c - 1];   // H[i0][0] feeds the diagonal of (i0 + 1, 1)
					pre_u = Ub[c];
					pre_tr = Tb[c] | (rr << 8);
				}
			};
			auto flush_out = [&](uint32_t c0) {                    // columns c0 .. c0 + 31 of lane 31's row, as far as they exist
				__syncwarp();
				uint32_t const c = c0 + lane, k = (c - 1) & 63u;
				if (c <= R) { Hb[c] = S.out_h[k]; Ub[c] = S.out_u[k]; Tb[c] = S.out_t[k]; }
				__syncwarp();
			};
			load_window(1);
			for (uint32_t ub = 0; ub < n_steps; ub += 32) {
				__syncwarp();
				S.in_h[lane] = pre_h; S.in_u[lane] = pre_u; S.in_tr[lane] = (uint16_t) pre_tr;
				__syncwarp();
				load_window(ub + 33);
				#pragma unroll 1
				for (uint32_t g = 0; g < 8; ++g) {
					uint32_t acc = 0;
					#pragma unroll
					for (uint32_t k = 0; k < 4; ++k) {
						uint32_t const u = ub + g * 4 + k;
						int32_t dH = __shfl_up_sync(FULL, Hprev, 1);
						int32_t uI = __shfl_up_sync(FULL, up_out, 1);
						uint32_t pk = __shfl_up_sync(FULL, tup_out | (my_ref << 8), 1);
						if (lane == 0) { uint32_t const w = g * 4 + k; dH = S.in_h[w]; uI = S.in_u[w]; pk = S.in_tr[w]; }
						if (u - lane < my_cols) {
							my_ref = pk >> 8;
							uint32_t const tI = pk & 0xffu;
							int32_t const sc = (my_ref == qi) ? s_match : s_mis;
							uint32_t tr;
							if (EDIT) {
								int32_t const lft = Hcur + 1, upp = uI + 1;
								int32_t best = dH + sc;
								if (upp < best) best = upp;
								if (lft < best) best = lft;
								tr = (best == lft ? 1u : 0u) | (best == upp ? 2u : 0u);
								Hprev = Hcur; Hcur = best; up_out = best;
								acc |= tr << (2 * k);
							} else {
								int32_t tmp = dH + sc;
								if (tmp < uI) { tr = tI; tmp = uI; } else tr = T_D | tI;
								if (tmp < left) { tr = tleft; tmp = left; } else tr |= tleft;
								Hprev = Hcur; Hcur = tmp;
								tmp += gopen;
								int32_t const upe = uI + gext, wl = left + gext;
								if (upe < tmp) { tup_out = T_UO; up_out = tmp; } else { tup_out = T_U; up_out = upe; }
								if (wl < tmp) { tleft = T_LO; left = tmp; } else { tleft = T_L; left = wl; }
								acc |= tr << (8 * k);
							}
							if (feeds && lane == 31) { uint32_t const w = (u - 31u) & 63u; S.out_h[w] = Hcur; S.out_u[w] = up_out; S.out_t[w] = (uint8_t) tup_out; }   // column u - 30
						}
					}
					size_t const w = strip_base + (size_t) ((ub >> 2) + g) * 32 + lane;
					if (EDIT) T[w] = (uint8_t) acc; else ((uint32_t *) T)[w] = acc;
				}
				if (feeds && ub) flush_out(ub - 31);               // lane 31 reached column ub + 1: columns ub - 31 .. ub are complete
			}
			if (feeds && n_steps) flush_out(((n_steps - 1) & ~31u) - 31u + 32u);
			__syncwarp();
		};

		__syncwarp();
		for (uint32_t i0 = 0; i0 < Q; i0 += 32) {
			if (edit) fill_strip(std::true_type{}, i0);
			else fill_strip(std::false_type{}, i0);
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
		auto diag_op = [&]() -> uint32_t { return (dna5_rank_of_char(__ldg(rbase + j - 1)) == qrank(i - 1)) ? OP_EQ : OP_X; };   // align.hh:104-107

		if (edit) {
			while (i > 0 || j > 0) {
				uint32_t const tr = (i > 0 && j > 0) ? trace_at(i, j) : 0u;
				if (j > 0 && (i == 0 || (tr & 1u))) { emit(OP_D); --j; }          // Left  ⇒ gap in query ⇒ D
				else if (i > 0 && (j == 0 || (tr & 2u))) { emit(OP_I); --i; }     // Up    ⇒ gap in reference ⇒ I
				else { emit(diag_op()); --i; --j; }
			}
		} else {
			// row 0 and column 0 of the trace matrix are not stored (affine_gap_init_policy.hpp: open at the first cell, extension after it)
			auto trace_full = [&](uint32_t ii, uint32_t jj) -> uint32_t {
				if (ii == 0) return jj ? (jj == 1 ? (uint32_t) T_LO : (uint32_t) T_L) : 0u;
				if (jj == 0) return ii == 1 ? (uint32_t) T_UO : (uint32_t) T_U;
				return trace_at(ii, jj);
			};
			uint32_t tr = trace_full(i, j);
			int dir = trace_pick(tr);
			uint32_t steps = 0;
			while (tr != 0) {
				if ((dir == DIR_UP && i == 0) || (dir == DIR_LEFT && j == 0) || (dir == DIR_DIAG && (i == 0 || j == 0))
					|| dir == DIR_NONE || steps > reserve) { ub = true; break; }
				uint32_t const old = tr;
				if (dir == DIR_UP) { emit(OP_I); --i; tr = trace_full(i, j); if (old & T_UO) dir = trace_pick(tr); }
				else if (dir == DIR_LEFT) { emit(OP_D); --j; tr = trace_full(i, j); if (old & T_LO) dir = trace_pick(tr); }
				else { emit(diag_op()); --i; --j; tr = trace_full(i, j); dir = trace_pick(tr); }
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
	size_t const bytes = (size_t) K3_WARPS * (SMEM_POOL + sizeof(StripStage));
	if (P.k1.seq_len) {
		cudaFuncSetAttribute(k3_realign_kernel <true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes);      // (per device: cheap, every launch)
		k3_realign_kernel <true><<<blocks, K3_WARPS * 32, bytes, stream>>>(P);
	} else {
		cudaFuncSetAttribute(k3_realign_kernel <false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes);
		k3_realign_kernel <false><<<blocks, K3_WARPS * 32, bytes, stream>>>(P);
	}
}

} // namespace pv3
