// panvc3_b200 — kernel K3T: realignment of SMALL indel runs, one THREAD per run (sm_100a).
//
// Same three algorithms and tie-breaks as the warp-per-run kernel in pv3_realign.cu (reference include/panvc3/align.hh:66-113
// → seqan3 edit distance / affine_gap_policy.hpp:87-131 with trace_iterator_base.hpp:150-179,254-274 traceback).
// Indel runs that come from MSA gap columns are a few bases wide (bench workloads: 17-26 DP cells per run): a warp per run
// leaves most lanes idle and pays three atomics per run. Here a lane runs the whole recurrence (which only needs row i-1)
// out of its local arrays, in FLAT loops — one over the cells, one over the traceback steps — so that the lanes of a warp,
// which hold runs of different shapes, stay converged; a warp pays one atomic for the op space of its 32 runs.
// Larger runs (dp_job_is_small: a side above 15 or more than 64 trace bytes) are left to the warp kernel — a lane waits for
// the largest run of its warp, and run sizes are heavy-tailed — through a compact list this kernel writes.
#include "pv3_device.cuh"
#include "pv3_kernels.h"

namespace pv3 {

namespace {

enum : uint32_t { TT_D = 1, TT_UO = 2, TT_U = 4, TT_LO = 8, TT_L = 16 };
enum : int { TD_NONE = -1, TD_DIAG = 0, TD_UP = 1, TD_LEFT = 2 };

// seqan3 dna5::assign_char: case-insensitive, U→T, everything else → N (dna5.hpp:99-110)
__device__ __forceinline__ uint8_t rank_of_char(char c)
{
	switch (c) {
		case 'A': case 'a': return 0;
		case 'C': case 'c': return 1;
		case 'G': case 'g': return 2;
		case 'T': case 't': case 'U': case 'u': return 4;
		default: return 3;
	}
}

__device__ __forceinline__ int pick(uint32_t t)
{
	return (t & TT_D) ? TD_DIAG : (t & (TT_U | TT_UO)) ? TD_UP : (t & (TT_L | TT_LO)) ? TD_LEFT : TD_NONE;
}

} // namespace

constexpr unsigned K3T_THREADS = 256;

template <bool PACKED>   // PACKED: SEQ in BAM's 4-bit packing (K1Params::seq_len)
__global__ void __launch_bounds__(K3T_THREADS)
k3_thread_kernel(K3Params P)
{
	// A warp runs as long as its largest run (round-2 capture: 9.2 of 32 lanes active, runs of 1 to 49 cells side by side), so
	// the CTA first sorts its 256 consecutive runs by cell count (counting sort in shared memory): the warps then hold
	// octiles of the size distribution instead of random mixtures.
	__shared__ uint32_t s_hist[64];
	__shared__ uint16_t s_perm[K3T_THREADS];
	K1Params const &B = P.k1;
	unsigned const lane = lane_id();
	int32_t const gext = B.gap_ext, gopen = B.gap_open + B.gap_ext;   // affine_gap_policy.hpp:183-190
	// The run count is read on the device (the rewrite kernels have finished on this stream): no host round trip between the
	// rewrite and the DP. Few small runs: the warp-per-run kernel takes them all (it reads the same counters and decides alike).
	uint64_t const n_jobs = dp_n_jobs(B.counters, P.n_jobs);
	if (!dp_use_thread_kernel(B.counters, n_jobs, P.dp_thread_min)) return;
	uint64_t const stride = (uint64_t) gridDim.x * K3T_THREADS;
	for (uint64_t cbase = (uint64_t) blockIdx.x * K3T_THREADS; cbase < n_jobs; cbase += stride) {
	{
		uint64_t const j0 = cbase + threadIdx.x;
		uint32_t key = 63;                                           // past the end
		if (j0 < n_jobs) {
			uint32_t const r = B.jobs[j0].ref_len, q = B.jobs[j0].qry_len;
			key = dp_job_is_small(r, q) ? r * q : 62u;                 // (R + 1)(Q + 1) <= 64 ⇒ R Q <= 49; the warp kernel's runs together
		}
		if (threadIdx.x < 64) s_hist[threadIdx.x] = 0;
		__syncthreads();
		uint32_t const in_bin = atomicAdd(&s_hist[key], 1u);
		__syncthreads();
		if (threadIdx.x < 32) {
			uint32_t const a = s_hist[2 * lane], b = s_hist[2 * lane + 1];
			uint32_t incl = a + b;
			#pragma unroll
			for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
			s_hist[2 * lane] = incl - a - b; s_hist[2 * lane + 1] = incl - b;
		}
		__syncthreads();
		s_perm[s_hist[key] + in_bin] = (uint16_t) threadIdx.x;
		__syncthreads();
	}
	uint64_t const jid = cbase + s_perm[threadIdx.x];

	uint32_t R = 0, Q = 0, qry_loc = 0;
	uint64_t rec = 0, ref_loc = 0;
	bool mine = false;
	if (jid < n_jobs) {
		Job const &job = B.jobs[jid];
		R = job.ref_len; Q = job.qry_len; rec = job.rec; ref_loc = job.ref_loc; qry_loc = job.qry_loc;
		mine = dp_job_is_small(R, Q);
	}
	// runs left to the warp kernel: compact list, one atomic per warp
	{
		uint32_t const m_big = __ballot_sync(FULL, jid < n_jobs && !mine);
		if (m_big && P.big_list) {
			unsigned long long b = 0;
			if (lane == 0) b = atomicAdd(&B.counters->big_cursor, (unsigned long long) __popc(m_big));
			b = shfl_u64(b, 0);
			if (jid < n_jobs && !mine) P.big_list[b + __popc(m_big & lanemask_lt())] = jid;
		}
	}
	// op space for the warp's runs (R + Q words each, filled from the end): one atomic
	uint32_t const reserve = mine ? R + Q : 0u;
	uint32_t incl = reserve;
	#pragma unroll
	for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
	uint32_t const total = __shfl_sync(FULL, incl, 31);
	unsigned long long wbase = 0;
	if (lane == 0 && total) {
		wbase = atomicAdd(&B.counters->dp_arena_cursor, (unsigned long long) total);
		if (wbase + total > P.dp_arena_cap) { atomicOr(&B.counters->overflow, 4u); wbase = ~0ull; }
	}
	wbase = shfl_u64(wbase, 0);
	if (!mine || wbase == ~0ull) continue;
	uint32_t *out = P.dp_arena + wbase + (incl - reserve);

	uint32_t const cols = R + 1;
	const char *rbase = B.refs[__ldg(B.dst_ref + rec)].bases + ref_loc;
	const uint8_t *qptr = B.seq + __ldg(B.seq_off + rec) + (PACKED ? 0u : qry_loc);
	auto qrank = [&](uint32_t w) -> uint8_t {                       // rank of base w of the run's query window
		if (!PACKED) return __ldg(qptr + w);
		uint32_t const g = qry_loc + w, b = __ldg(qptr + (g >> 1));
		return (uint8_t) bam4_to_rank((g & 1u) ? b : (b >> 4));
	};
	bool has_qual = false;
	const uint8_t *qlptr = nullptr;
	if (B.qual_off) {
		uint64_t const q0 = __ldg(B.qual_off + rec);
		has_qual = __ldg(B.qual_off + rec + 1) > q0;
		qlptr = P.job_qual_off ? (P.packed_qual + P.job_qual_off[jid]) : (B.qual + q0 + qry_loc);
	}
	bool const edit = !has_qual && B.gap_open == 0;

	uint8_t  T[K3T_MAX_TRACE];                                   // trace, row-major (Q + 1) x (R + 1)
	int32_t  Hr[K3T_MAX_DIM + 1], Ur[K3T_MAX_DIM + 1];           // row i-1: score, vertical gap score
	uint8_t  Tu[K3T_MAX_DIM + 1], rs[K3T_MAX_DIM + 1];           // row i-1: vertical trace bits; reference ranks

	for (uint32_t j = 0; j <= R; ++j) {
		rs[j] = (j < R) ? rank_of_char(__ldg(rbase + j)) : 0;                           // alignment_projector.cc:78-82
		int32_t const h = j ? gopen + (int32_t) (j - 1) * gext : 0;                    // row 0 of the score matrix
		Hr[j] = edit ? (int32_t) j : h;
		Ur[j] = edit ? (int32_t) j : h + gopen;
		Tu[j] = edit ? 0 : TT_UO;
		T[j] = edit ? 0 : (j ? (j == 1 ? TT_LO : TT_L) : 0);
	}

	// ---- all cells in one flat loop (row-major; the recurrence is order-independent given row i-1 and the cell to the left) --
	{
		uint32_t i = 1, j = 1, tleft = TT_LO;
		int32_t diag = 0, Hcur = 0, left = 0, s_match = 0, s_mis = 0;
		uint8_t qi = 0;
		uint32_t const n_cells = R * Q;
		for (uint32_t c = 0; c < n_cells; ++c) {
			if (j == 1) {                                                                 // row start
				qi = qrank(i - 1);
				if (edit) { s_match = 0; s_mis = 1; Hcur = (int32_t) i; }
				else {
					if (has_qual) { uint32_t ql = __ldg(qlptr + i - 1); if (ql > 41) ql = 41; s_match = P.lut[2 * ql]; s_mis = P.lut[2 * ql + 1]; }
					else { s_match = 0; s_mis = -1; }
					Hcur = gopen + (int32_t) (i - 1) * gext;                                // column 0 (affine_gap_init_policy.hpp)
					left = Hcur + gopen; tleft = TT_LO;
					T[i * cols] = (i == 1) ? TT_UO : TT_U;
				}
				diag = Hr[0]; Hr[0] = Hcur;
			}
			int32_t const dH = diag, uI = Ur[j];
			uint32_t const tI = Tu[j];
			diag = Hr[j];
			int32_t const s = (rs[j - 1] == qi) ? s_match : s_mis;
			uint32_t tr;
			if (edit) {
				int32_t const lft = Hcur + 1, upp = uI + 1;
				int32_t best = dH + s;
				if (upp < best) best = upp;
				if (lft < best) best = lft;
				tr = (best == lft ? 1u : 0u) | (best == upp ? 2u : 0u);
				Hcur = best; Ur[j] = best;
			} else {
				int32_t tmp = dH + s;
				if (tmp < uI) { tr = tI; tmp = uI; } else tr = TT_D | tI;
				if (tmp < left) { tr = tleft; tmp = left; } else tr |= tleft;
				Hcur = tmp;
				tmp += gopen;
				int32_t const upe = uI + gext, wl = left + gext;
				if (upe < tmp) { Tu[j] = TT_UO; Ur[j] = tmp; } else { Tu[j] = TT_U; Ur[j] = upe; }
				if (wl < tmp) { tleft = TT_LO; left = tmp; } else { tleft = TT_L; left = wl; }
			}
			Hr[j] = Hcur;
			T[i * cols + j] = (uint8_t) tr;
			if (++j > R) { j = 1; ++i; }
		}
	}

	// ---- traceback: one flat loop, ops run-length encoded backwards into the reserved words -----------------------------
	uint32_t wp = reserve, cur_op = OP_NONE, cur_cnt = 0;
	uint32_t i = Q, j = R;
	bool ub = false;
	{
		uint32_t tr = T[i * cols + j];
		int dir = edit ? TD_NONE : pick(tr);
		for (uint32_t steps = 0; ; ++steps) {
			uint32_t op;
			if (edit) {
				if (i == 0 && j == 0) break;
				uint32_t const t = (i > 0 && j > 0) ? T[i * cols + j] : 0u;
				if (j > 0 && (i == 0 || (t & 1u))) { op = OP_D; --j; }                      // Left  ⇒ gap in query ⇒ D
				else if (i > 0 && (j == 0 || (t & 2u))) { op = OP_I; --i; }                 // Up    ⇒ gap in reference ⇒ I
				else { op = (rs[j - 1] == qrank(i - 1)) ? OP_EQ : OP_X; --i; --j; }   // align.hh:104-107
			} else {
				if (tr == 0) break;
				if ((dir == TD_UP && i == 0) || (dir == TD_LEFT && j == 0) || (dir == TD_DIAG && (i == 0 || j == 0))
					|| dir == TD_NONE || steps > reserve) { ub = true; break; }
				uint32_t const old = tr;
				if (dir == TD_UP) { op = OP_I; --i; tr = T[i * cols + j]; if (old & TT_UO) dir = pick(tr); }
				else if (dir == TD_LEFT) { op = OP_D; --j; tr = T[i * cols + j]; if (old & TT_LO) dir = pick(tr); }
				else { op = (rs[j - 1] == qrank(i - 1)) ? OP_EQ : OP_X; --i; --j; tr = T[i * cols + j]; dir = pick(tr); }
			}
			if (op == cur_op) ++cur_cnt;
			else {
				if (cur_op != OP_NONE) out[--wp] = (cur_cnt << 4) | cur_op;
				cur_op = op; cur_cnt = 1;
			}
		}
		if (!edit && !ub && (i || j)) ub = true;
	}
	if (cur_op != OP_NONE) out[--wp] = (cur_cnt << 4) | cur_op;
	B.jobs[jid].out_off = wbase + (incl - reserve) + wp;
	B.jobs[jid].out_len = reserve - wp;
	if (ub) {   // reference UB: flag the record, do not emulate
		uint32_t const old = atomicOr(&B.flags[rec], PV3_FLAG_TRACE_UB);
		if (!old) atomicAdd(&B.counters->n_flagged, 1ull);
	}
	}   // grid-stride loop over the runs
}

void launch_k3_thread(const K3Params &P, cudaStream_t stream)
{
	// P.n_jobs known (> 0): one pass over the runs; ~0: read on the device, a resident grid strides over them
	if (!P.n_jobs) return;
	unsigned const threads = K3T_THREADS;
	uint64_t const want = P.n_jobs == ~0ull ? (uint64_t) (P.grid_hint + 1) / 2 : (P.n_jobs + threads - 1) / threads;
	unsigned const blocks = (unsigned) (want ? want : 1);
	if (P.k1.seq_len) k3_thread_kernel <true><<<blocks, threads, 0, stream>>>(P);
	else k3_thread_kernel <false><<<blocks, threads, 0, stream>>>(P);
}

} // namespace pv3
