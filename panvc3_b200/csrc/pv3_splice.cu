// panvc3_b200 — kernel K4 (splice realigned runs + collapse) and the position-only projection kernels (sm_100a).
//
//   K4          libpanvc3/alignment_projector.cc:57-151 (copy untouched ops, append realigned ops, copy tail) followed by
//               collapse_cigar_operations (libpanvc3/cigar.cc:13-51)
//   positions   msa_index::sequence_entry::project_position (include/panvc3/msa_index.hh:253-258), used for mates in
//               project-alignments/project_alignments.cc:880-906
#include "pv3_device.cuh"
#include "pv3_kernels.h"

namespace pv3 {

// One thread per run; the thread holding the first run of a record rebuilds that record's CIGAR. The arena space of a
// warp's records is taken with one atomic (per-record atomics on the cursor serialise when most records realign).
__global__ void __launch_bounds__(128)
k4_splice_kernel(K1Params P, uint64_t n_jobs_host, const uint32_t *dp_arena)
{
	unsigned const lane = lane_id();
	uint64_t const n_jobs = dp_n_jobs(P.counters, n_jobs_host);
	uint64_t const stride = (uint64_t) gridDim.x * blockDim.x;
	for (uint64_t jbase = ((uint64_t) blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; jbase < n_jobs; jbase += stride) {
	uint64_t const jid = jbase + lane;
	uint64_t rec = 0;
	uint32_t n_runs = 0, n_pre = 0, bound = 0;
	const uint32_t *pre = nullptr;
	bool owner = false;
	if (jid < n_jobs) {
		rec = P.jobs[jid].rec;
		owner = (P.rr_off[rec] == jid + P.rr_base);
	}
	if (owner) {
		n_runs = P.rr_len[rec];
		if (P.flags[rec]) { P.cig_out_len[rec] = 0; P.rr_len[rec] = 0; owner = false; }   // flagged by a DP kernel
	}
	if (owner) {
		pre = P.arena + (P.cig_out_off[rec] - P.arena_off);
		n_pre = P.cig_out_len[rec];
		bound = n_pre;
		for (uint32_t r = 0; r < n_runs; ++r) {
			Job const &j = P.jobs[jid + r];
			bound += j.out_len;
			bound -= (j.op_end - j.op_begin);
		}
	}
	uint32_t incl = bound;
	#pragma unroll
	for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
	uint32_t const total = __shfl_sync(FULL, incl, 31);
	unsigned long long wbase = 0;
	if (lane == 0 && total) {
		wbase = atomicAdd(&P.counters->arena_cursor, (unsigned long long) total);
		if (wbase + total > P.arena_cap) { atomicOr(&P.counters->overflow, 1u); wbase = ~0ull; }
	}
	wbase = shfl_u64(wbase, 0);
	if (!owner) continue;
	if (wbase == ~0ull) { P.cig_out_len[rec] = 0; continue; }
	unsigned long long const base = wbase + (incl - bound);
	uint32_t *dst = P.arena + base;
	// collapse_cigar_operations (cigar.cc:33-38): the open op stays in registers until a different op arrives
	uint32_t w = 0, cur_op = OP_NONE, cur_cnt = 0;
	auto push = [&](uint32_t v) {
		uint32_t const o = v & 0xfu;
		if (o == cur_op) { cur_cnt += v >> 4; return; }
		if (cur_op != OP_NONE) dst[w++] = (cur_cnt << 4) | cur_op;
		cur_op = o; cur_cnt = v >> 4;
	};
	uint32_t k = 0;
	for (uint32_t r = 0; r < n_runs; ++r) {
		Job const &j = P.jobs[jid + r];
		for (; k < j.op_begin; ++k) push(pre[k]);                       // alignment_projector.cc:61-66
		const uint32_t *ra = dp_arena + j.out_off;
		for (uint32_t t = 0; t < j.out_len; ++t) push(ra[t]);           // :140-141
		k = j.op_end;
	}
	for (; k < n_pre; ++k) push(pre[k]);                                // :145-148
	if (cur_op != OP_NONE) dst[w++] = (cur_cnt << 4) | cur_op;
	P.cig_out_off[rec] = base + P.arena_off;
	P.cig_out_len[rec] = w;
	}   // grid-stride loop over the runs
}

void launch_k4(const K1Params &P, uint64_t n_jobs, const uint32_t *dp_arena, cudaStream_t stream)
{
	if (!n_jobs) return;
	unsigned const threads = 128;
	// n_jobs == ~0: the count is read on the device and a resident grid strides over the runs
	unsigned const blocks = n_jobs == ~0ull ? 148u * 8u : (unsigned) ((n_jobs + threads - 1) / threads);
	k4_splice_kernel<<<blocks, threads, 0, stream>>>(P, n_jobs, dp_arena);
}

__global__ void positions_kernel(const RowDesc *rows, uint32_t n_rows, uint64_t n, const uint32_t *src_row,
	const uint32_t *dst_row, const uint64_t *pos, uint64_t *out)
{
	uint64_t const i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t const s = src_row[i], d = dst_row[i];
	if (s >= n_rows || d >= n_rows) { out[i] = ~0ull; return; }
	RowDesc const S = rows[s], D = rows[d];
	out[i] = rank0_thread(D, select0_thread(S, pos[i] + 1));
}

void launch_positions(const RowDesc *rows, uint32_t n_rows, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row,
	const uint64_t *pos, uint64_t *out, cudaStream_t stream)
{
	if (!n) return;
	unsigned const threads = 256;
	positions_kernel<<<(unsigned) ((n + threads - 1) / threads), threads, 0, stream>>>(rows, n_rows, n, src_row, dst_row, pos, out);
}

// Parity probe. Odd queries go through the warp-cooperative code path (as the rewrite kernel uses it), even ones through
// the single-thread path, so both are checked against the oracle.
__global__ void index_query_kernel(const RowDesc *rows, uint32_t n_rows, int op, uint64_t n, const uint32_t *row,
	const uint64_t *x, uint64_t *out)
{
	uint64_t const warp = ((uint64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	unsigned const lane = lane_id();
	for (uint64_t i = warp * 32; i < warp * 32 + 32 && i < n; ++i) {
		uint32_t const r = row[i];
		if (r >= n_rows) { if (lane == 0) out[i] = ~0ull; continue; }
		RowDesc const R = rows[r];
		uint64_t v;
		if (op == 2) v = (x[i] < R.n_bits) ? ((R.words[x[i] >> 6] >> (x[i] & 63)) & 1ull) : 1ull;
		else if (i & 1) v = (op == 0) ? rank0_warp(R, x[i]) : select0_warp(R, x[i]);
		else v = (op == 0) ? rank0_thread(R, x[i]) : select0_thread(R, x[i]);
		if (lane == 0) out[i] = v;
	}
}

void launch_index_query(const RowDesc *rows, uint32_t n_rows, int op, uint64_t n, const uint32_t *row, const uint64_t *x,
	uint64_t *out, cudaStream_t stream)
{
	if (!n) return;
	uint64_t const warps = (n + 31) / 32;
	unsigned const threads = 128;
	unsigned const blocks = (unsigned) ((warps * 32 + threads - 1) / threads);
	index_query_kernel<<<blocks, threads, 0, stream>>>(rows, n_rows, op, n, row, x, out);
}

} // namespace pv3
