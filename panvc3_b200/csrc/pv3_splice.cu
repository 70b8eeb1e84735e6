// panvc3_b200 — kernel K4 (splice realigned runs + collapse) and the position-only projection kernels (sm_100a).
//
//   K4          libpanvc3/alignment_projector.cc:57-151 (copy untouched ops, append realigned ops, copy tail) followed by
//               collapse_cigar_operations (libpanvc3/cigar.cc:13-51)
//   positions   msa_index::sequence_entry::project_position (include/panvc3/msa_index.hh:253-258), used for mates in
//               project-alignments/project_alignments.cc:880-906
#include "pv3_device.cuh"
#include "pv3_kernels.h"

namespace pv3 {

// One thread per run; the thread holding the first run of a record rebuilds that record's CIGAR.
__global__ void k4_splice_kernel(K1Params P, uint64_t n_jobs, const uint32_t *dp_arena)
{
	uint64_t const jid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (jid >= n_jobs) return;
	uint64_t const rec = P.jobs[jid].rec;
	if (P.rr_off[rec] != jid + P.rr_base) return;
	uint32_t const n_runs = P.rr_len[rec];
	if (P.flags[rec]) { P.cig_out_len[rec] = 0; P.rr_len[rec] = 0; return; }   // flagged by the DP kernel

	const uint32_t *pre = P.arena + (P.cig_out_off[rec] - P.arena_off);
	uint32_t const n_pre = P.cig_out_len[rec];
	unsigned long long bound = n_pre;
	for (uint32_t r = 0; r < n_runs; ++r) {
		Job const &j = P.jobs[jid + r];
		bound += j.out_len;
		bound -= (j.op_end - j.op_begin);
	}
	unsigned long long const base = atomicAdd(&P.counters->arena_cursor, bound);
	if (base + bound > P.arena_cap) { atomicOr(&P.counters->overflow, 1u); P.cig_out_len[rec] = 0; return; }
	uint32_t *dst = P.arena + base;
	long long w = -1;
	auto push = [&](uint32_t v) {
		if (w >= 0 && (dst[w] & 0xfu) == (v & 0xfu)) dst[w] += (v >> 4) << 4;   // cigar.cc:33-38
		else dst[++w] = v;
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
	P.cig_out_off[rec] = base + P.arena_off;
	P.cig_out_len[rec] = (uint32_t) (w + 1);
}

void launch_k4(const K1Params &P, uint64_t n_jobs, const uint32_t *dp_arena, cudaStream_t stream)
{
	if (!n_jobs) return;
	unsigned const threads = 128;
	unsigned const blocks = (unsigned) ((n_jobs + threads - 1) / threads);
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
