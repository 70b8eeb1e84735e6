// panvc3_b200 — the two ends of the host pipeline on the device (sm_100a): unpack_in turns the COMPACT input encoding
// (lengths instead of offsets, 16-bit ids, 32-bit positions: what crosses PCIe) into the arrays the kernels read;
// pack_out turns the kernels' arena-addressed results into the ORDERED output (arenas in record order, so that offsets need
// not travel) and writes it where the caller wants it — device memory or, zero-copy, the caller's pinned host buffers.
//
// Reference contract being served: records leave in input order (project-alignments/project_alignments.cc:958-970); the
// per-record fields are the arguments / results of alignment_projector::project_alignment
// (include/panvc3/alignment_projector.hh:58-72). HBM-bound streaming work: three passes (tile sums, one-CTA scan, write),
// 16-byte or lane-consecutive accesses.
#include "pv3_device.cuh"
#include "pv3_kernels.h"
#include <algorithm>

namespace pv3 {

namespace {

constexpr int PK_THREADS = 256;
constexpr int PK_PER = 4;                           // records per thread
constexpr int PK_TILE = PK_THREADS * PK_PER;        // records per CTA

__device__ __forceinline__ uint64_t warp_incl_scan(uint64_t v)
{
	#pragma unroll
	for (int d = 1; d < 32; d <<= 1) { uint64_t const t = shfl_u64(v, (int) lane_id() - d); if ((int) lane_id() >= d) v += t; }
	return v;
}

// exclusive scan of one value per thread over the CTA; returns the exclusive prefix, total in `total`
__device__ uint64_t block_excl_scan(uint64_t v, uint64_t *smem /* 9 words */, uint64_t &total)
{
	unsigned const lane = lane_id(), warp = threadIdx.x >> 5;
	uint64_t const incl = warp_incl_scan(v);
	__syncthreads();
	if (lane == 31) smem[warp] = incl;
	__syncthreads();
	if (warp == 0) {
		uint64_t x = lane < PK_THREADS / 32 ? smem[lane] : 0;
		uint64_t const xi = warp_incl_scan(x);
		if (lane < PK_THREADS / 32) smem[lane] = xi - x;
		if (lane == PK_THREADS / 32 - 1) smem[8] = xi;
	}
	__syncthreads();
	total = smem[8];
	return smem[warp] + incl - v;
}

// ---- input side ---------------------------------------------------------------------------------------------------

__device__ __forceinline__ void in_lengths(const UnpackParams &P, uint64_t i, uint64_t &seq_b, uint64_t &cig_n, uint64_t &qual_n)
{
	uint32_t const bases = P.seq_len[i];
	seq_b = P.seq_packed == 2 ? 2ull * ((bases + 3u) / 4u) : P.seq_packed ? (uint64_t) ((bases + 1u) / 2u) : (uint64_t) bases;
	cig_n = P.cig_len16[i];
	qual_n = (P.has_qual && P.has_qual[i >> 3] >> (i & 7) & 1u) ? bases : 0u;
}

__global__ void __launch_bounds__(PK_THREADS) unpack_sums_kernel(UnpackParams P)
{
	__shared__ uint64_t sm[3][9];
	uint64_t a = 0, b = 0, c = 0;
	uint64_t const i0 = (uint64_t) blockIdx.x * PK_TILE + (uint64_t) threadIdx.x * PK_PER;
	#pragma unroll
	for (int k = 0; k < PK_PER; ++k) if (i0 + k < P.n) { uint64_t x, y, z; in_lengths(P, i0 + k, x, y, z); a += x; b += y; c += z; }
	uint64_t ta, tb, tc;
	block_excl_scan(a, sm[0], ta); block_excl_scan(b, sm[1], tb); block_excl_scan(c, sm[2], tc);
	if (threadIdx.x == 0) { P.tile_sums[3 * blockIdx.x] = ta; P.tile_sums[3 * blockIdx.x + 1] = tb; P.tile_sums[3 * blockIdx.x + 2] = tc; }
}

// one CTA: exclusive scan of the tile sums, in place (three interleaved series)
__global__ void __launch_bounds__(PK_THREADS) scan3_kernel(uint64_t *sums, uint64_t n_tiles, uint64_t *totals /* [3], may be null */)
{
	__shared__ uint64_t sm[3][9];
	uint64_t carry[3] = { 0, 0, 0 };
	for (uint64_t t0 = 0; t0 < n_tiles; t0 += PK_THREADS) {
		uint64_t const t = t0 + threadIdx.x;
		#pragma unroll
		for (int s = 0; s < 3; ++s) {
			uint64_t const v = t < n_tiles ? sums[3 * t + s] : 0;
			uint64_t tot;
			uint64_t const ex = block_excl_scan(v, sm[s], tot);
			if (t < n_tiles) sums[3 * t + s] = carry[s] + ex;
			carry[s] += tot;
		}
	}
	if (totals && threadIdx.x == 0) { totals[0] = carry[0]; totals[1] = carry[1]; totals[2] = carry[2]; }
}

__global__ void __launch_bounds__(PK_THREADS) unpack_write_kernel(UnpackParams P)
{
	__shared__ uint64_t sm[3][9];
	uint64_t const i0 = (uint64_t) blockIdx.x * PK_TILE + (uint64_t) threadIdx.x * PK_PER;
	uint64_t la[PK_PER], lb[PK_PER], lc[PK_PER], a = 0, b = 0, c = 0;
	#pragma unroll
	for (int k = 0; k < PK_PER; ++k) {
		la[k] = lb[k] = lc[k] = 0;
		if (i0 + k < P.n) in_lengths(P, i0 + k, la[k], lb[k], lc[k]);
		a += la[k]; b += lb[k]; c += lc[k];
	}
	uint64_t ta, tb, tc;
	uint64_t ea = block_excl_scan(a, sm[0], ta) + P.tile_sums[3 * blockIdx.x];
	uint64_t eb = block_excl_scan(b, sm[1], tb) + P.tile_sums[3 * blockIdx.x + 1];
	uint64_t ec = block_excl_scan(c, sm[2], tc) + P.tile_sums[3 * blockIdx.x + 2];
	#pragma unroll
	for (int k = 0; k < PK_PER; ++k) {
		uint64_t const i = i0 + k;
		if (i < P.n) {
			P.seq_off[i] = ea; P.cig_off[i] = eb;
			if (P.qual_off) P.qual_off[i] = ec;
			P.src_row[i] = P.src_row16[i]; P.dst_row[i] = P.dst_row16[i]; P.dst_ref[i] = P.dst_ref16[i];
			P.src_pos[i] = P.src_pos32[i];
			if (P.mate_row16) { P.mate_row[i] = P.mate_row16[i]; P.mate_pos[i] = P.mate_pos32[i]; }
			ea += la[k]; eb += lb[k]; ec += lc[k];
			if (i + 1 == P.n) { P.seq_off[P.n] = ea; P.cig_off[P.n] = eb; if (P.qual_off) P.qual_off[P.n] = ec; }
		}
	}
}

// ---- output side --------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(PK_THREADS) pack_sums_kernel(PackParams P)
{
	__shared__ uint64_t sm[2][9];
	uint64_t a = 0, b = 0;
	uint64_t const i0 = (uint64_t) blockIdx.x * PK_TILE + (uint64_t) threadIdx.x * PK_PER;
	#pragma unroll
	for (int k = 0; k < PK_PER; ++k) if (i0 + k < P.n) { a += P.cig_len[i0 + k]; b += P.rr_len[i0 + k]; }
	uint64_t ta, tb;
	block_excl_scan(a, sm[0], ta); block_excl_scan(b, sm[1], tb);
	if (threadIdx.x == 0) { P.tile_sums[3 * blockIdx.x] = ta; P.tile_sums[3 * blockIdx.x + 1] = tb; P.tile_sums[3 * blockIdx.x + 2] = 0; }
}

// One CTA after the scan of the tile sums: takes this chunk's place in the batch-wide output (running totals of the
// chunks before it, kept on the device so that chunks chain without the host), checks the caller's capacities and
// publishes the chunk's counters (to mapped pinned memory in the host path).
__global__ void pack_place_kernel(PackParams P, const uint64_t *totals /* [3] */)
{
	if (threadIdx.x || blockIdx.x) return;
	Counters const c = *P.counters;
	uint64_t const cig_base = P.running[0], rr_base = P.running[1];
	uint64_t const cig_n = totals[0], rr_n = totals[1];
	unsigned ovf = c.overflow;
	if (cig_base + cig_n > P.out_cig_cap) ovf |= 16u;
	if (rr_base + rr_n > P.out_rr_cap) ovf |= 32u;
	uint64_t const pad = P.staged ? ((P.cig_dst_word0 + cig_base) & 3ull) : 0ull;
	P.place[0] = cig_base; P.place[1] = rr_base; P.place[2] = ovf; P.place[3] = cig_n; P.place[4] = rr_n; P.place[5] = pad;
	P.running[0] = cig_base + cig_n; P.running[1] = rr_base + rr_n;
	ChunkReport r;
	r.cig_used = cig_n; r.rr_used = rr_n; r.n_flagged = c.n_flagged; r.n_realigned = c.n_realigned; r.dp_cells = c.dp_cells;
	r.dp_path = c.dp_path; r.dp_max_pool = c.dp_max_pool; r.arena_need = c.arena_cursor + c.dp_path; r.jobs = c.job_cursor;
	r.n_deferred = c.n_deferred; r.cig_base = cig_base; r.rr_base = rr_base; r.cig_pad = pad; r.seg_need = c.seg_cursor; r.seg_ops_need = c.seg_ops_cursor; r.overflow = ovf; r.done = 1;
	*P.report = r;
}

__global__ void __launch_bounds__(PK_THREADS) pack_write_kernel(PackParams P)
{
	__shared__ uint64_t sm[2][9];
	if (P.place[2] & (1u | 2u | 4u | 8u | 16u | 32u | 128u | 256u)) return;       // a capacity was exceeded: the host grows and re-runs
	uint64_t const i0 = (uint64_t) blockIdx.x * PK_TILE + (uint64_t) threadIdx.x * PK_PER;
	uint32_t la[PK_PER], lb[PK_PER];
	uint64_t a = 0, b = 0;
	#pragma unroll
	for (int k = 0; k < PK_PER; ++k) {
		la[k] = lb[k] = 0;
		if (i0 + k < P.n) { la[k] = P.cig_len[i0 + k]; lb[k] = P.rr_len[i0 + k]; }
		a += la[k]; b += lb[k];
	}
	uint64_t ta, tb;
	uint64_t ea = block_excl_scan(a, sm[0], ta) + P.tile_sums[3 * blockIdx.x] + P.place[0];
	uint64_t eb = block_excl_scan(b, sm[1], tb) + P.tile_sums[3 * blockIdx.x + 1] + P.place[1];
	uint64_t const wa = P.staged ? P.place[0] - P.place[5] : 0, wb = P.staged ? P.place[1] : 0;   // staging buffers start at the chunk's place (+ pad)
	#pragma unroll
	for (int k = 0; k < PK_PER; ++k) {
		uint64_t const i = i0 + k;
		if (i >= P.n) break;
		uint64_t const o = P.rec0 + i;                                   // record index in the caller's batch
		uint32_t const fl = P.flags[i];
		if (P.out_dst_pos) P.out_dst_pos[o] = P.dst_pos[i];
		if (P.out_dst_pos32) P.out_dst_pos32[o] = (uint32_t) P.dst_pos[i];
		if (P.out_flags) P.out_flags[o] = fl;
		if (P.out_flags8) P.out_flags8[o] = (uint8_t) fl;
		if (P.out_cig_off) P.out_cig_off[o] = ea + P.cig_off_bias;
		if (P.out_cig_len) P.out_cig_len[o] = la[k];
		if (P.out_cig_len16) P.out_cig_len16[o] = (uint16_t) (la[k] > 0xffffu ? 0xffffu : la[k]);
		if (P.out_rr_off) P.out_rr_off[o] = eb + P.rr_off_bias;
		if (P.out_rr_len) P.out_rr_len[o] = lb[k];
		if (P.out_rr_len8) P.out_rr_len8[o] = (uint8_t) (lb[k] > 0xffu ? 0xffu : lb[k]);
		if (P.mate_dst_pos && P.out_mate_dst_pos) P.out_mate_dst_pos[o] = P.mate_dst_pos[i];
		if (P.mate_dst_pos && P.out_mate_dst_pos32) P.out_mate_dst_pos32[o] = (uint32_t) P.mate_dst_pos[i];
		if ((la[k] > 0xffffu && P.out_cig_len16) || (lb[k] > 0xffu && P.out_rr_len8) || (fl > 0xffu && P.out_flags8)
			|| ((P.dst_pos[i] >> 32) && P.out_dst_pos32))
			atomicOr(&P.report->overflow, 64u);                         // a value does not fit the compact field the caller chose
		const uint32_t *src = P.arena + (P.cig_off[i] - P.arena_off);
		for (uint32_t j = 0; j < la[k]; ++j) P.out_cig[ea - wa + j] = src[j];
		const uint64_t *rs = P.rr + 4 * (P.rr_off[i] - P.rr_base);
		for (uint32_t j = 0; j < 4 * lb[k]; ++j) P.out_rr[4 * (eb - wb) + j] = rs[j];
		ea += la[k]; eb += lb[k];
	}
}

// bytes from src to dst where (dst - src) is a multiple of 16: byte head, 16-byte body, byte tail; any other pair: bytes
__device__ void stream_copy(uint8_t *dst, const uint8_t *src, uint64_t bytes, uint64_t tid, uint64_t n_threads)
{
	if (!bytes) return;
	if (((uintptr_t) dst - (uintptr_t) src) & 15u) { for (uint64_t i = tid; i < bytes; i += n_threads) dst[i] = src[i]; return; }
	uint64_t head = (16u - ((uintptr_t) dst & 15u)) & 15u;
	if (head > bytes) head = bytes;
	uint64_t const body = (bytes - head) / 16, tail0 = head + body * 16;
	if (tid < head) dst[tid] = src[tid];
	const uint4 *s4 = reinterpret_cast <const uint4 *>(src + head);
	uint4 *d4 = reinterpret_cast <uint4 *>(dst + head);
	for (uint64_t i = tid; i < body; i += n_threads) d4[i] = s4[i];
	if (tid < bytes - tail0) dst[tail0 + tid] = src[tail0 + tid];
}

__global__ void __launch_bounds__(256) stream_out_kernel(StreamOutParams P)
{
	if (P.place[2] & (1u | 2u | 4u | 8u | 16u | 32u | 128u | 256u)) return;
	uint64_t const tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x, nt = (uint64_t) gridDim.x * blockDim.x;
	for (int s = 0; s < P.n_seg; ++s) stream_copy((uint8_t *) P.seg[s].dst, (const uint8_t *) P.seg[s].src, P.seg[s].bytes, tid, nt);
	if (P.cig_dst) stream_copy((uint8_t *) (P.cig_dst + P.place[0]), (const uint8_t *) (P.cig_src + P.place[5]), P.place[3] * 4, tid, nt);
	if (P.rr_dst) stream_copy((uint8_t *) (P.rr_dst + 4 * P.place[1]), (const uint8_t *) P.rr_src, P.place[4] * 32, tid, nt);
}

} // namespace

void launch_stream_out(const StreamOutParams &P, cudaStream_t stream)
{
	stream_out_kernel<<<96, 256, 0, stream>>>(P);
}

uint64_t pack_tiles(uint64_t n) { return (n + PK_TILE - 1) / PK_TILE; }

// Two-bit SEQ → the 4-bit codes of a BAM record (A = 1, C = 2, G = 4, T = 8: 1 << code), first base in the high nibble.
// A thread takes 4 input bytes (16 bases) and writes 8 output bytes; a record of ceil(L / 4) input bytes becomes
// 2 ceil(L / 4) output bytes, so output offsets are twice the input offsets and no per-record work is needed.
__global__ void __launch_bounds__(256) expand_seq2_kernel(const uint8_t *seq2, uint64_t n_bytes2, uint8_t *seq4)
{
	uint64_t const n_words = (n_bytes2 + 3) / 4;
	for (uint64_t w = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; w < n_words; w += (uint64_t) gridDim.x * blockDim.x) {
		uint32_t const x = reinterpret_cast <const uint32_t *>(seq2)[w];       // (buffers are padded to 64 bytes)
		uint32_t o[2];
		#pragma unroll
		for (int h = 0; h < 2; ++h) {
			uint32_t r = 0;
			#pragma unroll
			for (int b = 0; b < 2; ++b) {                                        // input byte 2 h + b → output bytes 2 b, 2 b + 1 of word h
				uint32_t const v = (x >> (16 * h + 8 * b)) & 0xffu;
				uint32_t const c0 = 1u << (v & 3u), c1 = 1u << ((v >> 2) & 3u), c2 = 1u << ((v >> 4) & 3u), c3 = 1u << ((v >> 6) & 3u);
				r |= (((c0 << 4) | c1) | (((c2 << 4) | c3) << 8)) << (16 * b);
			}
			o[h] = r;
		}
		reinterpret_cast <uint2 *>(seq4)[w] = make_uint2(o[0], o[1]);
	}
}

// base exc_off[j] of record exc_rec[j] is N (BAM code 15): OR the nibble with 0xf (atomic on the aligned word: neighbours may be hit at once)
__global__ void __launch_bounds__(256) seq_exceptions_kernel(uint8_t *seq4, const uint32_t *exc_rec, const uint32_t *exc_off, uint64_t n_exc,
	uint64_t rec_base, const uint64_t *seq_off)
{
	uint64_t const j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= n_exc) return;
	uint64_t const r = (uint64_t) exc_rec[j] - rec_base;
	uint32_t const o = exc_off[j];
	uint64_t const byte = seq_off[r] + (o >> 1);
	uint32_t const shift = 8u * (uint32_t) (byte & 3u) + ((o & 1u) ? 0u : 4u);
	atomicOr(reinterpret_cast <unsigned int *>(seq4 + (byte & ~3ull)), 0xfu << shift);
}

void launch_expand_seq2(const uint8_t *seq2, uint64_t n_bytes2, uint8_t *seq4, const uint32_t *exc_rec, const uint32_t *exc_off, uint64_t n_exc,
	uint64_t rec_base, const uint64_t *seq_off, cudaStream_t stream)
{
	if (n_bytes2) {
		uint64_t const n_words = (n_bytes2 + 3) / 4;
		unsigned const blocks = (unsigned) std::min <uint64_t>((n_words + 255) / 256, 148u * 16u);
		expand_seq2_kernel<<<blocks, 256, 0, stream>>>(seq2, n_bytes2, seq4);
	}
	if (n_exc) seq_exceptions_kernel<<<(unsigned) ((n_exc + 255) / 256), 256, 0, stream>>>(seq4, exc_rec, exc_off, n_exc, rec_base, seq_off);
}

void launch_unpack_in(const UnpackParams &P, cudaStream_t stream)
{
	if (!P.n) return;
	unsigned const tiles = (unsigned) pack_tiles(P.n);
	unpack_sums_kernel<<<tiles, PK_THREADS, 0, stream>>>(P);
	scan3_kernel<<<1, PK_THREADS, 0, stream>>>(P.tile_sums, tiles, nullptr);
	unpack_write_kernel<<<tiles, PK_THREADS, 0, stream>>>(P);
}

void launch_pack_out(const PackParams &P, cudaStream_t stream)
{
	unsigned const tiles = (unsigned) pack_tiles(P.n ? P.n : 1);
	pack_sums_kernel<<<tiles, PK_THREADS, 0, stream>>>(P);
	scan3_kernel<<<1, PK_THREADS, 0, stream>>>(P.tile_sums, tiles, P.tile_sums + 3 * (uint64_t) tiles);
	pack_place_kernel<<<1, 32, 0, stream>>>(P, P.tile_sums + 3 * (uint64_t) tiles);
	pack_write_kernel<<<tiles, PK_THREADS, 0, stream>>>(P);
}

} // namespace pv3
