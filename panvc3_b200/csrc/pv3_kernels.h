// panvc3_b200 — kernel parameter blocks and host-side launchers (internal; the public surface is include/panvc3_b200.h).
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/panvc3_b200.h"
#include "pv3_device.cuh"


namespace pv3 {

struct K1Params {
	uint64_t        n;
	const uint32_t *src_row, *dst_row, *dst_ref;
	const uint64_t *src_pos, *seq_off;
	const uint8_t  *seq;
	const uint32_t *seq_len;     // null: seq holds one dna5 rank per byte; else seq is BAM 4-bit packed, seq_off counts bytes, seq_len bases
	const uint64_t *qual_off;    // may be null
	const uint8_t  *qual;
	const uint64_t *cig_off;
	const uint32_t *cig;
	const RowDesc  *rows; uint32_t n_rows;
	const RefDesc  *refs; uint32_t n_refs;
	uint64_t       *dst_pos;
	uint32_t       *flags;
	uint64_t       *cig_out_off;
	uint32_t       *cig_out_len;
	uint32_t       *arena; uint64_t arena_cap;
	uint64_t       *rr_off;
	uint32_t       *rr_len;
	uint64_t       *rr; uint64_t rr_cap;
	Job            *jobs;
	Counters       *counters;
	int32_t         gap_open, gap_ext;
	int             stage;       // 0 = full pipeline, 1 = rewrite only (parity probe of rewrite_cigar())
	const uint64_t *rec_list;    // warp kernel: process records rec_list[0..n_list) instead of 0..n (null ⇒ all)
	uint64_t        n_list;
	const unsigned long long *n_list_dev;   // warp kernel: when set, the list length is read on the device (no host round trip)
	uint64_t       *deferred;    // thread kernel: records it could not handle (capacity n)
	uint64_t       *complex_list;// word kernel: records with events, handled by its second launch (capacity n)
	uint64_t       *complex_ce;  // word kernel: start column and projected position of those records (2 per entry)
	// long records as segments (pv3_segment_core.h): plan → segment kernel → assembly; what fails there goes to deferred2
	uint64_t       *plan_rec, *plan_seg0; uint32_t *plan_nseg;     // per planned record: record, first segment, segment count (capacity n)
	void           *segs;        // seg::SegDesc[seg_cap]
	void           *seg_out;     // seg::SegOut[seg_cap]
	uint64_t        seg_cap;
	uint32_t       *seg_ops; uint64_t seg_ops_cap;                 // segment op arena
	uint64_t       *deferred2;   // capacity n
	uint64_t        arena_off;   // added to every stored cig_off (chunked host pipeline: arena of chunk c starts here)
	uint64_t        rr_base;     // added to every stored rr_off
	// thread kernel only: mate positions projected in its phase A (the mate sits a few hundred columns from the read: the
	// directory and word sectors are the ones just fetched; a separate pass misses DRAM for all of them again at chromosome scale)
	const uint32_t *mate_row;    // null: no mates here
	const uint64_t *mate_pos;
	uint64_t       *mate_dst_pos;
};

// BAM 4-bit base code ("=ACMGRSVTWYHKDBN") → seqan3::dna5 rank, as seqan3 converts when it reads a BAM SEQ into dna5
// (A C G T keep their letter, everything else becomes N): a 16 x 4-bit table.
__host__ __device__ inline uint32_t bam4_to_rank(uint32_t code) { return (uint32_t) (0x3333333433323103ull >> (4 * (code & 15u))) & 15u; }

#ifdef __CUDACC__
// rank of base i of the record whose SEQ starts at seq_off value qbase
__device__ __forceinline__ uint8_t query_rank(K1Params const &P, uint64_t qbase, uint64_t i)
{
	if (!P.seq_len) return __ldg(P.seq + qbase + i);
	uint32_t const b = __ldg(P.seq + qbase + (i >> 1));
	return (uint8_t) bam4_to_rank((i & 1u) ? b : (b >> 4));
}
__device__ __forceinline__ uint64_t query_length(K1Params const &P, uint64_t rec)
{
	return P.seq_len ? (uint64_t) __ldg(P.seq_len + rec) : __ldg(P.seq_off + rec + 1) - __ldg(P.seq_off + rec);
}
#endif

// Scratch of one DP job of the warp-per-run kernel. Trace: per strip of 32 query rows, R + 31 steps rounded up to blocks of 32 steps, in groups of four steps, 32 lanes
// per group, a 32-bit word (affine: one byte per cell) or a byte (edit distance: two bits per cell) per lane and group. Strip
// boundary arrays (H, vertical gap state, its trace bit: 4 bytes each per column) only when there is more than one strip.
__host__ __device__ inline uint32_t dp_trace_groups(uint32_t R) { return ((R + 62u) >> 5) << 3; }
__host__ __device__ inline unsigned long long dp_trace_bytes(uint32_t R, uint32_t Q, bool affine)
{
	unsigned long long const strips = ((unsigned long long) Q + 31) >> 5;
	return strips * dp_trace_groups(R) * 32ull * (affine ? 4ull : 1ull);
}
__host__ __device__ inline unsigned long long dp_pool_bytes(uint32_t R, uint32_t Q, bool affine)
{
	return dp_trace_bytes(R, Q, affine) + (Q > 32u ? 12ull * ((unsigned long long) R + 1) : 0ull);
}

// Runs small enough for the thread-per-run DP kernel (K3T): the whole DP state of one run fits a thread's local arrays.
// The warp-per-run kernel (K3) skips them, K3T skips the others; counters.n_big_jobs tells the host whether K3 must run.
constexpr uint32_t K3T_MAX_DIM   = 15;
constexpr uint32_t K3T_MAX_TRACE = 64;           // (R + 1) * (Q + 1) trace bytes: a lane waits for the largest run of its warp
__host__ __device__ inline bool dp_job_is_small(uint32_t R, uint32_t Q)
{
	return R <= K3T_MAX_DIM && Q <= K3T_MAX_DIM && (R + 1) * (Q + 1) <= K3T_MAX_TRACE;
}

#ifdef __CUDACC__
// Run count of a batch: given by the host, or (~0) read from the counters the rewrite kernels filled on the same stream.
__device__ __forceinline__ uint64_t dp_n_jobs(const Counters *c, uint64_t host_value)
{
	if (host_value != ~0ull) return host_value;
	// an arena or the run list overflowed in the rewrite: runs beyond the list's capacity were not written — nothing to do
	// here, the host sizes the buffers from the counters and repeats the batch
	if (*((volatile const unsigned int *) &c->overflow) & 3u) return 0;
	return (uint64_t) *((volatile const unsigned long long *) &c->job_cursor);
}
// Small runs go to the thread-per-run kernel once a batch holds at least dp_thread_min of them; both DP kernels evaluate this.
__device__ __forceinline__ bool dp_use_thread_kernel(const Counters *c, uint64_t n_jobs, uint64_t dp_thread_min)
{
	uint64_t const n_big = *((volatile const unsigned long long *) &c->n_big_jobs);
	return n_jobs > n_big && n_jobs - n_big >= dp_thread_min;
}
#endif

struct K3Params {
	K1Params        k1;             // inputs, index, flags, counters, jobs
	uint64_t        n_jobs;         // ~0: read counters->job_cursor on the device
	uint64_t        dp_thread_min;  // pv3_set_dp_thread_min
	uint32_t        grid_hint;      // thread-per-run kernel: CTAs of the resident grid when n_jobs is read on the device
	uint64_t        pool_slots;     // HBM trace slots allocated (one per resident warp of the warp-per-run kernel)
	uint32_t       *dp_arena; uint64_t dp_arena_cap;    // realigned ops, R+Q words reserved per job
	uint8_t        *pool; uint64_t pool_slot_bytes;     // global scratch, one slot per resident warp (for jobs too large for smem)
	int32_t         lut[84];        // quality LUT: [2*q] match, [2*q+1] mismatch
	const uint8_t  *packed_qual;    // lazy-QUAL path: phred ranks of the runs' query windows only, run j at job_qual_off[j]
	const uint64_t *job_qual_off;   // null ⇒ read k1.qual
	uint64_t       *big_list;       // ids of the runs left to the warp kernel, written by the thread kernel (capacity: jobs)
};

struct K4Params {
	K1Params        k1;
	K3Params const *unused;
	uint64_t        n_jobs;
	const uint32_t *dp_arena;
	const uint64_t *realigned_recs; // unused (records are found through jobs)
};

// ---- host pipeline ends (pv3_pack.cu) ----------------------------------------------------------------------------------
// Compact input (device copies of what crossed PCIe) → the arrays of K1Params.
struct UnpackParams {
	uint64_t        n;
	const uint16_t *src_row16, *dst_row16, *dst_ref16;
	const uint32_t *src_pos32;
	const uint32_t *seq_len;      // bases per record
	const uint16_t *cig_len16;
	const uint8_t  *has_qual;     // bit i: record i has QUAL (as long as its SEQ); null: none has
	const uint16_t *mate_row16;   // null: no mates
	const uint32_t *mate_pos32;
	int             seq_packed;   // SEQ bytes per record on the device: 0 = bases, 1 = (bases + 1) / 2 (BAM 4-bit), 2 = 2 * ceil(bases / 4) (BAM 4-bit expanded from two-bit input)
	uint32_t       *src_row, *dst_row, *dst_ref, *mate_row;
	uint64_t       *src_pos, *mate_pos, *seq_off, *cig_off, *qual_off;   // qual_off null when has_qual is null
	uint64_t       *tile_sums;    // scratch: 3 * pack_tiles(n) + 3 words
};

// What a chunk reports to the host (mapped pinned memory in the host path): read once, after the call's only synchronisation.
struct ChunkReport {
	unsigned long long cig_used, rr_used, n_flagged, n_realigned, dp_cells, dp_path, dp_max_pool, arena_need, jobs, n_deferred;
	unsigned long long cig_base, rr_base;   // the chunk's place in the call's ordered arenas
	unsigned long long cig_pad;             // staged output: words in front of the chunk's ops in the staging buffer
	unsigned long long seg_need, seg_ops_need;
	unsigned int       overflow;  // bits 0-3 as Counters::overflow (8: HBM trace slots), 16 / 32: caller's cig / rr capacity, 64: compact field
	unsigned int       done;
};

// Arena-addressed results of one chunk → ordered output at the chunk's place in the caller's batch.
struct PackParams {
	uint64_t        n, rec0;      // records of the chunk; index of its first record in the caller's batch
	const uint64_t *dst_pos; const uint32_t *flags;
	const uint64_t *cig_off; const uint32_t *cig_len; const uint32_t *arena; uint64_t arena_off;
	const uint64_t *rr_off;  const uint32_t *rr_len;  const uint64_t *rr;    uint64_t rr_base;
	const uint64_t *mate_dst_pos; // null: no mates
	const Counters *counters;
	uint64_t       *running;      // [2] device: output ops / runs of the chunks before this one (chains the chunks of a call)
	uint64_t       *place;        // [3] device scratch of this chunk: its bases and the overflow word
	uint64_t       *tile_sums;    // scratch: 3 * pack_tiles(n) + 3 words
	ChunkReport    *report;
	// destination (device or mapped host pointers; null = not wanted)
	uint64_t *out_dst_pos; uint32_t *out_dst_pos32; uint32_t *out_flags; uint8_t *out_flags8;
	uint64_t *out_cig_off; uint32_t *out_cig_len; uint16_t *out_cig_len16; uint32_t *out_cig; uint64_t out_cig_cap;
	uint64_t *out_rr_off; uint32_t *out_rr_len; uint8_t *out_rr_len8; uint64_t *out_rr; uint64_t out_rr_cap;
	uint64_t *out_mate_dst_pos; uint32_t *out_mate_dst_pos32;
	uint64_t  cig_off_bias, rr_off_bias;   // added to the stored offsets (pv3_multi: where the shard's arenas start in the caller's arrays)
	int       staged;         // 1: out_cig / out_rr are staging buffers of this chunk (index 0 = the chunk's place)
	uint64_t  cig_dst_word0;  // staged: (address of the destination's word 0 of this context's arena) / 4 — the staging copy keeps the
	                          // destination's alignment modulo 16 bytes so that it can be streamed out with 16-byte stores
};

// Staged ordered output of a chunk → the caller's pinned arrays, with coalesced 16-byte stores (full PCIe write
// transactions; per-record stores straight into mapped memory reached 4 GB/s, this reaches the link's rate).
struct StreamOutParams {
	struct Seg { void *dst; const void *src; uint64_t bytes; } seg[12];
	int             n_seg;
	uint32_t       *cig_dst;  const uint32_t *cig_src;      // dst: word 0 of this context's arena in the caller's array
	uint64_t       *rr_dst;   const uint64_t *rr_src;
	const uint64_t *place;    // [0] cig base, [1] rr base, [2] overflow, [3] cig words, [4] runs, [5] cig pad
};
void launch_stream_out(const StreamOutParams &P, cudaStream_t stream);
uint64_t pack_tiles(uint64_t n);
// two-bit SEQ (pv3_batch_in.seq_2bit) → BAM 4-bit layout, records padded to four bases; then the N exceptions
void launch_expand_seq2(const uint8_t *seq2, uint64_t n_bytes2, uint8_t *seq4, const uint32_t *exc_rec, const uint32_t *exc_off, uint64_t n_exc,
	uint64_t rec_base, const uint64_t *seq_off, cudaStream_t stream);
void launch_unpack_in(const UnpackParams &P, cudaStream_t stream);
void launch_pack_out(const PackParams &P, cudaStream_t stream);

void launch_k1(const K1Params &P, int n_sms, cudaStream_t stream);          // warp per record (general)
void launch_k1_thread(const K1Params &P, int n_sms, uint64_t target_code_bytes, cudaStream_t stream, int layout = -1);
bool k1_thread_layout_is_tunable(uint64_t target_code_bytes);                 // large target and no PV3_K1T_VARIANT override
   // thread per record, run walk (v4-v13)
void launch_k1_word(const K1Params &P, int n_sms, cudaStream_t stream);
// the deferred list of either thread kernel cut into segments of <= 64 reference bases, one thread per segment (long reads)
// `expected`: the deferred records the host expects (last batch's count, read lengths): sizes the resident grids, which read
// the real counts on the device — a wrong guess costs time, never results
void launch_k1_segments(const K1Params &P, int n_sms, uint64_t expected, cudaStream_t stream);     // plan + segment kernels (pv3_rewrite_seg.cu)
void launch_k1_assemble(const K1Params &P, int n_sms, uint64_t expected, cudaStream_t stream);     // concatenation, run detection, outputs (pv3_rewrite.cu)
void launch_k1_list(const K1Params &P, int n_sms, uint64_t expected, cudaStream_t stream);         // warp-per-record kernel over a device-counted list     // thread per record, word-parallel (v14; default)
void launch_k3(const K3Params &P, int n_sms, uint64_t grid_warps, cudaStream_t stream);
void launch_k3_thread(const K3Params &P, cudaStream_t stream);               // thread per run (small runs)
void launch_k4(const K1Params &P, uint64_t n_jobs, const uint32_t *dp_arena, cudaStream_t stream);
void launch_positions(const RowDesc *rows, uint32_t n_rows, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row,
	const uint64_t *pos, uint64_t *out, cudaStream_t stream);
void launch_index_query(const RowDesc *rows, uint32_t n_rows, int op, uint64_t n, const uint32_t *row, const uint64_t *x,
	uint64_t *out, cudaStream_t stream);
// index-msa on the device (pv3_msa_build.cu): scratch of msa_extract_scratch_words(n) words; word msa_extract_total_word(n)
// receives the non-gap count
uint64_t msa_extract_tiles(uint64_t n);
uint64_t msa_extract_scratch_words(uint64_t n);
uint64_t msa_extract_total_word(uint64_t n);
cudaError_t launch_msa_extract(const char *text, uint64_t n, uint64_t *gap_words, char *ungapped, unsigned long long *scratch,
	cudaStream_t stream, unsigned *launches);
// recalculate-mapq (pv3_mapq.cu)
struct RescoreParams {
	uint64_t        n;
	const uint64_t *seq_off;  const uint8_t *seq;       // dna5 ranks
	const uint64_t *qual_off; const uint8_t *qual;      // phred ranks
	const uint64_t *cig_off;  const uint32_t *cig;
	const uint64_t *refn_off; const uint8_t *refn;      // XN tag: bit (p % 8) of byte p / 8 set ⇔ reference position p of the alignment is N; refn_off null ⇒ no record has it
	const int32_t  *mismatch_penalties, *n_penalties;   // [256] by quality rank
	int32_t         gap_open, gap_ext;
	double         *score;
	uint32_t       *flags;
};
void launch_rescore(const RescoreParams &P, cudaStream_t stream);
void launch_mapq(uint64_t n, const uint32_t *read_len, const uint32_t *other_len, const double *score, const double *next_score,
	uint8_t *mapq, uint32_t *flags, cudaStream_t stream);

// exclusive scan of 32-bit counts into 64-bit offsets, one CTA (cnt 16-byte aligned and padded to a multiple of four; off[n] = total)
void launch_scan_u32(const uint32_t *cnt, uint64_t n, unsigned long long *off, cudaStream_t stream);

// Record annotation text (pv3_annotate.cu): the OA entry of a record, or the SAM text of a CIGAR
struct OaParams {
	uint64_t        n;
	const uint32_t *ref_id;      // null ⇒ CIGAR text only
	const uint64_t *pos;
	const uint16_t *flag;
	const uint8_t  *mapq;
	const int32_t  *nm;
	const uint8_t  *has_nm;      // may be null
	const uint64_t *cig_off;     // [n] (with cig_len) or [n + 1]
	const uint32_t *cig_len;     // null ⇒ cig_off[i + 1] - cig_off[i]
	const uint32_t *cig;
	uint32_t        n_names;
	const uint64_t *name_off;
	const char     *names;
	uint32_t       *len;         // [n] scratch: text length of every record
	uint32_t       *cta_len;     // [ceil(n / 256)] padded to 4
	unsigned long long *cta_off; // [ceil(n / 256) + 1]
	uint64_t       *text_off;    // [n + 1]
	char           *text;
};
unsigned oa_ctas(uint64_t n);
void launch_oa_lengths(const OaParams &P, cudaStream_t stream);   // len, cta_len
void launch_oa_write(const OaParams &P, cudaStream_t stream);     // text_off, text (after launch_scan_u32 over cta_len)
uint64_t k3_default_grid_warps(int n_sms);
uint32_t k3_smem_pool_bytes();

} // namespace pv3
