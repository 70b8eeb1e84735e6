// panvc3_b200 — device-side data layout and warp-cooperative rank/select primitives (sm_100a).
//
// Replaces msa_index::sequence_entry over sdsl::rrr_vector<15> (reference include/panvc3/msa_index.hh:47-74,239-258;
// lib/sdsl-lite/include/sdsl/rrr_vector_15.hpp:460-559 rank, :664-701 select).  rank0/select0/access are pure functions
// of the bit sequence, so the H0-compressed blocks are replaced by plain words + a block directory that one warp
// reads with coalesced loads:
//
//   words      uint64, LSB first, bit = 1 ⇔ gap; bits >= n_bits are 1; PAD_WORDS all-ones words follow so that a
//              32-word window load never needs a bounds check
//   blk_zeros  uint64 per 512-bit block: zeros in [0, 512*b); n_blocks+1 entries + 32 entries of padding (= n_zeros)
//   sel        uint32 per SEL_RATE zeros: index of the block holding zero number s*SEL_RATE+1 (1-based); one sentinel
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace pv3 {

constexpr int      BLK_WORDS = 8;            // 512-bit blocks
constexpr int      BLK_BITS  = 512;
constexpr uint32_t SEL_RATE  = 512;            // one sample per block's worth of zeros: the search below is 1-2 loads deep
constexpr int      PAD_WORDS = 72;           // >= 64 + BLK_WORDS
constexpr unsigned FULL      = 0xffffffffu;

struct RowDesc {
	const uint64_t *words;
	const uint64_t *blk_zeros;
	const uint32_t *sel;
	uint64_t        n_bits;
	uint64_t        n_zeros;
	uint64_t        n_blocks;   // ceil(n_bits / 512), at least 1
	uint64_t        n_words_alloc;
};

struct RefDesc {
	const char    *bases;   // raw FASTA bytes (rewrite compares raw bytes)
	const uint8_t *codes4;  // two bases per byte (even base in the low nibble): dna5 rank of assign_char(byte) in bits 0-2 | 8 unless the
	                        // byte is exactly one of "ACGNT"; (query rank == nibble) ⇔ (dna5::to_char(query) == byte). 512 readable
	                        // pad bytes (0xff) follow.
	uint64_t       n;
};

// One realigned indel run (indel_run_checker.cc:104-118 + alignment_projector.cc:61-73).
struct Job {
	uint64_t rec;
	uint64_t ref_loc;      // DP window start in the target contig (true window, m_ref_range.location)
	uint32_t ref_len;
	uint32_t qry_loc;
	uint32_t qry_len;
	uint32_t op_begin;     // [op_begin, op_end) of the record's rewritten ops is replaced
	uint32_t op_end;
	uint32_t out_len;      // realigned ops (filled by the DP kernel)
	uint64_t out_off;      // offset into the DP op arena
};

struct Counters {
	unsigned long long arena_cursor;   // CIGAR arena words handed out
	unsigned long long job_cursor;     // runs found
	unsigned long long n_flagged;
	unsigned long long n_realigned;    // records with >= 1 run
	unsigned long long dp_cells;       // sum R*Q
	unsigned long long dp_path;        // sum R+Q   (DP op arena bound)
	unsigned long long dp_pool;        // sum of per-job scratch bytes
	unsigned long long dp_max_pool;    // max per-job scratch bytes
	unsigned long long dp_arena_cursor;
	unsigned long long job_ticket;     // work-stealing counter of the DP kernel
	unsigned long long n_deferred;     // records the thread-per-record kernel handed to the warp-per-record kernel
	unsigned int       overflow;       // bit0: CIGAR arena, bit1: rr/jobs, bit2: DP op arena, bit3: HBM trace slots, bit7: segment list, bit8: segment op arena
	unsigned int       pad;
	unsigned long long n_big_jobs;     // runs too large for the thread-per-run DP kernel (the warp-per-run kernel takes them)
	unsigned long long big_cursor;     // entries of the big-run list written by the thread-per-run kernel
	unsigned long long n_complex;      // records with I S H P D N ops the first launch of the word kernel left to the second
	unsigned long long n_planned;      // deferred records cut into segments (pv3_rewrite_seg.cu)
	unsigned long long seg_cursor;     // segments planned (the number NEEDED, also when the list overflowed)
	unsigned long long seg_ops_cursor; // words of the segment op arena handed out (likewise)
	unsigned long long n_deferred2;    // records left to the warp-per-record kernel after the segment pass
};

// BAM op codes
enum : uint32_t { OP_M = 0, OP_I = 1, OP_D = 2, OP_N = 3, OP_S = 4, OP_H = 5, OP_P = 6, OP_EQ = 7, OP_X = 8, OP_NONE = 15 };

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemask_lt() { unsigned m; asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m)); return m; }

__device__ __forceinline__ uint64_t shfl_u64(uint64_t v, int src)
{
	unsigned lo = __shfl_sync(FULL, (unsigned) v, src);
	unsigned hi = __shfl_sync(FULL, (unsigned) (v >> 32), src);
	return ((uint64_t) hi << 32) | lo;
}

__device__ __forceinline__ uint64_t shfl_up_u64(uint64_t v, int delta)
{
	unsigned lo = __shfl_up_sync(FULL, (unsigned) v, delta);
	unsigned hi = __shfl_up_sync(FULL, (unsigned) (v >> 32), delta);
	return ((uint64_t) hi << 32) | lo;
}
__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, int mask)
{
	unsigned lo = __shfl_xor_sync(FULL, (unsigned) v, mask);
	unsigned hi = __shfl_xor_sync(FULL, (unsigned) (v >> 32), mask);
	return ((uint64_t) hi << 32) | lo;
}

// Position (0..31) of the k-th (1-based) set bit of x; caller guarantees popc(x) >= k. Branch-free popcount descent.
__device__ __forceinline__ unsigned select_in_u32(unsigned x, unsigned k)
{
	unsigned pos = 0, c;
	c = __popc(x & 0xffffu); if (k > c) { k -= c; pos += 16; x >>= 16; }
	c = __popc(x & 0xffu);   if (k > c) { k -= c; pos += 8;  x >>= 8; }
	c = __popc(x & 0xfu);    if (k > c) { k -= c; pos += 4;  x >>= 4; }
	c = __popc(x & 0x3u);    if (k > c) { k -= c; pos += 2;  x >>= 2; }
	c = x & 1u;              if (k > c) { pos += 1; }
	return pos;
}

// Position (0..63) of the k-th (1-based) set bit of x; caller guarantees popc(x) >= k.
__device__ __forceinline__ unsigned select_in_word(uint64_t x, unsigned k)
{
	unsigned lo = (unsigned) x, c = __popc(lo);
	if (k <= c) return select_in_u32(lo, k);
	return 32u + select_in_u32((unsigned) (x >> 32), k - c);
}

// ---- warp-cooperative queries: every lane passes the same arguments and receives the same result -----------------

// zeros in [0, pos)  — project_aligned_position (msa_index.hh:247-251)
__device__ __forceinline__ uint64_t rank0_warp(const RowDesc &r, uint64_t pos)
{
	if (pos >= r.n_bits) return r.n_zeros;
	unsigned const lane = lane_id();
	uint64_t const b = pos / BLK_BITS;
	unsigned const wi = (unsigned) ((pos % BLK_BITS) / 64), bit = (unsigned) (pos % 64);
	unsigned contrib = 0;
	if (lane <= wi) {
		uint64_t inv = ~__ldg(r.words + b * BLK_WORDS + lane);
		if (lane == wi) inv &= (bit ? ((1ull << bit) - 1ull) : 0ull);
		contrib = __popcll(inv);
	}
	contrib = __reduce_add_sync(FULL, contrib);
	return __ldg(r.blk_zeros + b) + contrib;
}

// position of the k-th zero, k >= 1 — aligned_position(p) = select0(p + 1) (msa_index.hh:239-244);
// returns n_bits when the row has fewer than k zeros (rrr_vector_15.hpp:666).
__device__ __forceinline__ uint64_t select0_warp(const RowDesc &r, uint64_t k)
{
	if (k == 0 || k > r.n_zeros) return r.n_bits;
	unsigned const lane = lane_id();
	uint64_t const s = (k - 1) / SEL_RATE;
	uint64_t lo = __ldg(r.sel + s), hi = __ldg(r.sel + s + 1);
	while (hi - lo >= 32) {
		uint64_t const mid = lo + (hi - lo + 1) / 2;
		if (__ldg(r.blk_zeros + mid) < k) lo = mid; else hi = mid - 1;
	}
	uint64_t const v = __ldg(r.blk_zeros + lo + lane);
	unsigned const m = __ballot_sync(FULL, (lo + lane <= hi) && (v < k));
	uint64_t const b = lo + __popc(m) - 1;                      // lane 0 always qualifies
	uint64_t const base = shfl_u64(v, __popc(m) - 1);
	unsigned need = (unsigned) (k - base);                      // 1..512
	uint64_t inv = 0;
	unsigned pc = 0;
	if (lane < BLK_WORDS) { inv = ~__ldg(r.words + b * BLK_WORDS + lane); pc = __popcll(inv); }
	unsigned incl = pc;
	#pragma unroll
	for (int d = 1; d < BLK_WORDS; d <<= 1) { unsigned t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
	unsigned const hit = __ballot_sync(FULL, lane < BLK_WORDS && incl >= need);
	int const wl = __ffs(hit) - 1;
	unsigned const before = __shfl_sync(FULL, incl - pc, wl);
	uint64_t const w = shfl_u64(inv, wl);
	return b * BLK_BITS + (uint64_t) wl * 64 + select_in_word(w, need - before);
}

// ---- single-thread versions (position-only queries: mates, parity probes) -----------------------------------------

__device__ inline uint64_t rank0_thread(const RowDesc &r, uint64_t pos)
{
	if (pos >= r.n_bits) return r.n_zeros;
	uint64_t const b = pos / BLK_BITS;
	unsigned const wi = (unsigned) ((pos % BLK_BITS) / 64), bit = (unsigned) (pos % 64);
	uint64_t z = __ldg(r.blk_zeros + b);
	for (unsigned w = 0; w < wi; ++w) z += __popcll(~__ldg(r.words + b * BLK_WORDS + w));
	if (bit) z += __popcll(~__ldg(r.words + b * BLK_WORDS + wi) & ((1ull << bit) - 1ull));
	return z;
}

__device__ inline uint64_t select0_thread(const RowDesc &r, uint64_t k)
{
	if (k == 0 || k > r.n_zeros) return r.n_bits;
	uint64_t const s = (k - 1) / SEL_RATE;
	uint64_t lo = __ldg(r.sel + s), hi = __ldg(r.sel + s + 1);
	while (hi - lo > 3) {                                           // long gap stretches only
		uint64_t const mid = lo + (hi - lo + 1) / 2;
		if (__ldg(r.blk_zeros + mid) < k) lo = mid; else hi = mid - 1;
	}
	// the block is one of lo..lo+3: blk_zeros is monotone and padded with n_zeros (>= k), so entries past `hi` never count
	uint64_t const v0 = __ldg(r.blk_zeros + lo), v1 = __ldg(r.blk_zeros + lo + 1), v2 = __ldg(r.blk_zeros + lo + 2), v3 = __ldg(r.blk_zeros + lo + 3);
	unsigned const adv = (unsigned) (v1 < k) + (unsigned) (v2 < k) + (unsigned) (v3 < k);
	uint64_t const base = adv == 0 ? v0 : adv == 1 ? v1 : adv == 2 ? v2 : v3;
	uint64_t const blk = lo + adv;
	unsigned need = (unsigned) (k - base);
	const ulonglong2 *wp = reinterpret_cast <const ulonglong2 *>(r.words + blk * BLK_WORDS);   // 64-byte aligned
	ulonglong2 const a = __ldg(wp), b = __ldg(wp + 1), c = __ldg(wp + 2), d = __ldg(wp + 3);
	uint64_t const inv[BLK_WORDS] = { ~a.x, ~a.y, ~b.x, ~b.y, ~c.x, ~c.y, ~d.x, ~d.y };
	uint64_t word = inv[BLK_WORDS - 1];
	unsigned wi = BLK_WORDS - 1;
	bool found = false;
	#pragma unroll
	for (unsigned w = 0; w < (unsigned) BLK_WORDS; ++w) {
		unsigned const pc = __popcll(inv[w]);
		if (!found) { if (need <= pc) { found = true; word = inv[w]; wi = w; } else need -= pc; }
	}
	if (!found) return r.n_bits; // unreachable for a consistent directory
	return blk * BLK_BITS + wi * 64 + select_in_word(word, need);
}

} // namespace pv3
