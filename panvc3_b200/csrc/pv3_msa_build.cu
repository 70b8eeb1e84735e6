// panvc3_b200 — index-msa on the device: gap bits and the unaligned sequence of one MSA row in ONE pass over the text (sm_100a).
//
// Reference: index-msa/index_vector_builder.cc:28-50 — bit i = ('-' == text[i]); every other byte goes to the unaligned
// FASTA in order. HBM-bound byte → bit compaction plus a stream compaction: per row of n columns the kernel reads n bytes,
// writes n/8 bytes of gap words and one byte per non-gap column (≈ 2.1 n for a typical MSA).
//
// Two passes over 16 KiB tiles, no dependency between CTAs. Pass 1: gap masks from a SWAR byte compare, four lanes
// assemble one 64-bit word by shuffles, one non-gap count per tile. A one-CTA scan turns the counts into output offsets.
// Pass 2 (only when the sequence is wanted) compacts each tile's non-gap bytes in shared memory and stores them as aligned
// 32-bit words; it walks the row backwards so that the tail of pass 1's reads is still in the L2.
// (A single-pass variant with a decoupled look-back measured 0.30 ms per 256 MiB row on B200 against 0.07 ms for pass 1
// alone: each CTA sat on its registers for the look-back's L2 round trips. Two passes re-read the text but never wait.)
#include <mutex>

#include "pv3_device.cuh"
#include "pv3_kernels.h"
#include "pv3_ctx.h"

namespace pv3 {

namespace {

constexpr unsigned EX_THREADS = 256;
constexpr unsigned EX_ITERS   = 4;
constexpr unsigned EX_TILE    = EX_THREADS * 16 * EX_ITERS;      // 16 KiB of text per CTA
constexpr unsigned EX_SEGS    = EX_ITERS * (EX_THREADS / 32);    // one 512-byte segment per (iteration, warp)

// one bit per byte of x that equals '-' (exact: no borrow between bytes)
__device__ __forceinline__ uint32_t gap_nibble(uint32_t x)
{
	uint32_t const y = x ^ 0x2D2D2D2Du;
	uint32_t const z = ~(((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y | 0x7F7F7F7Fu);    // 0x80 where the byte of y is zero
	return ((z >> 7) * 0x10204080u) >> 28;
}

} // namespace

// Pass 1 — gap bits, and the number of non-gap bytes of every tile: 16 bytes per lane, four lanes assemble one word.
template <bool WITH_COUNTS>
__global__ void __launch_bounds__(EX_THREADS)
msa_bits_kernel(const uint4 *__restrict__ text, uint64_t n, uint64_t *__restrict__ gap_words, uint32_t *__restrict__ tile_cnt)
{
	__shared__ uint32_t warp_cnt[EX_THREADS / 32];
	unsigned const tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	uint64_t const tile_byte = (uint64_t) blockIdx.x * EX_TILE;
	uint64_t const n_words = (n + 63) / 64;
	uint4    v[EX_ITERS];
	uint32_t valid[EX_ITERS];
	#pragma unroll
	for (unsigned k = 0; k < EX_ITERS; ++k) {
		uint64_t const off = tile_byte + (uint64_t) (k * EX_THREADS + tid) * 16;
		valid[k] = off >= n ? 0u : (uint32_t) (n - off < 16 ? n - off : 16);
		v[k] = valid[k] ? __ldg(text + off / 16) : make_uint4(0, 0, 0, 0);
	}
	uint32_t non_gap = 0;
	#pragma unroll
	for (unsigned k = 0; k < EX_ITERS; ++k) {
		uint32_t m = gap_nibble(v[k].x) | (gap_nibble(v[k].y) << 4) | (gap_nibble(v[k].z) << 8) | (gap_nibble(v[k].w) << 12);
		m &= (1u << valid[k]) - 1u;
		non_gap += valid[k] - __popc(m);
		uint64_t w = (uint64_t) m << (16 * (lane & 3u));
		w |= shfl_u64(w, (int) (lane ^ 1u));
		w |= shfl_u64(w, (int) (lane ^ 2u));
		uint64_t const word = (tile_byte + (uint64_t) (k * EX_THREADS + warp * 32) * 16) / 64 + lane / 4;
		if ((lane & 3u) == 0 && word < n_words) __stcs(gap_words + word, w);
	}
	if (WITH_COUNTS) {
		non_gap = __reduce_add_sync(FULL, non_gap);
		if (lane == 0) warp_cnt[warp] = non_gap;
		__syncthreads();
		if (tid == 0) {
			uint32_t t = 0;
			#pragma unroll
			for (unsigned w = 0; w < EX_THREADS / 32; ++w) t += warp_cnt[w];
			tile_cnt[blockIdx.x] = t;
		}
	}
}

// Exclusive scan of the tile counts: tile_off[0 .. n_tiles], the last element the total. One CTA, 4096 counts per step
// (tile_cnt is padded to a multiple of four and 16-byte aligned).
__global__ void __launch_bounds__(1024)
msa_scan_kernel(const uint4 *__restrict__ tile_cnt4, uint64_t n_tiles, unsigned long long *__restrict__ tile_off)
{
	__shared__ unsigned long long warp_sum[32];
	__shared__ unsigned long long carry_s;
	unsigned const tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	unsigned long long carry = 0;
	for (uint64_t base = 0; base < n_tiles; base += 4096) {
		uint64_t const i0 = base + 4 * tid;
		uint4 c = i0 < n_tiles ? tile_cnt4[i0 / 4] : make_uint4(0, 0, 0, 0);
		if (i0 + 1 >= n_tiles) c.y = 0;
		if (i0 + 2 >= n_tiles) c.z = 0;
		if (i0 + 3 >= n_tiles) c.w = 0;
		unsigned long long const mine = (unsigned long long) c.x + c.y + c.z + c.w;
		unsigned long long incl = mine;
		#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { unsigned long long const t = shfl_u64(incl, (int) (lane >= (unsigned) d ? lane - d : lane)); if ((int) lane >= d) incl += t; }
		if (lane == 31) warp_sum[warp] = incl;
		__syncthreads();
		if (warp == 0) {
			unsigned long long const ws = warp_sum[lane];
			unsigned long long wi = ws;
			#pragma unroll
			for (int d = 1; d < 32; d <<= 1) { unsigned long long const t = shfl_u64(wi, (int) (lane >= (unsigned) d ? lane - d : lane)); if ((int) lane >= d) wi += t; }
			warp_sum[lane] = wi - ws;
			if (lane == 31) carry_s = wi;
		}
		__syncthreads();
		unsigned long long run = carry + warp_sum[warp] + incl - mine;
		if (i0 < n_tiles) tile_off[i0] = run;
		run += c.x; if (i0 + 1 < n_tiles) tile_off[i0 + 1] = run;
		run += c.y; if (i0 + 2 < n_tiles) tile_off[i0 + 2] = run;
		run += c.z; if (i0 + 3 < n_tiles) tile_off[i0 + 3] = run;
		carry += carry_s;
		__syncthreads();
	}
	if (tid == 0) tile_off[n_tiles] = carry;
}

// Pass 2 — the unaligned sequence. A warp owns 2 KiB of the tile as four 512-byte segments; in a segment lane l holds the
// 32-bit words l, l + 32, l + 64, l + 96, so that the lanes' compacted bytes are neighbours in shared memory (byte and word
// stores of a warp then fall into distinct banks). Tiles are taken from the END of the row first: that is the text pass 1
// read last, so a good part of this pass is served by the L2.
template <bool WHOLE>   // WHOLE: no column of the tile is past the end of the row
__device__ __forceinline__ void msa_compact_tile(const uint32_t *__restrict__ text, uint64_t n, uint64_t tile, uint8_t *__restrict__ ungapped,
	const unsigned long long *__restrict__ tile_off, uint8_t *comp, uint32_t *seg_cnt, uint32_t *sel_lut)
{
	unsigned const tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	if (tid < 16) {
		uint32_t sel = 0x4444u, at = 0;                                  // 4 = a byte of the second (zero) operand
		for (unsigned b = 0; b < 4; ++b) if ((tid >> b) & 1u) { sel = (sel & ~(0xFu << (4 * at))) | (b << (4 * at)); ++at; }
		sel_lut[tid] = sel;
	}
	#pragma unroll
	for (unsigned i = tid; i < (EX_TILE + 16) / 16; i += EX_THREADS) ((uint4 *) comp)[i] = make_uint4(0, 0, 0, 0);
	uint64_t const warp_byte = tile * EX_TILE + (uint64_t) warp * (EX_TILE / (EX_THREADS / 32));
	const uint32_t *src = text + warp_byte / 4 + lane;

	uint32_t x[4][4];          // [segment][word j]: bytes warp_byte + 512 k + 128 j + 4 lane ...
	#pragma unroll
	for (unsigned k = 0; k < 4; ++k) {
		#pragma unroll
		for (unsigned j = 0; j < 4; ++j)
			x[k][j] = (WHOLE || warp_byte + 512 * k + 128 * j + 4 * lane < n) ? __ldcs(src + 128 * k + 32 * j) : 0u;
	}
	uint64_t const g = tile_off[tile];
	uint32_t nm16[4];          // per segment: non-gap nibbles of the four words
	uint32_t excl[4], tot[4];  // per segment: packed (8 bits per word) lane-exclusive prefix / warp total of non-gap bytes
	#pragma unroll
	for (unsigned k = 0; k < 4; ++k) {
		uint32_t packed = 0;
		nm16[k] = 0;
		#pragma unroll
		for (unsigned j = 0; j < 4; ++j) {
			uint32_t vmask = 0xFu;
			if (!WHOLE) {
				uint64_t const off = warp_byte + 512 * k + 128 * j + 4 * lane;
				vmask = off >= n ? 0u : (n - off >= 4 ? 0xFu : (1u << (uint32_t) (n - off)) - 1u);
			}
			uint32_t const nm = ~gap_nibble(x[k][j]) & vmask;
			nm16[k] |= nm << (4 * j);
			packed |= (uint32_t) __popc(nm) << (8 * j);
		}
		uint32_t incl = packed;
		#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
		excl[k] = incl - packed;
		tot[k] = __shfl_sync(FULL, incl, 31);
		if (lane == 0) seg_cnt[warp * 4 + k] = (tot[k] & 0xFFu) + ((tot[k] >> 8) & 0xFFu) + ((tot[k] >> 16) & 0xFFu) + (tot[k] >> 24);
	}
	__syncthreads();

	// exclusive scan of the 32 segment counts (every warp for itself: cheaper than another barrier)
	uint32_t const seg_mine = seg_cnt[lane];
	uint32_t seg_incl = seg_mine;
	#pragma unroll
	for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, seg_incl, d); if ((int) lane >= d) seg_incl += t; }
	uint32_t const tile_cnt = __shfl_sync(FULL, seg_incl, 31);
	uint32_t const seg_excl = seg_incl - seg_mine;

	// compaction into shared memory
	#pragma unroll
	for (unsigned k = 0; k < 4; ++k) {
		uint32_t base = __shfl_sync(FULL, seg_excl, (int) (warp * 4 + k));
		#pragma unroll
		for (unsigned j = 0; j < 4; ++j) {
			// the word's non-gap bytes, packed low, OR-ed into the (zeroed) one or two aligned words they fall into
			uint32_t const pos = base + ((excl[k] >> (8 * j)) & 0xFFu);
			uint32_t const nm = (nm16[k] >> (4 * j)) & 0xFu;
			uint32_t const cw = __byte_perm(x[k][j], 0u, sel_lut[nm]);
			unsigned long long const t = (unsigned long long) cw << (8 * (pos & 3u));
			uint32_t *dst = (uint32_t *) comp + (pos >> 2);
			atomicOr(dst, (uint32_t) t);
			if ((uint32_t) (t >> 32)) atomicOr(dst + 1, (uint32_t) (t >> 32));
			base += (tot[k] >> (8 * j)) & 0xFFu;
		}
	}
	__syncthreads();

	// tile_cnt bytes of comp → ungapped[g ...): head bytes up to a 4-byte boundary, aligned words, tail bytes
	uint32_t head = (uint32_t) ((4 - (g & 3)) & 3);
	if (head > tile_cnt) head = tile_cnt;
	uint32_t const nw = (tile_cnt - head) / 4, tail = tile_cnt - head - 4 * nw;
	if (tid < head) ungapped[g + tid] = comp[tid];
	if (tid < tail) ungapped[g + head + 4 * nw + tid] = comp[head + 4 * nw + tid];
	uint32_t *out32 = (uint32_t *) (ungapped + g + head);
	uint32_t const *c32 = (uint32_t const *) comp;
	uint32_t const sel = 0x3210u + 0x1111u * (head & 3u);
	for (uint32_t i = tid; i < nw; i += EX_THREADS) {
		uint32_t const s = head + 4 * i;
		__stcs(out32 + i, __byte_perm(c32[s >> 2], c32[(s >> 2) + 1], sel));
	}
}

__global__ void __launch_bounds__(EX_THREADS)
msa_compact_kernel(const uint32_t *__restrict__ text, uint64_t n, uint64_t n_tiles, uint8_t *__restrict__ ungapped,
	const unsigned long long *__restrict__ tile_off)
{
	__shared__ __align__(16) uint8_t comp[EX_TILE + 16];
	__shared__ uint32_t seg_cnt[EX_SEGS];
	__shared__ uint32_t sel_lut[16];   // PRMT selector that moves the bytes named by a 4-bit mask to the low end, zeros above
	uint64_t const tile = n_tiles - 1 - blockIdx.x;
	if ((tile + 1) * EX_TILE <= n) msa_compact_tile <true>(text, n, tile, ungapped, tile_off, comp, seg_cnt, sel_lut);
	else msa_compact_tile <false>(text, n, tile, ungapped, tile_off, comp, seg_cnt, sel_lut);
}

// cnt: 16-byte aligned, readable up to the next multiple of four entries; off: n + 1 entries
void launch_scan_u32(const uint32_t *cnt, uint64_t n, unsigned long long *off, cudaStream_t stream)
{
	msa_scan_kernel<<<1, 1024, 0, stream>>>((const uint4 *) cnt, n, off);
}

uint64_t msa_extract_tiles(uint64_t n) { return (n + EX_TILE - 1) / EX_TILE; }

// scratch: msa_extract_scratch_words(n) 64-bit words, 16-byte aligned — the 32-bit tile counts (padded to a multiple of
// four), then the tile offsets [0 .. tiles]; the last offset, at word msa_extract_total_word(n), is the number of non-gap
// bytes (0 when no sequence is wanted). text must be 16-byte aligned and readable up to the next multiple of 16.
static uint64_t cnt_words(uint64_t n_tiles) { return (n_tiles + 3) / 4 * 2; }
uint64_t msa_extract_total_word(uint64_t n) { uint64_t const t = msa_extract_tiles(n); return cnt_words(t) + t; }
uint64_t msa_extract_scratch_words(uint64_t n) { return msa_extract_total_word(n) + 1; }

cudaError_t launch_msa_extract(const char *text, uint64_t n, uint64_t *gap_words, char *ungapped, unsigned long long *scratch,
	cudaStream_t stream, unsigned *launches)
{
	uint64_t const n_tiles = msa_extract_tiles(n);
	if (launches) *launches = 0;
	if (!n_tiles || !ungapped) {
		cudaError_t e = cudaMemsetAsync(scratch + msa_extract_total_word(n), 0, 8, stream);
		if (e != cudaSuccess || !n_tiles) return e;
		msa_bits_kernel <false><<<(unsigned) n_tiles, EX_THREADS, 0, stream>>>((const uint4 *) text, n, gap_words, nullptr);
		if (launches) *launches = 1;
		return cudaGetLastError();
	}
	uint32_t *tile_cnt = (uint32_t *) scratch;
	unsigned long long *tile_off = scratch + cnt_words(n_tiles);
	msa_bits_kernel <true><<<(unsigned) n_tiles, EX_THREADS, 0, stream>>>((const uint4 *) text, n, gap_words, tile_cnt);
	msa_scan_kernel<<<1, 1024, 0, stream>>>((const uint4 *) tile_cnt, n_tiles, tile_off);
	msa_compact_kernel<<<(unsigned) n_tiles, EX_THREADS, 0, stream>>>((const uint32_t *) text, n, n_tiles, (uint8_t *) ungapped, tile_off);
	if (launches) *launches = 3;
	return cudaGetLastError();
}

// ---- allocation-free entry point (no counterpart in the reference) ------------
// Callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it) get the
// row extraction with its scratch — tile counts and tile offsets — in a file-scope device array instead of a context buffer.
// Every GPU that loads the library holds the array until exit; every call on a device shares it (calls that overlap on the
// GPU race) and it keeps the last call's contents: each call rewrites all words it reads (the tile counts of all its tiles,
// the offsets, the total) before reading them.
namespace {

constexpr uint64_t NOALLOC_MAX_COLUMNS = 1ull << 30;                       // 65536 tiles of 16 KiB
constexpr uint64_t NOALLOC_SCRATCH_WORDS = ((NOALLOC_MAX_COLUMNS / EX_TILE + 3) / 4 * 2) + NOALLOC_MAX_COLUMNS / EX_TILE + 1 + 3;   // 768 KiB
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NOALLOC_SCRATCH_WORDS];

constexpr int NOALLOC_MAX_DEVICES = 64;
unsigned long long *g_noalloc_addr[NOALLOC_MAX_DEVICES] = {};              // device address of the array, per device
std::mutex g_noalloc_mutex;

// In host code `g_noalloc_scratch` is only a handle: its device address must come from cudaGetSymbolAddress, once per
// device, with that device current.
cudaError_t noalloc_scratch(int device, unsigned long long **out)
{
	std::lock_guard <std::mutex> lock(g_noalloc_mutex);
	if (!g_noalloc_addr[device]) {
		void *p = nullptr;
		cudaError_t const e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
		if (e != cudaSuccess) return e;
		g_noalloc_addr[device] = (unsigned long long *) p;
	}
	*out = g_noalloc_addr[device];
	return cudaSuccess;
}

} // namespace

} // namespace pv3

extern "C" int pv3_msa_row_extract_noalloc(int device, void *cuda_stream, const char *text, uint64_t n, uint64_t *gap_words, char *ungapped,
	uint64_t *n_ungapped)
{
	using namespace pv3;
	if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(nullptr, PV3_ERR_INVALID, "pv3_msa_row_extract_noalloc: device %d out of range", device);
	if (n > NOALLOC_MAX_COLUMNS)
		return fail(nullptr, PV3_ERR_CAPACITY, "pv3_msa_row_extract_noalloc: %llu columns exceed the fixed scratch (at most %llu columns per call); use pv3_msa_row_extract_device",
			(unsigned long long) n, (unsigned long long) NOALLOC_MAX_COLUMNS);
	if (n && (!text || !gap_words)) return fail(nullptr, PV3_ERR_INVALID, "pv3_msa_row_extract_noalloc: NULL argument");
	if ((uintptr_t) text % 16 || (uintptr_t) gap_words % 8 || (uintptr_t) ungapped % 4 || (uintptr_t) n_ungapped % 8)
		return fail(nullptr, PV3_ERR_INVALID, "pv3_msa_row_extract_noalloc: text must be 16-byte, gap_words and n_ungapped 8-byte, ungapped 4-byte aligned");
	cudaError_t e = cudaSetDevice(device);
	unsigned long long *scratch = nullptr;
	if (e == cudaSuccess) e = noalloc_scratch(device, &scratch);
	if (e != cudaSuccess) return fail(nullptr, PV3_ERR_CUDA, "pv3_msa_row_extract_noalloc: %s", cudaGetErrorString(e));
	cudaStream_t const stream = (cudaStream_t) cuda_stream;
	e = launch_msa_extract(text, n, gap_words, ungapped, scratch, stream, nullptr);
	if (e == cudaSuccess && n_ungapped)
		e = cudaMemcpyAsync(n_ungapped, scratch + msa_extract_total_word(n), 8, cudaMemcpyDeviceToDevice, stream);
	if (e != cudaSuccess) return fail(nullptr, PV3_ERR_CUDA, "pv3_msa_row_extract_noalloc: %s", cudaGetErrorString(e));
	return PV3_OK;
}

