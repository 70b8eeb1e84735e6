// panvc3_b200 — record annotation text on the device (sm_100a): the OA entry project_alignments appends to every projected
// record, and the SAM text of a CIGAR.
//
// Reference: project-alignments/project_alignments.cc:803-845 — per record
//     RNAME ',' POS ',' ('-' if FLAG & 0x10 else '+') ',' CIGAR ',' MAPQ ',' [NM] ';'
// built with an std::ostringstream (decimal integers, POS as seqan3 stores it, the ORIGINAL CIGAR as count + op char).
// With the kernels at 10⁹ records/s this per-record string building is the caller's bottleneck.
//
// Variable-length output: pass 1 computes every record's length and one sum per 256-record CTA, a one-CTA scan turns the
// CTA sums into offsets, pass 2 repeats the CTA-local scan and writes the bytes (a thread writes its own record: ≈ 35
// bytes; the text is small against the batch's other traffic).
#include "pv3_device.cuh"
#include "pv3_kernels.h"

namespace pv3 {

namespace {

constexpr unsigned OA_THREADS = 256;

__device__ __forceinline__ uint32_t digits10(uint64_t v)
{
	uint32_t d = 1;
	while (v >= 10000) { v /= 10000; d += 4; }
	return d + (v >= 10) + (v >= 100) + (v >= 1000);
}

// decimal of v at dst (no terminator); returns the number of bytes
__device__ __forceinline__ uint32_t put_u64(char *dst, uint64_t v)
{
	uint32_t const d = digits10(v);
	for (uint32_t i = d; i-- > 0; ) { dst[i] = (char) ('0' + (uint32_t) (v % 10)); v /= 10; }
	return d;
}

__device__ __forceinline__ void record_cigar(OaParams const &P, uint64_t i, uint64_t &c0, uint64_t &c1)
{
	c0 = P.cig_off[i];
	c1 = P.cig_len ? c0 + P.cig_len[i] : P.cig_off[i + 1];
}

__device__ uint32_t record_length(OaParams const &P, uint64_t i)
{
	uint64_t c0, c1; record_cigar(P, i, c0, c1);
	uint32_t len = 0;
	for (uint64_t c = c0; c < c1; ++c) len += digits10(__ldg(P.cig + c) >> 4) + 1;
	if (!P.ref_id) return len;
	uint32_t const rid = P.ref_id[i];
	if (rid < P.n_names) len += (uint32_t) (P.name_off[rid + 1] - P.name_off[rid]);
	len += 1 + digits10(P.pos[i]) + 1 + 1 + 1 + 1 + digits10(P.mapq[i]) + 1 + 1;
	if (P.has_nm && P.has_nm[i]) { int32_t const nm = P.nm[i]; len += digits10(nm < 0 ? (uint64_t) -(int64_t) nm : (uint64_t) nm) + (nm < 0); }
	return len;
}

__device__ uint32_t block_exclusive_scan(uint32_t mine, uint32_t *warp_tot, uint32_t &block_total)
{
	unsigned const lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	uint32_t incl = mine;
	#pragma unroll
	for (int d = 1; d < 32; d <<= 1) { uint32_t const t = __shfl_up_sync(FULL, incl, d); if ((int) lane >= d) incl += t; }
	if (lane == 31) warp_tot[warp] = incl;
	__syncthreads();
	uint32_t before = 0, total = 0;
	#pragma unroll
	for (unsigned w = 0; w < OA_THREADS / 32; ++w) { uint32_t const t = warp_tot[w]; if (w < warp) before += t; total += t; }
	block_total = total;
	return before + incl - mine;
}

} // namespace

__global__ void __launch_bounds__(OA_THREADS)
oa_length_kernel(OaParams P)
{
	__shared__ uint32_t warp_tot[OA_THREADS / 32];
	uint64_t const i = (uint64_t) blockIdx.x * OA_THREADS + threadIdx.x;
	uint32_t const len = i < P.n ? record_length(P, i) : 0u;
	if (i < P.n) P.len[i] = len;
	uint32_t total;
	block_exclusive_scan(len, warp_tot, total);
	if (threadIdx.x == 0) P.cta_len[blockIdx.x] = total;
}

__global__ void __launch_bounds__(OA_THREADS)
oa_write_kernel(OaParams P)
{
	__shared__ uint32_t warp_tot[OA_THREADS / 32];
	uint64_t const i = (uint64_t) blockIdx.x * OA_THREADS + threadIdx.x;
	uint32_t const len = i < P.n ? P.len[i] : 0u;
	uint32_t total;
	uint64_t const at = P.cta_off[blockIdx.x] + block_exclusive_scan(len, warp_tot, total);
	if (i >= P.n) return;
	P.text_off[i] = at;
	if (i == P.n - 1) P.text_off[P.n] = at + len;
	char *dst = P.text + at;
	if (P.ref_id) {
		uint32_t const rid = P.ref_id[i];
		if (rid < P.n_names) { for (uint64_t k = P.name_off[rid]; k < P.name_off[rid + 1]; ++k) *dst++ = P.names[k]; }
		*dst++ = ',';
		dst += put_u64(dst, P.pos[i]);
		*dst++ = ',';
		*dst++ = (P.flag[i] & 0x10u) ? '-' : '+';
		*dst++ = ',';
	}
	uint64_t c0, c1; record_cigar(P, i, c0, c1);
	for (uint64_t c = c0; c < c1; ++c) {
		uint32_t const op = __ldg(P.cig + c);
		dst += put_u64(dst, op >> 4);
		*dst++ = "MIDNSHP=X???????"[op & 15u];
	}
	if (P.ref_id) {
		*dst++ = ',';
		dst += put_u64(dst, P.mapq[i]);
		*dst++ = ',';
		if (P.has_nm && P.has_nm[i]) {
			int32_t const nm = P.nm[i];
			if (nm < 0) *dst++ = '-';
			dst += put_u64(dst, nm < 0 ? (uint64_t) -(int64_t) nm : (uint64_t) nm);
		}
		*dst++ = ';';
	}
}

unsigned oa_ctas(uint64_t n) { return (unsigned) ((n + OA_THREADS - 1) / OA_THREADS); }
void launch_oa_lengths(const OaParams &P, cudaStream_t stream) { if (P.n) oa_length_kernel<<<oa_ctas(P.n), OA_THREADS, 0, stream>>>(P); }
void launch_oa_write(const OaParams &P, cudaStream_t stream) { if (P.n) oa_write_kernel<<<oa_ctas(P.n), OA_THREADS, 0, stream>>>(P); }

} // namespace pv3
