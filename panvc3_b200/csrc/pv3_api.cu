// panvc3_b200 — host side of the C-ABI (include/panvc3_b200.h): context, index upload, batch pipeline.
//
// The context replaces, for the hot path, what project_task / input_processor own in the reference
// (project-alignments/project_alignments.cc:219-330,485-654): per-task alignment_projector instances and their reusable
// vectors become device buffers owned here and reused across batches; the loaded msa_index and the reference contigs
// (sequence_buffer_store) become one immutable device-resident index.
// There is no CPU fallback: every entry point fails with PV3_ERR_CUDA if the device is unusable.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pv3_kernels.h"

using namespace pv3;

namespace {

std::string g_create_error;

struct DevBuf {
	void  *p = nullptr;
	size_t cap = 0;
	cudaError_t ensure(size_t bytes)
	{
		if (bytes <= cap) return cudaSuccess;
		if (p) cudaFree(p);
		p = nullptr; cap = 0;
		size_t const want = bytes + bytes / 4 + 256;
		cudaError_t e = cudaMalloc(&p, want);
		if (e == cudaSuccess) cap = want;
		return e;
	}
	void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
	template <typename T> T *as() const { return (T *) p; }
};

struct HostRow { std::vector <uint64_t> words; uint64_t n_bits; };
struct HostRef { std::string bases; };

} // namespace

// One of two buffer sets of the chunked host pipeline (device mirrors of a chunk's inputs and outputs).
struct HostSet {
	DevBuf   in_src_row, in_dst_row, in_dst_ref, in_src_pos, in_seq_off, in_seq, in_qual_off, in_qual, in_cig_off, in_cig, in_seq_len;
	DevBuf   out_dst_pos, out_flags, out_cig_off, out_cig_len, out_arena, out_rr_off, out_rr_len, out_rr;
	uint64_t arena_cap_hint = 0, rr_cap_hint = 0;
	void release()
	{
		for (DevBuf *b : { &in_src_row, &in_dst_row, &in_dst_ref, &in_src_pos, &in_seq_off, &in_seq, &in_qual_off, &in_qual, &in_cig_off, &in_cig, &in_seq_len,
				&out_dst_pos, &out_flags, &out_cig_off, &out_cig_len, &out_arena, &out_rr_off, &out_rr_len, &out_rr }) b->release();
	}
};

struct pv3_ctx {
	int          device = 0;
	int          n_sms = 148;
	cudaStream_t stream = nullptr;       // the stream all work is issued on
	cudaStream_t own_stream = nullptr;
	cudaStream_t stream2 = nullptr;      // second stream of the chunked host pipeline
	cudaEvent_t  cev[2][4] = {};         // per buffer set: H2D begin/end, D2H begin/end
	HostSet      sets[2];
	cudaEvent_t  ev[8] = {};             // 0 start, 1 rewrite end, 2 realign end, 3 splice end, 4 thread-kernel end
	std::string  error;

	std::vector <HostRow> rows;
	std::vector <HostRef> refs;
	bool         committed = false;
	DevBuf       index_mem;              // words, directories, contigs — one allocation
	DevBuf       row_desc, ref_desc;
	uint64_t     index_bytes = 0;
	uint64_t     target_code_bytes = 0;   // packed target codes of all references (chooses the thread kernel's layout)

	int32_t      gap_open = 0, gap_ext = -1;   // project-alignments/cmdline.ggo defaults
	int32_t      lut[84];
	int          stage = 0;
	uint64_t     chunk_records = 500000;     // records per chunk of the host pipeline (PV3_CHUNK_RECORDS overrides)
	uint64_t     chunk_edge = 125000;        // records of the first and the last chunk: a short pipeline fill and drain (PV3_CHUNK_EDGE; 0 = equal chunks)
	uint64_t     dp_thread_min = 16384;      // fewer small runs than this: the warp-per-run kernel takes them all
	bool         many_deferred = false;      // last batch: more than 1/16 of the records went to the warp-per-record kernel
	int          kernel_path = 0;            // 0: thread-per-record run-walk kernel + warp-per-record for deferred records; 1: warp-per-record only;
	                                         // 2: thread-per-record word-parallel kernel + warp-per-record for deferred records

	// device work buffers
	DevBuf       counters, jobs, dp_arena, pool, deferred, complex_list, complex_ce, big_list, d_packed, d_job_qual_off;
	DevBuf       ex_text, ex_words, ex_seq, ex_status;   // pv3_msa_row_extract
	DevBuf       an[16];                                 // pv3_format_oa / pv3_format_cigars
	Job         *h_jobs = nullptr; uint64_t *h_job_qual_off = nullptr; uint64_t h_jobs_cap = 0;   // pinned (lazy QUAL)
	uint8_t     *h_packed = nullptr; uint64_t h_packed_cap = 0;
	Counters    *h_counters = nullptr;   // pinned
	// host-path device mirrors
	DevBuf       in_src_row, in_dst_row, in_dst_ref, in_src_pos, in_seq_off, in_seq, in_qual_off, in_qual, in_cig_off, in_cig;
	DevBuf       out_dst_pos, out_flags, out_cig_off, out_cig_len, out_arena, out_rr_off, out_rr_len, out_rr;
	uint64_t     arena_cap_hint = 0, rr_cap_hint = 0;

	pv3_timing   timing = {};
	uint64_t     total_launches = 0;
};

namespace {

int fail(pv3_ctx *ctx, int code, const char *fmt, ...)
{
	char buf[512];
	va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
	if (ctx) ctx->error = buf; else g_create_error = buf;
	return code;
}

#define PV3_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
	return fail((ctx), PV3_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); } while (0)

// Malde quality scores truncated to int32 (align.hh:28-37); reference base at phred 41
// (alignment_projector.cc:108). Same expression order as the reference.
void default_quality_lut(int32_t *lut)
{
	for (int q = 0; q < 42; ++q) {
		double const lhs_e = 1.0 / std::pow(10.0, 41 / 10.0);
		double const rhs_e = 1.0 / std::pow(10.0, q / 10.0);
		double const combined_e = lhs_e + rhs_e - lhs_e / 3.0 * rhs_e * 4.0;
		lut[2 * q] = (int32_t) (2.0 + std::log2(1.0 - combined_e));
		lut[2 * q + 1] = (int32_t) (2.0 - std::log2(3.0) + std::log2(combined_e));
	}
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

} // namespace

extern "C" {

int pv3_abi_version(void) { return PV3_ABI_VERSION; }

int pv3_create(pv3_ctx **out, int device)
{
	if (!out) return fail(nullptr, PV3_ERR_INVALID, "pv3_create: out is NULL");
	*out = nullptr;
	int n_dev = 0;
	cudaError_t e = cudaGetDeviceCount(&n_dev);
	if (e != cudaSuccess || n_dev == 0)
		return fail(nullptr, PV3_ERR_CUDA, "pv3_create: no CUDA device (%s); panvc3_b200 has no CPU fallback",
			e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
	if (device < 0 || device >= n_dev) return fail(nullptr, PV3_ERR_INVALID, "pv3_create: device %d out of range (%d devices)", device, n_dev);
	PV3_CUDA(nullptr, cudaSetDevice(device));
	cudaDeviceProp prop;
	PV3_CUDA(nullptr, cudaGetDeviceProperties(&prop, device));
	if (prop.major < 10)
		return fail(nullptr, PV3_ERR_CUDA, "pv3_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
	pv3_ctx *ctx = new pv3_ctx;
	ctx->device = device;
	ctx->n_sms = prop.multiProcessorCount;
	default_quality_lut(ctx->lut);
	if (const char *e = getenv("PV3_CHUNK_RECORDS")) { long long v = atoll(e); if (v >= 1024) ctx->chunk_records = (uint64_t) v; }
	if (const char *e = getenv("PV3_CHUNK_EDGE")) { long long v = atoll(e); if (v >= 0) ctx->chunk_edge = (uint64_t) v; }
	if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return fail(nullptr, PV3_ERR_CUDA, "cudaStreamCreate failed"); }
	ctx->stream = ctx->own_stream;
	for (auto &ev : ctx->ev) if (cudaEventCreate(&ev) != cudaSuccess) { delete ctx; return fail(nullptr, PV3_ERR_CUDA, "cudaEventCreate failed"); }
	if (cudaMallocHost((void **) &ctx->h_counters, sizeof(Counters)) != cudaSuccess) { delete ctx; return fail(nullptr, PV3_ERR_CUDA, "cudaMallocHost failed"); }
	if (ctx->counters.ensure(sizeof(Counters)) != cudaSuccess) { delete ctx; return fail(nullptr, PV3_ERR_CUDA, "cudaMalloc failed"); }
	*out = ctx;
	return PV3_OK;
}

void pv3_destroy(pv3_ctx *ctx)
{
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	if (ctx->stream) cudaStreamSynchronize(ctx->stream);
	for (DevBuf *b : { &ctx->index_mem, &ctx->row_desc, &ctx->ref_desc, &ctx->counters, &ctx->jobs, &ctx->dp_arena, &ctx->pool, &ctx->deferred, &ctx->complex_list, &ctx->complex_ce, &ctx->big_list, &ctx->d_packed, &ctx->d_job_qual_off,
			&ctx->ex_text, &ctx->ex_words, &ctx->ex_seq, &ctx->ex_status,
			&ctx->in_src_row, &ctx->in_dst_row, &ctx->in_dst_ref, &ctx->in_src_pos, &ctx->in_seq_off, &ctx->in_seq, &ctx->in_qual_off,
			&ctx->in_qual, &ctx->in_cig_off, &ctx->in_cig, &ctx->out_dst_pos, &ctx->out_flags, &ctx->out_cig_off, &ctx->out_cig_len,
			&ctx->out_arena, &ctx->out_rr_off, &ctx->out_rr_len, &ctx->out_rr })
		b->release();
	for (auto &b : ctx->an) b.release();
	if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
	if (ctx->h_jobs) cudaFreeHost(ctx->h_jobs);
	if (ctx->h_job_qual_off) cudaFreeHost(ctx->h_job_qual_off);
	if (ctx->h_packed) cudaFreeHost(ctx->h_packed);
	for (auto &ev : ctx->ev) if (ev) cudaEventDestroy(ev);
	for (auto &hs : ctx->sets) hs.release();
	for (auto &row : ctx->cev) for (auto &ev : row) if (ev) cudaEventDestroy(ev);
	if (ctx->stream2) { cudaStreamSynchronize(ctx->stream2); cudaStreamDestroy(ctx->stream2); }
	if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
	delete ctx;
}

int pv3_set_stream(pv3_ctx *ctx, void *cuda_stream)
{
	if (!ctx) return PV3_ERR_INVALID;
	cudaSetDevice(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	ctx->stream = cuda_stream ? (cudaStream_t) cuda_stream : ctx->own_stream;
	return PV3_OK;
}

const char *pv3_last_error(const pv3_ctx *ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

int pv3_index_add_row(pv3_ctx *ctx, const uint64_t *gap_words, uint64_t n_bits)
{
	if (!ctx || (!gap_words && n_bits)) return fail(ctx, PV3_ERR_INVALID, "pv3_index_add_row: bad argument");
	ctx->committed = false;   // rows may be added later (lazily registered entries); the next commit re-uploads
	HostRow row;
	row.n_bits = n_bits;
	row.words.assign(gap_words, gap_words + (n_bits + 63) / 64);
	ctx->rows.emplace_back(std::move(row));
	return (int) ctx->rows.size() - 1;
}

int pv3_index_add_ref(pv3_ctx *ctx, const char *bases, uint64_t n)
{
	if (!ctx || (!bases && n)) return fail(ctx, PV3_ERR_INVALID, "pv3_index_add_ref: bad argument");
	ctx->committed = false;
	HostRef ref; ref.bases.assign(bases, bases + n);
	ctx->refs.emplace_back(std::move(ref));
	return (int) ctx->refs.size() - 1;
}

int pv3_index_commit(pv3_ctx *ctx)
{
	if (!ctx) return PV3_ERR_INVALID;
	if (ctx->committed) return PV3_OK;
	if (ctx->stream) cudaStreamSynchronize(ctx->stream);
	PV3_CUDA(ctx, cudaSetDevice(ctx->device));

	// Layout of the single index allocation: per row [words | blk_zeros | sel], then the contigs; 256-byte aligned parts.
	struct RowLayout { size_t words_off, blk_off, sel_off; uint64_t n_blocks, n_words_alloc, n_zeros, n_sel; };
	std::vector <RowLayout> lay(ctx->rows.size());
	size_t total = 0;
	for (size_t r = 0; r < ctx->rows.size(); ++r) {
		auto const &row = ctx->rows[r];
		RowLayout &L = lay[r];
		L.n_blocks = (row.n_bits + BLK_BITS - 1) / BLK_BITS; if (!L.n_blocks) L.n_blocks = 1;
		L.n_words_alloc = L.n_blocks * BLK_WORDS + PAD_WORDS;
		L.words_off = total; total = align_up(total + L.n_words_alloc * 8, 256);
		L.blk_off = total; total = align_up(total + (L.n_blocks + 1 + 32) * 8, 256);
		L.sel_off = total; // size known after counting zeros
		uint64_t zeros = 0;
		for (uint64_t w = 0; w < row.words.size(); ++w) {
			uint64_t v = row.words[w];
			if (w == row.words.size() - 1 && row.n_bits % 64) v |= ~0ull << (row.n_bits % 64);
			zeros += (uint64_t) __builtin_popcountll(~v);
		}
		L.n_zeros = zeros;
		L.n_sel = (zeros ? (zeros - 1) / SEL_RATE + 1 : 0) + 1;
		total = align_up(total + L.n_sel * 4, 256);
	}
	std::vector <size_t> ref_off(ctx->refs.size());
	std::vector <size_t> code_off(ctx->refs.size());
	for (size_t r = 0; r < ctx->refs.size(); ++r) {
		ref_off[r] = total; total = align_up(total + ctx->refs[r].bases.size() + 512, 256);
		code_off[r] = total; total = align_up(total + (ctx->refs[r].bases.size() + 1) / 2 + 512, 256);
	}
	if (!total) total = 256;

	std::vector <uint8_t> host(total, 0);
	for (size_t r = 0; r < ctx->rows.size(); ++r) {
		auto const &row = ctx->rows[r];
		RowLayout const &L = lay[r];
		uint64_t *words = (uint64_t *) (host.data() + L.words_off);
		uint64_t *blk = (uint64_t *) (host.data() + L.blk_off);
		uint32_t *sel = (uint32_t *) (host.data() + L.sel_off);
		for (uint64_t w = 0; w < L.n_words_alloc; ++w) words[w] = (w < row.words.size()) ? row.words[w] : ~0ull;
		if (row.n_bits % 64) words[row.words.size() - 1] |= ~0ull << (row.n_bits % 64);
		uint64_t z = 0, next_sel = 0;   // next_sel: index of the next sample to write
		for (uint64_t b = 0; b < L.n_blocks; ++b) {
			blk[b] = z;
			uint64_t zb = 0;
			for (int w = 0; w < BLK_WORDS; ++w) zb += (uint64_t) __builtin_popcountll(~words[b * BLK_WORDS + w]);
			// zero numbers (1-based) in this block: z+1 .. z+zb; sample s covers zero number s*SEL_RATE+1
			while (next_sel + 1 < L.n_sel && next_sel * (uint64_t) SEL_RATE + 1 <= z + zb) sel[next_sel++] = (uint32_t) b;
			z += zb;
		}
		for (uint64_t b = L.n_blocks; b < L.n_blocks + 1 + 32; ++b) blk[b] = z;
		for (uint64_t s = next_sel; s < L.n_sel; ++s) sel[s] = (uint32_t) (L.n_blocks - 1);   // sentinel(s)
	}
	for (size_t r = 0; r < ctx->refs.size(); ++r) {
		auto const &bases = ctx->refs[r].bases;
		memcpy(host.data() + ref_off[r], bases.data(), bases.size());
		uint8_t *codes4 = host.data() + code_off[r];
		for (size_t i = 0; i < bases.size(); ++i) {
			uint8_t c;
			switch (bases[i]) {   // dna5::assign_char (dna5.hpp:99-110); exact = byte equals dna5::to_char(rank)
				case 'A': c = 0; break; case 'C': c = 1; break; case 'G': c = 2; break; case 'N': c = 3; break; case 'T': c = 4; break;
				case 'a': c = 8; break; case 'c': c = 9; break; case 'g': c = 10; break;
				case 't': case 'U': case 'u': c = 12; break;
				default: c = 11; break;
			}
			if (i & 1) codes4[i >> 1] |= (uint8_t) (c << 4); else codes4[i >> 1] = c;
		}
		if (bases.size() & 1) codes4[bases.size() >> 1] |= 0xf0;
		memset(codes4 + (bases.size() + 1) / 2, 0xff, 512);
	}

	PV3_CUDA(ctx, ctx->index_mem.ensure(total));
	PV3_CUDA(ctx, cudaMemcpy(ctx->index_mem.p, host.data(), total, cudaMemcpyHostToDevice));
	ctx->index_bytes = total;
	ctx->target_code_bytes = 0;
	for (auto const &rf : ctx->refs) ctx->target_code_bytes += (rf.bases.size() + 1) / 2;

	std::vector <RowDesc> rd(ctx->rows.size() ? ctx->rows.size() : 1);
	for (size_t r = 0; r < ctx->rows.size(); ++r) {
		uint8_t *base = ctx->index_mem.as <uint8_t>();
		rd[r].words = (const uint64_t *) (base + lay[r].words_off);
		rd[r].blk_zeros = (const uint64_t *) (base + lay[r].blk_off);
		rd[r].sel = (const uint32_t *) (base + lay[r].sel_off);
		rd[r].n_bits = ctx->rows[r].n_bits;
		rd[r].n_zeros = lay[r].n_zeros;
		rd[r].n_blocks = lay[r].n_blocks;
		rd[r].n_words_alloc = lay[r].n_words_alloc;
	}
	std::vector <RefDesc> fd(ctx->refs.size() ? ctx->refs.size() : 1);
	for (size_t r = 0; r < ctx->refs.size(); ++r) {
		fd[r].bases = (const char *) (ctx->index_mem.as <uint8_t>() + ref_off[r]);
		fd[r].codes4 = ctx->index_mem.as <uint8_t>() + code_off[r];
		fd[r].n = ctx->refs[r].bases.size();
	}
	PV3_CUDA(ctx, ctx->row_desc.ensure(rd.size() * sizeof(RowDesc)));
	PV3_CUDA(ctx, ctx->ref_desc.ensure(fd.size() * sizeof(RefDesc)));
	PV3_CUDA(ctx, cudaMemcpy(ctx->row_desc.p, rd.data(), rd.size() * sizeof(RowDesc), cudaMemcpyHostToDevice));
	PV3_CUDA(ctx, cudaMemcpy(ctx->ref_desc.p, fd.data(), fd.size() * sizeof(RefDesc), cudaMemcpyHostToDevice));
	ctx->committed = true;
	return PV3_OK;
}

int pv3_set_params(pv3_ctx *ctx, int32_t gap_open, int32_t gap_ext)
{
	if (!ctx) return PV3_ERR_INVALID;
	ctx->gap_open = gap_open; ctx->gap_ext = gap_ext;
	return PV3_OK;
}

int pv3_set_quality_lut(pv3_ctx *ctx, const int32_t *lut84)
{
	if (!ctx || !lut84) return PV3_ERR_INVALID;
	memcpy(ctx->lut, lut84, sizeof ctx->lut);
	return PV3_OK;
}

int pv3_get_quality_lut(const pv3_ctx *ctx, int32_t *lut84)
{
	if (!ctx || !lut84) return PV3_ERR_INVALID;
	memcpy(lut84, ctx->lut, sizeof ctx->lut);
	return PV3_OK;
}

int pv3_set_kernel_path(pv3_ctx *ctx, int path)
{
	if (!ctx || path < 0 || path > 2) return PV3_ERR_INVALID;
	ctx->kernel_path = path;
	return PV3_OK;
}

int pv3_set_dp_thread_min(pv3_ctx *ctx, uint64_t n)
{
	if (!ctx) return PV3_ERR_INVALID;
	ctx->dp_thread_min = n;
	return PV3_OK;
}

int pv3_set_stage(pv3_ctx *ctx, int stage)
{
	if (!ctx) return PV3_ERR_INVALID;
	ctx->stage = stage;
	return PV3_OK;
}

} // extern "C"

namespace {

constexpr size_t POOL_BUDGET = (size_t) 8 << 30;   // HBM scratch for traces of runs too large for shared memory

K1Params make_k1(pv3_ctx *ctx, const pv3_batch_in *in, pv3_batch_out *out)
{
	K1Params P;
	memset(&P, 0, sizeof P);
	P.n = in->n;
	P.src_row = in->src_row; P.dst_row = in->dst_row; P.dst_ref = in->dst_ref; P.src_pos = in->src_pos;
	P.seq_off = in->seq_off; P.seq = in->seq; P.seq_len = in->seq_len; P.qual_off = in->qual_off; P.qual = in->qual;
	P.cig_off = in->cig_off; P.cig = in->cig;
	P.rows = ctx->row_desc.as <RowDesc>(); P.n_rows = (uint32_t) ctx->rows.size();
	P.refs = ctx->ref_desc.as <RefDesc>(); P.n_refs = (uint32_t) ctx->refs.size();
	P.dst_pos = out->dst_pos; P.flags = out->flags; P.cig_out_off = out->cig_off; P.cig_out_len = out->cig_len;
	P.arena = out->cig; P.arena_cap = out->cig_cap;
	P.rr_off = out->rr_off; P.rr_len = out->rr_len; P.rr = out->rr; P.rr_cap = out->rr_cap;
	P.jobs = ctx->jobs.as <Job>();
	P.counters = ctx->counters.as <Counters>();
	P.gap_open = ctx->gap_open; P.gap_ext = ctx->gap_ext;
	P.stage = ctx->stage;
	return P;
}

// Small transfers between pinned host memory and the device done by a kernel (the pointers of cudaMallocHost memory are
// valid on the device under unified addressing). A cudaMemcpyAsync of a few hundred bytes would queue on the copy engine
// behind the OTHER stream's bulk chunk transfer and hold this chunk's next kernel for a millisecond; these loads and
// stores go over PCIe directly.
__global__ void small_copy_kernel(unsigned long long *dst, const unsigned long long *src, size_t n_words)
{
	for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (size_t) gridDim.x * blockDim.x) dst[i] = src[i];
}

cudaError_t small_copy(void *dst, const void *src, size_t bytes, cudaStream_t stream)   // bytes rounded up to 8: buffers have the slack
{
	size_t const n_words = (bytes + 7) / 8;
	if (!n_words) return cudaSuccess;
	unsigned const blocks = (unsigned) std::min <size_t>((n_words + 255) / 256, 64);
	small_copy_kernel<<<blocks, 256, 0, stream>>>((unsigned long long *) dst, (const unsigned long long *) src, n_words);
	return cudaGetLastError();
}

int fetch_counters(pv3_ctx *ctx, cudaStream_t stream)
{
	PV3_CUDA(ctx, small_copy(ctx->h_counters, ctx->counters.p, sizeof(Counters), stream));
	ctx->timing.launches += 1;
	PV3_CUDA(ctx, cudaStreamSynchronize(stream));
	return PV3_OK;
}

// The kernel pipeline on device-resident buffers. Events: 0 start, 1 after K1, 2 after K3, 3 after K4.
// Host-side description of where a chunk's QUAL lives when it has NOT been uploaded (lazy QUAL): the rewrite kernel never
// reads base qualities, only the realignment of the (few) indel runs does, so the host path uploads just those windows.
struct PipeOpts {
	cudaStream_t    stream = nullptr;
	uint64_t        arena_off = 0, rr_base = 0;
	bool            lazy_qual = false;
	const uint64_t *h_qual_off = nullptr;   // host, indexed by chunk-local record, values are offsets into h_qual
	const uint8_t  *h_qual = nullptr;       // host
	uint8_t        *d_qual_chunk = nullptr; // device buffer for the whole chunk's QUAL (bulk fallback)
	uint64_t        chunk_q0 = 0, chunk_q1 = 0;
};

int run_pipeline(pv3_ctx *ctx, const pv3_batch_in *in, pv3_batch_out *out, PipeOpts const &opt)
{
	cudaStream_t const stream = opt.stream ? opt.stream : ctx->stream;
	out->cig_used = out->rr_used = out->n_flagged = out->n_realigned = out->dp_cells = 0;
	ctx->timing.ms_rewrite = ctx->timing.ms_realign = ctx->timing.ms_splice = ctx->timing.ms_total = ctx->timing.ms_rewrite_main = 0;
	ctx->timing.launches = 0; ctx->timing.n_deferred = 0;
	if (!in->n) return PV3_OK;
	PV3_CUDA(ctx, ctx->jobs.ensure((out->rr_cap ? out->rr_cap : 1) * sizeof(Job)));
	PV3_CUDA(ctx, ctx->deferred.ensure(in->n * 8));
	PV3_CUDA(ctx, ctx->complex_list.ensure(in->n * 8));
	PV3_CUDA(ctx, ctx->complex_ce.ensure(in->n * 16));
	K1Params P = make_k1(ctx, in, out);
	P.deferred = ctx->deferred.as <uint64_t>();
	P.complex_list = ctx->complex_list.as <uint64_t>();
	P.complex_ce = ctx->complex_ce.as <uint64_t>();
	P.arena_off = opt.arena_off; P.rr_base = opt.rr_base;
	PV3_CUDA(ctx, cudaMemsetAsync(ctx->counters.p, 0, sizeof(Counters), stream));
	PV3_CUDA(ctx, cudaEventRecord(ctx->ev[0], stream));
	int st = PV3_OK;
	if (ctx->kernel_path != 1) {
		// short records: one thread each; whatever exceeds its fixed quotas comes back in the deferred list
		{
			// path 0: the run-walk kernel (v13); path 2: the word-parallel kernel (v14). PV3_K1_KERNEL=word|thread overrides path 0
			// (A/B runs of bench.py); both are parity-tested.
			const char *e = getenv("PV3_K1_KERNEL");
			bool const word = ctx->kernel_path == 2 || (ctx->kernel_path == 0 && e && e[0] == 'w');
			if (word) launch_k1_word(P, ctx->n_sms, stream);
			else launch_k1_thread(P, ctx->n_sms, ctx->target_code_bytes, stream);
		}
		PV3_CUDA(ctx, cudaGetLastError());
		PV3_CUDA(ctx, cudaEventRecord(ctx->ev[4], stream));
		ctx->timing.launches += 1;
		if (ctx->many_deferred) {
			// the previous batch deferred a large share (long reads): read the count and size the general kernel for it
			st = fetch_counters(ctx, stream);
			if (st) return st;
			if (ctx->h_counters->n_deferred && !ctx->h_counters->overflow) {
				K1Params P2 = P;
				P2.rec_list = ctx->deferred.as <uint64_t>();
				P2.n_list = ctx->h_counters->n_deferred;
				launch_k1(P2, ctx->n_sms, stream);
				PV3_CUDA(ctx, cudaGetLastError());
				ctx->timing.launches += 1;
			}
		} else {
			// usual case, a handful of deferred records: the general kernel reads the count on the device (no host round trip)
			K1Params P2 = P;
			P2.rec_list = ctx->deferred.as <uint64_t>();
			P2.n_list_dev = &ctx->counters.as <Counters>()->n_deferred;
			launch_k1(P2, ctx->n_sms, stream);
			PV3_CUDA(ctx, cudaGetLastError());
			ctx->timing.launches += 1;
		}
		PV3_CUDA(ctx, cudaEventRecord(ctx->ev[1], stream));
		st = fetch_counters(ctx, stream);
		if (st) return st;
		ctx->many_deferred = ctx->h_counters->n_deferred * 16 > in->n;
	} else {
		launch_k1(P, ctx->n_sms, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		PV3_CUDA(ctx, cudaEventRecord(ctx->ev[1], stream));
		ctx->timing.launches += 1;
		st = fetch_counters(ctx, stream);
		if (st) return st;
	}
	Counters c = *ctx->h_counters;
	out->rr_used = c.job_cursor;
	out->cig_used = c.arena_cursor + c.dp_path;   // upper bound of what the splice kernel may still append
	if (c.overflow) {
		ctx->total_launches += ctx->timing.launches;
		return fail(ctx, PV3_ERR_CAPACITY, "output capacity too small: need cig_cap >= %llu, rr_cap >= %llu",
			(unsigned long long) out->cig_used, (unsigned long long) out->rr_used);
	}
	uint64_t const n_jobs = c.job_cursor;
	bool ran_dp = false;
	if (n_jobs) {
		if (c.arena_cursor + c.dp_path > out->cig_cap) {
			ctx->total_launches += ctx->timing.launches;
			return fail(ctx, PV3_ERR_CAPACITY, "output capacity too small: need cig_cap >= %llu", (unsigned long long) out->cig_used);
		}
		K3Params K;
		memset(&K, 0, sizeof K);
		K.k1 = P; K.n_jobs = n_jobs;
		uint64_t grid_warps = k3_default_grid_warps(ctx->n_sms);
		PV3_CUDA(ctx, ctx->dp_arena.ensure((c.dp_path + 1) * 4));
		K.dp_arena = ctx->dp_arena.as <uint32_t>(); K.dp_arena_cap = c.dp_path;
		if (c.dp_max_pool > k3_smem_pool_bytes()) {
			size_t const slot = align_up((size_t) c.dp_max_pool, 256);
			uint64_t const fit = POOL_BUDGET / slot ? POOL_BUDGET / slot : 1;
			if (grid_warps > fit) grid_warps = fit;
			if (grid_warps > n_jobs) grid_warps = n_jobs;
			grid_warps = (grid_warps + 7) / 8 * 8;
			PV3_CUDA(ctx, ctx->pool.ensure(slot * grid_warps));
			K.pool = ctx->pool.as <uint8_t>(); K.pool_slot_bytes = slot;
		}
		memcpy(K.lut, ctx->lut, sizeof K.lut);
		if (opt.lazy_qual) {
			if (n_jobs * 16 > in->n) {
				// many runs (indel-heavy input): one bulk upload of the chunk's QUAL beats gathering windows
				PV3_CUDA(ctx, cudaMemcpyAsync(opt.d_qual_chunk, opt.h_qual + opt.chunk_q0, opt.chunk_q1 - opt.chunk_q0, cudaMemcpyHostToDevice, stream));
			} else {
				// gather the query windows of the runs on the host and upload only those bytes
				if (ctx->h_jobs_cap < n_jobs) {
					if (ctx->h_jobs) cudaFreeHost(ctx->h_jobs);
					if (ctx->h_job_qual_off) cudaFreeHost(ctx->h_job_qual_off);
					ctx->h_jobs_cap = n_jobs * 2 + 1024;
					PV3_CUDA(ctx, cudaMallocHost((void **) &ctx->h_jobs, ctx->h_jobs_cap * sizeof(Job)));
					PV3_CUDA(ctx, cudaMallocHost((void **) &ctx->h_job_qual_off, ctx->h_jobs_cap * 8));
				}
				PV3_CUDA(ctx, small_copy(ctx->h_jobs, ctx->jobs.p, n_jobs * sizeof(Job), stream));
				ctx->timing.launches += 1;
				PV3_CUDA(ctx, cudaStreamSynchronize(stream));
				uint64_t need = 0;
				for (uint64_t j = 0; j < n_jobs; ++j) need += ctx->h_jobs[j].qry_len;
				if (ctx->h_packed_cap < need + 16) {
					if (ctx->h_packed) cudaFreeHost(ctx->h_packed);
					ctx->h_packed_cap = need * 2 + 4096;
					PV3_CUDA(ctx, cudaMallocHost((void **) &ctx->h_packed, ctx->h_packed_cap));
				}
				uint64_t w = 0;
				for (uint64_t j = 0; j < n_jobs; ++j) {
					Job const &job = ctx->h_jobs[j];
					ctx->h_job_qual_off[j] = w;
					uint64_t const q0 = opt.h_qual_off[job.rec], q1 = opt.h_qual_off[job.rec + 1];
					if (q1 > q0) { memcpy(ctx->h_packed + w, opt.h_qual + q0 + job.qry_loc, job.qry_len); w += job.qry_len; }
				}
				PV3_CUDA(ctx, ctx->d_packed.ensure(w + 16));
				PV3_CUDA(ctx, ctx->d_job_qual_off.ensure(n_jobs * 8));
				PV3_CUDA(ctx, small_copy(ctx->d_packed.p, ctx->h_packed, w, stream));
				PV3_CUDA(ctx, small_copy(ctx->d_job_qual_off.p, ctx->h_job_qual_off, n_jobs * 8, stream));
				ctx->timing.launches += (w ? 1 : 0) + 1;
				K.packed_qual = ctx->d_packed.as <uint8_t>();
				K.job_qual_off = ctx->d_job_qual_off.as <uint64_t>();
			}
		}
		// small runs (the usual case: a few bases wide): one thread per run; the rest: one warp per run
		// (a few thousand runs: the warp kernel alone finishes within its launch latency, a second launch only adds to it)
		bool const use_thread_kernel = n_jobs > c.n_big_jobs && n_jobs - c.n_big_jobs >= ctx->dp_thread_min;
		if (use_thread_kernel) {
			if (c.n_big_jobs) {
				PV3_CUDA(ctx, ctx->big_list.ensure(c.n_big_jobs * 8));
				K.big_list = ctx->big_list.as <uint64_t>(); K.n_big = c.n_big_jobs;
			}
			launch_k3_thread(K, stream);
			PV3_CUDA(ctx, cudaGetLastError());
			ctx->timing.launches += 1;
		}
		if (c.n_big_jobs || !use_thread_kernel) {
			launch_k3(K, ctx->n_sms, grid_warps, stream);
			PV3_CUDA(ctx, cudaGetLastError());
			ctx->timing.launches += 1;
		}
		PV3_CUDA(ctx, cudaEventRecord(ctx->ev[2], stream));
		launch_k4(P, n_jobs, K.dp_arena, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		ctx->timing.launches += 1;
		PV3_CUDA(ctx, cudaEventRecord(ctx->ev[3], stream));
		st = fetch_counters(ctx, stream);
		if (st) return st;
		c = *ctx->h_counters;
		ran_dp = true;
		if (c.overflow) {
			ctx->total_launches += ctx->timing.launches;
			return fail(ctx, PV3_ERR_CAPACITY, "output capacity too small (splice): need cig_cap >= %llu", (unsigned long long) c.arena_cursor);
		}
	}
	out->cig_used = c.arena_cursor;
	out->rr_used = c.job_cursor;
	out->n_flagged = c.n_flagged;
	out->n_realigned = c.n_realigned;
	out->dp_cells = c.dp_cells;
	cudaEventElapsedTime(&ctx->timing.ms_rewrite, ctx->ev[0], ctx->ev[1]);
	if (ctx->kernel_path != 1) cudaEventElapsedTime(&ctx->timing.ms_rewrite_main, ctx->ev[0], ctx->ev[4]);
	else ctx->timing.ms_rewrite_main = ctx->timing.ms_rewrite;
	ctx->timing.n_deferred = c.n_deferred;
	if (ran_dp) {
		cudaEventElapsedTime(&ctx->timing.ms_realign, ctx->ev[1], ctx->ev[2]);
		cudaEventElapsedTime(&ctx->timing.ms_splice, ctx->ev[2], ctx->ev[3]);
		cudaEventElapsedTime(&ctx->timing.ms_total, ctx->ev[0], ctx->ev[3]);
	} else ctx->timing.ms_total = ctx->timing.ms_rewrite;
	ctx->total_launches += ctx->timing.launches;
	return PV3_OK;
}

int check_ready(pv3_ctx *ctx, const char *who)
{
	if (!ctx) return PV3_ERR_INVALID;
	if (!ctx->committed) return fail(ctx, PV3_ERR_STATE, "%s: pv3_index_commit has not been called", who);
	cudaError_t e = cudaSetDevice(ctx->device);
	if (e != cudaSuccess) return fail(ctx, PV3_ERR_CUDA, "%s: cudaSetDevice failed: %s", who, cudaGetErrorString(e));
	return PV3_OK;
}

} // namespace

extern "C" {

int pv3_project_device(pv3_ctx *ctx, const pv3_batch_in *in, pv3_batch_out *out)
{
	int st = check_ready(ctx, "pv3_project_device");
	if (st) return st;
	if (!in || !out) return fail(ctx, PV3_ERR_INVALID, "pv3_project_device: NULL batch");
	ctx->timing.ms_h2d = ctx->timing.ms_d2h = 0;
	return run_pipeline(ctx, in, out, PipeOpts());
}

int pv3_project(pv3_ctx *ctx, const pv3_batch_in *in, pv3_batch_out *out)
{
	// Host pipeline (replaces the 16-slot SPSC ring of 4-record tasks, include/panvc3/spsc_queue.hh): the batch is cut
	// into chunks; chunk i+1's H2D copy runs on the other stream while chunk i's kernels and chunk i-1's D2H copy run,
	// so PCIe transfers in both directions overlap the compute. QUAL is uploaded lazily (see PipeOpts).
	int st = check_ready(ctx, "pv3_project");
	if (st) return st;
	if (!in || !out) return fail(ctx, PV3_ERR_INVALID, "pv3_project: NULL batch");
	uint64_t const n = in->n;
	out->cig_used = out->rr_used = out->n_flagged = out->n_realigned = out->dp_cells = 0;
	memset(&ctx->timing, 0, sizeof ctx->timing);
	if (!n) return PV3_OK;

	bool const own_streams = (ctx->stream == ctx->own_stream);
	uint64_t n_chunks = own_streams ? (n + ctx->chunk_records / 2) / ctx->chunk_records : 1;
	if (n_chunks < 1) n_chunks = 1;
	if (n_chunks > 64) n_chunks = 64;
	if (own_streams && !ctx->stream2) PV3_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
	cudaStream_t const streams[2] = { ctx->stream, own_streams ? ctx->stream2 : ctx->stream };
	for (int k = 0; k < 2; ++k) for (int e = 0; e < 4; ++e)
		if (!ctx->cev[k][e]) PV3_CUDA(ctx, cudaEventCreate(&ctx->cev[k][e]));

	// Chunk cuts (tile-aligned). Only the first chunk's upload and the last chunk's kernels and download are not hidden
	// behind other chunks' transfers, so those two chunks are small and the ones between them large.
	uint64_t cuts[67];
	{
		uint64_t const edge = (n_chunks >= 2 && ctx->chunk_edge && n > 4 * ctx->chunk_edge) ? ctx->chunk_edge : 0;
		if (edge) {
			uint64_t n_mid = (n - 2 * edge + ctx->chunk_records / 2) / ctx->chunk_records;
			if (n_mid < 1) n_mid = 1;
			if (n_mid > 62) n_mid = 62;
			n_chunks = n_mid + 2;
			cuts[0] = 0;
			for (uint64_t c = 0; c <= n_mid; ++c) cuts[c + 1] = (edge + (n - 2 * edge) * c / n_mid) / 128 * 128;
			cuts[n_chunks] = n;
		} else {
			for (uint64_t c = 0; c < n_chunks; ++c) cuts[c] = (n * c / n_chunks) / 128 * 128;
			cuts[n_chunks] = n;
		}
	}
	auto chunk_begin = [&](uint64_t c) { return cuts[c]; };
	auto chunk_end = [&](uint64_t c) { return cuts[c + 1]; };

	// H2D of chunk c into buffer set c % 2 (QUAL bytes are not copied)
	auto enqueue_h2d = [&](uint64_t c) -> int {
		HostSet &hs = ctx->sets[c % 2];
		cudaStream_t const stream = streams[c % 2];
		uint64_t const b = chunk_begin(c), e = chunk_end(c), m = e - b;
		uint64_t const s0 = in->seq_off[b], s1 = in->seq_off[e], c0 = in->cig_off[b], c1 = in->cig_off[e];
		struct Copy { DevBuf *buf; const void *src; size_t bytes; };
		Copy const copies[] = {
			{ &hs.in_src_row, in->src_row + b, m * 4 }, { &hs.in_dst_row, in->dst_row + b, m * 4 }, { &hs.in_dst_ref, in->dst_ref + b, m * 4 },
			{ &hs.in_src_pos, in->src_pos + b, m * 8 }, { &hs.in_seq_off, in->seq_off + b, (m + 1) * 8 }, { &hs.in_seq, in->seq + s0, s1 - s0 },
			{ &hs.in_qual_off, in->qual_off ? in->qual_off + b : nullptr, in->qual_off ? (m + 1) * 8 : 0 },
			{ &hs.in_cig_off, in->cig_off + b, (m + 1) * 8 }, { &hs.in_cig, in->cig + c0, (c1 - c0) * 4 },
			{ &hs.in_seq_len, in->seq_len ? in->seq_len + b : nullptr, in->seq_len ? m * 4 : 0 },
		};
		for (auto const &cp : copies) PV3_CUDA(ctx, cp.buf->ensure((cp.bytes ? cp.bytes : 16) + 32));
		if (in->qual_off) PV3_CUDA(ctx, hs.in_qual.ensure(in->qual_off[e] - in->qual_off[b] + 32));
		PV3_CUDA(ctx, cudaEventRecord(ctx->cev[c % 2][0], stream));
		for (auto const &cp : copies)
			if (cp.bytes) PV3_CUDA(ctx, cudaMemcpyAsync(cp.buf->p, cp.src, cp.bytes, cudaMemcpyHostToDevice, stream));
		PV3_CUDA(ctx, cudaEventRecord(ctx->cev[c % 2][1], stream));
		return PV3_OK;
	};

	uint64_t arena_off = 0, rr_base = 0;
	bool host_short = false;
	float ms_h2d = 0, ms_d2h = 0, ms_k[4] = { 0, 0, 0, 0 };
	uint64_t launches = 0, n_deferred = 0;
	float ms_main = 0;
	int pending_d2h[2] = { 0, 0 };

	bool const trace = getenv("PV3_TRACE") != nullptr;     // timeline of the chunks on stderr, ms since the call began
	if (trace) cudaEventRecord(ctx->ev[7], streams[0]);
	auto since = [&](cudaEvent_t e) { float t = -1; cudaEventElapsedTime(&t, ctx->ev[7], e); return t; };
	auto harvest_copy_times = [&](int k) {
		float t = 0;
		if (trace) {
			fprintf(stderr, "[pv3 trace] set %d: h2d %.3f-%.3f  kernels %.3f-%.3f", k, since(ctx->cev[k][0]), since(ctx->cev[k][1]), since(ctx->ev[0]), since(ctx->ev[1]));
			if (pending_d2h[k]) fprintf(stderr, "  | previous d2h of this set %.3f-%.3f", since(ctx->cev[k][2]), since(ctx->cev[k][3]));
			fprintf(stderr, "\n");
		}
		if (cudaEventElapsedTime(&t, ctx->cev[k][0], ctx->cev[k][1]) == cudaSuccess) ms_h2d += t;
		if (pending_d2h[k] && cudaEventElapsedTime(&t, ctx->cev[k][2], ctx->cev[k][3]) == cudaSuccess) ms_d2h += t;
		pending_d2h[k] = 0;
	};

	st = enqueue_h2d(0);
	if (st) return st;
	for (uint64_t c = 0; c < n_chunks; ++c) {
		int const k = (int) (c % 2);
		HostSet &hs = ctx->sets[k];
		cudaStream_t const stream = streams[k];
		uint64_t const b = chunk_begin(c), e = chunk_end(c), m = e - b;
		if (c + 1 < n_chunks) {
			// the other set's previous D2H (chunk c-1) only reads its output buffers; its input buffers are free
			if (own_streams) { st = enqueue_h2d(c + 1); if (st) return st; }
		}
		if (!m) { if (!own_streams && c + 1 < n_chunks) { st = enqueue_h2d(c + 1); if (st) return st; } continue; }

		pv3_batch_in din = *in;
		din.n = m;
		din.src_row = hs.in_src_row.as <uint32_t>(); din.dst_row = hs.in_dst_row.as <uint32_t>(); din.dst_ref = hs.in_dst_ref.as <uint32_t>();
		din.src_pos = hs.in_src_pos.as <uint64_t>(); din.seq_off = hs.in_seq_off.as <uint64_t>();
		din.seq = hs.in_seq.as <uint8_t>() - in->seq_off[b];              // offsets in seq_off stay batch-global
		din.qual_off = in->qual_off ? hs.in_qual_off.as <uint64_t>() : nullptr;
		din.qual = in->qual_off ? hs.in_qual.as <uint8_t>() - in->qual_off[b] : nullptr;
		din.cig_off = hs.in_cig_off.as <uint64_t>();
		din.cig = hs.in_cig.as <uint32_t>() - in->cig_off[b];
		din.seq_len = in->seq_len ? hs.in_seq_len.as <uint32_t>() : nullptr;

		PV3_CUDA(ctx, hs.out_dst_pos.ensure(m * 8)); PV3_CUDA(ctx, hs.out_flags.ensure(m * 4));
		PV3_CUDA(ctx, hs.out_cig_off.ensure(m * 8)); PV3_CUDA(ctx, hs.out_cig_len.ensure(m * 4));
		PV3_CUDA(ctx, hs.out_rr_off.ensure(m * 8)); PV3_CUDA(ctx, hs.out_rr_len.ensure(m * 4));
		uint64_t arena_cap = (in->cig_off[e] - in->cig_off[b]) * 3 + m * 8 + 1024, rr_cap = m / 8 + 1024;
		if (arena_cap < hs.arena_cap_hint) arena_cap = hs.arena_cap_hint;
		if (rr_cap < hs.rr_cap_hint) rr_cap = hs.rr_cap_hint;

		PipeOpts opt;
		opt.stream = stream; opt.arena_off = arena_off; opt.rr_base = rr_base;
		if (in->qual_off) {
			opt.lazy_qual = true; opt.h_qual_off = in->qual_off + b; opt.h_qual = in->qual;
			opt.d_qual_chunk = hs.in_qual.as <uint8_t>(); opt.chunk_q0 = in->qual_off[b]; opt.chunk_q1 = in->qual_off[e];
		}
		pv3_batch_out dout;
		for (int attempt = 0; ; ++attempt) {
			PV3_CUDA(ctx, hs.out_arena.ensure(arena_cap * 4));
			PV3_CUDA(ctx, hs.out_rr.ensure(rr_cap * 32));
			memset(&dout, 0, sizeof dout);
			dout.dst_pos = hs.out_dst_pos.as <uint64_t>(); dout.flags = hs.out_flags.as <uint32_t>();
			dout.cig_off = hs.out_cig_off.as <uint64_t>(); dout.cig_len = hs.out_cig_len.as <uint32_t>();
			dout.cig = hs.out_arena.as <uint32_t>(); dout.cig_cap = arena_cap;
			dout.rr_off = hs.out_rr_off.as <uint64_t>(); dout.rr_len = hs.out_rr_len.as <uint32_t>();
			dout.rr = hs.out_rr.as <uint64_t>(); dout.rr_cap = rr_cap;
			st = run_pipeline(ctx, &din, &dout, opt);
			launches += ctx->timing.launches;
			if (st == PV3_ERR_CAPACITY && attempt < 6) {
				// grow from what the kernels reported and run the chunk again (results of the failed attempt are discarded)
				if (dout.cig_used + 1024 > arena_cap) arena_cap = dout.cig_used * 2 + 1024; else arena_cap *= 2;
				if (dout.rr_used + 64 > rr_cap) rr_cap = dout.rr_used * 2 + 64;
				continue;
			}
			if (st) return st;
			break;
		}
		hs.arena_cap_hint = arena_cap; hs.rr_cap_hint = rr_cap;
		ms_k[0] += ctx->timing.ms_rewrite; ms_k[1] += ctx->timing.ms_realign; ms_k[2] += ctx->timing.ms_splice; ms_k[3] += ctx->timing.ms_total;
		n_deferred += ctx->timing.n_deferred; ms_main += ctx->timing.ms_rewrite_main;
		harvest_copy_times(k);     // run_pipeline synchronized this stream: the events of its H2D (and older D2H) are complete

		out->n_flagged += dout.n_flagged; out->n_realigned += dout.n_realigned; out->dp_cells += dout.dp_cells;
		if (arena_off + dout.cig_used > out->cig_cap || rr_base + dout.rr_used > out->rr_cap) host_short = true;
		if (!host_short) {
			PV3_CUDA(ctx, cudaEventRecord(ctx->cev[k][2], stream));
			struct Back { void *dst; const void *src; size_t bytes; };
			Back const backs[] = {
				{ out->dst_pos + b, dout.dst_pos, m * 8 }, { out->flags + b, dout.flags, m * 4 }, { out->cig_off + b, dout.cig_off, m * 8 },
				{ out->cig_len + b, dout.cig_len, m * 4 }, { out->rr_off + b, dout.rr_off, m * 8 }, { out->rr_len + b, dout.rr_len, m * 4 },
				{ out->cig + arena_off, dout.cig, dout.cig_used * 4 }, { out->rr + 4 * rr_base, dout.rr, dout.rr_used * 32 },
			};
			for (auto const &bk : backs)
				if (bk.bytes) PV3_CUDA(ctx, cudaMemcpyAsync(bk.dst, bk.src, bk.bytes, cudaMemcpyDeviceToHost, stream));
			PV3_CUDA(ctx, cudaEventRecord(ctx->cev[k][3], stream));
			pending_d2h[k] = 1;
		}
		arena_off += dout.cig_used; rr_base += dout.rr_used;
		if (!own_streams && c + 1 < n_chunks) { st = enqueue_h2d(c + 1); if (st) return st; }
	}
	PV3_CUDA(ctx, cudaStreamSynchronize(streams[0]));
	if (own_streams) PV3_CUDA(ctx, cudaStreamSynchronize(streams[1]));
	for (int k = 0; k < 2; ++k) if (pending_d2h[k]) {
		float t = 0; if (cudaEventElapsedTime(&t, ctx->cev[k][2], ctx->cev[k][3]) == cudaSuccess) ms_d2h += t;
		if (trace) fprintf(stderr, "[pv3 trace] set %d: last d2h %.3f-%.3f\n", k, since(ctx->cev[k][2]), since(ctx->cev[k][3]));
	}

	out->cig_used = arena_off; out->rr_used = rr_base;
	ctx->timing.ms_rewrite = ms_k[0]; ctx->timing.ms_realign = ms_k[1]; ctx->timing.ms_splice = ms_k[2]; ctx->timing.ms_total = ms_k[3];
	ctx->timing.n_deferred = n_deferred; ctx->timing.ms_rewrite_main = ms_main;
	ctx->timing.ms_h2d = ms_h2d; ctx->timing.ms_d2h = ms_d2h;
	ctx->timing.launches = launches;
	if (host_short)
		return fail(ctx, PV3_ERR_CAPACITY, "host output capacity too small: need cig_cap >= %llu, rr_cap >= %llu",
			(unsigned long long) arena_off, (unsigned long long) rr_base);
	return PV3_OK;
}

int pv3_project_positions(pv3_ctx *ctx, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row, const uint64_t *pos, uint64_t *out_pos)
{
	int st = check_ready(ctx, "pv3_project_positions");
	if (st) return st;
	if (!n) return PV3_OK;
	if (!src_row || !dst_row || !pos || !out_pos) return fail(ctx, PV3_ERR_INVALID, "pv3_project_positions: NULL argument");
	PV3_CUDA(ctx, ctx->in_src_row.ensure(n * 4)); PV3_CUDA(ctx, ctx->in_dst_row.ensure(n * 4));
	PV3_CUDA(ctx, ctx->in_src_pos.ensure(n * 8)); PV3_CUDA(ctx, ctx->out_dst_pos.ensure(n * 8));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->in_src_row.p, src_row, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->in_dst_row.p, dst_row, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->in_src_pos.p, pos, n * 8, cudaMemcpyHostToDevice, ctx->stream));
	launch_positions(ctx->row_desc.as <RowDesc>(), (uint32_t) ctx->rows.size(), n, ctx->in_src_row.as <uint32_t>(),
		ctx->in_dst_row.as <uint32_t>(), ctx->in_src_pos.as <uint64_t>(), ctx->out_dst_pos.as <uint64_t>(), ctx->stream);
	PV3_CUDA(ctx, cudaGetLastError());
	ctx->total_launches += 1;
	PV3_CUDA(ctx, cudaMemcpyAsync(out_pos, ctx->out_dst_pos.p, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
	PV3_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PV3_OK;
}

int pv3_index_query(pv3_ctx *ctx, int op, uint64_t n, const uint32_t *row, const uint64_t *x, uint64_t *out)
{
	int st = check_ready(ctx, "pv3_index_query");
	if (st) return st;
	if (op < 0 || op > 2) return fail(ctx, PV3_ERR_INVALID, "pv3_index_query: op must be 0, 1 or 2");
	if (!n) return PV3_OK;
	PV3_CUDA(ctx, ctx->in_src_row.ensure(n * 4)); PV3_CUDA(ctx, ctx->in_src_pos.ensure(n * 8)); PV3_CUDA(ctx, ctx->out_dst_pos.ensure(n * 8));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->in_src_row.p, row, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->in_src_pos.p, x, n * 8, cudaMemcpyHostToDevice, ctx->stream));
	launch_index_query(ctx->row_desc.as <RowDesc>(), (uint32_t) ctx->rows.size(), op, n, ctx->in_src_row.as <uint32_t>(),
		ctx->in_src_pos.as <uint64_t>(), ctx->out_dst_pos.as <uint64_t>(), ctx->stream);
	PV3_CUDA(ctx, cudaGetLastError());
	ctx->total_launches += 1;
	PV3_CUDA(ctx, cudaMemcpyAsync(out, ctx->out_dst_pos.p, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
	PV3_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PV3_OK;
}

// index-msa's per-sequence work (index-msa/index_vector_builder.cc:28-50) on device-resident text
int pv3_msa_row_extract_device(pv3_ctx *ctx, const char *text, uint64_t n, uint64_t *gap_words, char *ungapped, uint64_t *n_ungapped)
{
	if (!ctx) return PV3_ERR_INVALID;
	cudaSetDevice(ctx->device);
	if (n && (!text || !gap_words)) return fail(ctx, PV3_ERR_INVALID, "pv3_msa_row_extract_device: NULL argument");
	if ((uintptr_t) text % 16 || (uintptr_t) gap_words % 8 || (uintptr_t) ungapped % 4)
		return fail(ctx, PV3_ERR_INVALID, "pv3_msa_row_extract_device: text must be 16-byte, gap_words 8-byte, ungapped 4-byte aligned");
	PV3_CUDA(ctx, ctx->ex_status.ensure(msa_extract_scratch_words(n) * 8));
	unsigned launches = 0;
	PV3_CUDA(ctx, cudaEventRecord(ctx->ev[5], ctx->stream));
	PV3_CUDA(ctx, launch_msa_extract(text, n, gap_words, ungapped, ctx->ex_status.as <unsigned long long>(), ctx->stream, &launches));
	PV3_CUDA(ctx, cudaEventRecord(ctx->ev[6], ctx->stream));
	ctx->total_launches += launches;
	unsigned long long total = 0;
	PV3_CUDA(ctx, cudaMemcpyAsync(&total, ctx->ex_status.as <unsigned long long>() + msa_extract_total_word(n), 8, cudaMemcpyDeviceToHost, ctx->stream));
	PV3_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	cudaEventElapsedTime(&ctx->timing.ms_msa_extract, ctx->ev[5], ctx->ev[6]);
	if (n_ungapped) *n_ungapped = ungapped ? total : 0;
	return PV3_OK;
}

int pv3_msa_row_extract(pv3_ctx *ctx, const char *text, uint64_t n, uint64_t *gap_words, char *ungapped, uint64_t *n_ungapped)
{
	if (!ctx) return PV3_ERR_INVALID;
	cudaSetDevice(ctx->device);
	if (n && (!text || !gap_words)) return fail(ctx, PV3_ERR_INVALID, "pv3_msa_row_extract: NULL argument");
	if (n_ungapped) *n_ungapped = 0;
	if (!n) return PV3_OK;
	uint64_t const n_words = (n + 63) / 64;
	PV3_CUDA(ctx, ctx->ex_text.ensure(n + 16)); PV3_CUDA(ctx, ctx->ex_words.ensure(n_words * 8));
	if (ungapped) PV3_CUDA(ctx, ctx->ex_seq.ensure(n + 16));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->ex_text.p, text, n, cudaMemcpyHostToDevice, ctx->stream));
	uint64_t total = 0;
	int st = pv3_msa_row_extract_device(ctx, ctx->ex_text.as <char>(), n, ctx->ex_words.as <uint64_t>(), ungapped ? ctx->ex_seq.as <char>() : nullptr, &total);
	if (st) return st;
	PV3_CUDA(ctx, cudaMemcpyAsync(gap_words, ctx->ex_words.p, n_words * 8, cudaMemcpyDeviceToHost, ctx->stream));
	if (ungapped && total) PV3_CUDA(ctx, cudaMemcpyAsync(ungapped, ctx->ex_seq.p, total, cudaMemcpyDeviceToHost, ctx->stream));
	PV3_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	if (n_ungapped) *n_ungapped = total;
	return PV3_OK;
}

} // extern "C"

namespace {

// Host buffers in, text + offsets out. P holds HOST pointers on entry (cig_len may be null, ref_id null ⇒ CIGAR text only).
int format_text(pv3_ctx *ctx, const char *who, OaParams P, uint64_t n_cig_words, uint64_t names_bytes,
	uint64_t *text_off, char *text, uint64_t text_cap, uint64_t *text_used)
{
	cudaSetDevice(ctx->device);
	if (text_used) *text_used = 0;
	if (!text_off) return fail(ctx, PV3_ERR_INVALID, "%s: NULL text_off", who);
	text_off[0] = 0;
	uint64_t const n = P.n;
	if (!n) return PV3_OK;
	cudaStream_t const st = ctx->stream;
	int slot = 0;
	auto up = [&](const void *src, size_t bytes, const void **dev) -> cudaError_t {
		DevBuf &b = ctx->an[slot++];
		cudaError_t e = b.ensure(bytes + 16);
		if (e != cudaSuccess) return e;
		*dev = b.p;
		return bytes ? cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, st) : cudaSuccess;
	};
	auto scratch = [&](size_t bytes, void **dev) -> cudaError_t {
		DevBuf &b = ctx->an[slot++];
		cudaError_t e = b.ensure(bytes + 16);
		*dev = b.p;
		return e;
	};
	OaParams D = P;
	if (P.ref_id) {
		PV3_CUDA(ctx, up(P.ref_id, n * 4, (const void **) &D.ref_id));
		PV3_CUDA(ctx, up(P.pos, n * 8, (const void **) &D.pos));
		PV3_CUDA(ctx, up(P.flag, n * 2, (const void **) &D.flag));
		PV3_CUDA(ctx, up(P.mapq, n, (const void **) &D.mapq));
		if (P.has_nm) {
			PV3_CUDA(ctx, up(P.nm, n * 4, (const void **) &D.nm));
			PV3_CUDA(ctx, up(P.has_nm, n, (const void **) &D.has_nm));
		} else slot += 2;
		PV3_CUDA(ctx, up(P.name_off, ((size_t) P.n_names + 1) * 8, (const void **) &D.name_off));
		PV3_CUDA(ctx, up(P.names, names_bytes, (const void **) &D.names));
	} else slot += 8;
	PV3_CUDA(ctx, up(P.cig_off, (n + (P.cig_len ? 0 : 1)) * 8, (const void **) &D.cig_off));
	if (P.cig_len) PV3_CUDA(ctx, up(P.cig_len, n * 4, (const void **) &D.cig_len)); else ++slot;
	PV3_CUDA(ctx, up(P.cig, n_cig_words * 4, (const void **) &D.cig));
	unsigned const ctas = oa_ctas(n);
	PV3_CUDA(ctx, scratch(n * 4, (void **) &D.len));
	PV3_CUDA(ctx, scratch(((size_t) ctas + 4) * 4, (void **) &D.cta_len));
	PV3_CUDA(ctx, scratch(((size_t) ctas + 1) * 8, (void **) &D.cta_off));
	PV3_CUDA(ctx, scratch((n + 1) * 8, (void **) &D.text_off));
	launch_oa_lengths(D, st);
	launch_scan_u32(D.cta_len, ctas, D.cta_off, st);
	PV3_CUDA(ctx, cudaGetLastError());
	unsigned long long total = 0;
	PV3_CUDA(ctx, cudaMemcpyAsync(&total, D.cta_off + ctas, 8, cudaMemcpyDeviceToHost, st));
	PV3_CUDA(ctx, cudaStreamSynchronize(st));
	ctx->total_launches += 2;
	if (text_used) *text_used = total;
	if (total > text_cap) return fail(ctx, PV3_ERR_CAPACITY, "%s: text needs %llu bytes, capacity is %llu", who, total, (unsigned long long) text_cap);
	if (total && !text) return fail(ctx, PV3_ERR_INVALID, "%s: NULL text", who);
	PV3_CUDA(ctx, scratch(total, (void **) &D.text));
	launch_oa_write(D, st);
	PV3_CUDA(ctx, cudaGetLastError());
	ctx->total_launches += 1;
	PV3_CUDA(ctx, cudaMemcpyAsync(text_off, D.text_off, (n + 1) * 8, cudaMemcpyDeviceToHost, st));
	if (total) PV3_CUDA(ctx, cudaMemcpyAsync(text, D.text, total, cudaMemcpyDeviceToHost, st));
	PV3_CUDA(ctx, cudaStreamSynchronize(st));
	return PV3_OK;
}

} // namespace

extern "C" {

int pv3_format_oa(pv3_ctx *ctx, const pv3_oa_in *in, uint64_t *text_off, char *text, uint64_t text_cap, uint64_t *text_used)
{
	if (!ctx) return PV3_ERR_INVALID;
	if (!in) return fail(ctx, PV3_ERR_INVALID, "pv3_format_oa: NULL argument");
	if (in->n && (!in->ref_id || !in->pos || !in->flag || !in->mapq || !in->cig_off || (in->has_nm && !in->nm) || !in->name_off))
		return fail(ctx, PV3_ERR_INVALID, "pv3_format_oa: NULL array");
	OaParams P;
	memset(&P, 0, sizeof P);
	P.n = in->n; P.ref_id = in->ref_id; P.pos = in->pos; P.flag = in->flag; P.mapq = in->mapq; P.nm = in->nm; P.has_nm = in->has_nm;
	P.cig_off = in->cig_off; P.cig = in->cig; P.n_names = in->n_names; P.name_off = in->name_off; P.names = in->names;
	uint64_t const n_cig = in->n ? in->cig_off[in->n] : 0;
	uint64_t const names_bytes = in->n_names ? in->name_off[in->n_names] : 0;
	return format_text(ctx, "pv3_format_oa", P, n_cig, names_bytes, text_off, text, text_cap, text_used);
}

// recalculate-mapq/main.cc:181-186 — cut-off at 40 like Bowtie 2; the double expression is truncated to int32
static void fill_penalties(int32_t *penalties, int32_t const min, int32_t const max)
{
	for (size_t i = 0; i < 256; ++i) penalties[i] = (int32_t) (std::min(i, size_t(40)) / double(40) * (max - min) + min);
}

int pv3_rescore_alignments(pv3_ctx *ctx, const pv3_scoring *sc, uint64_t n, const uint64_t *seq_off, const uint8_t *seq,
	const uint64_t *qual_off, const uint8_t *qual, const uint64_t *cig_off, const uint32_t *cig,
	const uint64_t *refn_off, const uint8_t *refn, double *score, uint32_t *flags)
{
	if (!ctx) return PV3_ERR_INVALID;
	cudaSetDevice(ctx->device);
	if (!n) return PV3_OK;
	if (!sc || !seq_off || !qual_off || !cig_off || !score || !flags) return fail(ctx, PV3_ERR_INVALID, "pv3_rescore_alignments: NULL argument");
	cudaStream_t const st = ctx->stream;
	int32_t lut[512];
	fill_penalties(lut, sc->min_mismatch_penalty, sc->max_mismatch_penalty);
	fill_penalties(lut + 256, sc->n_penalty, sc->n_penalty);
	int slot = 0;
	auto up = [&](const void *src, size_t bytes, const void **dev) -> cudaError_t {
		DevBuf &b = ctx->an[slot++];
		cudaError_t e = b.ensure(bytes + 16);
		if (e != cudaSuccess) return e;
		*dev = b.p;
		return bytes ? cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, st) : cudaSuccess;
	};
	RescoreParams P;
	memset(&P, 0, sizeof P);
	P.n = n; P.gap_open = sc->gap_opening_penalty; P.gap_ext = sc->gap_extension_penalty;
	const void *d_lut = nullptr;
	PV3_CUDA(ctx, up(seq_off, (n + 1) * 8, (const void **) &P.seq_off));
	PV3_CUDA(ctx, up(seq, seq_off[n], (const void **) &P.seq));
	PV3_CUDA(ctx, up(qual_off, (n + 1) * 8, (const void **) &P.qual_off));
	PV3_CUDA(ctx, up(qual, qual_off[n], (const void **) &P.qual));
	PV3_CUDA(ctx, up(cig_off, (n + 1) * 8, (const void **) &P.cig_off));
	PV3_CUDA(ctx, up(cig, cig_off[n] * 4, (const void **) &P.cig));
	if (refn_off) {
		PV3_CUDA(ctx, up(refn_off, (n + 1) * 8, (const void **) &P.refn_off));
		PV3_CUDA(ctx, up(refn, refn_off[n], (const void **) &P.refn));
	} else slot += 2;
	PV3_CUDA(ctx, up(lut, sizeof lut, &d_lut));
	P.mismatch_penalties = (const int32_t *) d_lut; P.n_penalties = (const int32_t *) d_lut + 256;
	PV3_CUDA(ctx, ctx->an[slot].ensure(n * 8)); P.score = ctx->an[slot++].as <double>();
	PV3_CUDA(ctx, ctx->an[slot].ensure(n * 4)); P.flags = ctx->an[slot++].as <uint32_t>();
	launch_rescore(P, st);
	PV3_CUDA(ctx, cudaGetLastError());
	ctx->total_launches += 1;
	PV3_CUDA(ctx, cudaMemcpyAsync(score, P.score, n * 8, cudaMemcpyDeviceToHost, st));
	PV3_CUDA(ctx, cudaMemcpyAsync(flags, P.flags, n * 4, cudaMemcpyDeviceToHost, st));
	PV3_CUDA(ctx, cudaStreamSynchronize(st));     // also keeps `lut` alive until its upload is done
	return PV3_OK;
}

int pv3_calculate_mapq(pv3_ctx *ctx, uint64_t n, const uint32_t *read_len, const uint32_t *other_len, const double *score,
	const double *next_score, uint8_t *mapq, uint32_t *flags)
{
	if (!ctx) return PV3_ERR_INVALID;
	cudaSetDevice(ctx->device);
	if (!n) return PV3_OK;
	if (!read_len || !other_len || !score || !next_score || !mapq || !flags) return fail(ctx, PV3_ERR_INVALID, "pv3_calculate_mapq: NULL argument");
	cudaStream_t const st = ctx->stream;
	struct { const void *src; size_t bytes; } const ins[4] = { { read_len, n * 4 }, { other_len, n * 4 }, { score, n * 8 }, { next_score, n * 8 } };
	for (int k = 0; k < 4; ++k) {
		PV3_CUDA(ctx, ctx->an[k].ensure(ins[k].bytes));
		PV3_CUDA(ctx, cudaMemcpyAsync(ctx->an[k].p, ins[k].src, ins[k].bytes, cudaMemcpyHostToDevice, st));
	}
	PV3_CUDA(ctx, ctx->an[4].ensure(n)); PV3_CUDA(ctx, ctx->an[5].ensure(n * 4));
	launch_mapq(n, ctx->an[0].as <uint32_t>(), ctx->an[1].as <uint32_t>(), ctx->an[2].as <double>(), ctx->an[3].as <double>(),
		ctx->an[4].as <uint8_t>(), ctx->an[5].as <uint32_t>(), st);
	PV3_CUDA(ctx, cudaGetLastError());
	ctx->total_launches += 1;
	PV3_CUDA(ctx, cudaMemcpyAsync(mapq, ctx->an[4].p, n, cudaMemcpyDeviceToHost, st));
	PV3_CUDA(ctx, cudaMemcpyAsync(flags, ctx->an[5].p, n * 4, cudaMemcpyDeviceToHost, st));
	PV3_CUDA(ctx, cudaStreamSynchronize(st));
	return PV3_OK;
}

int pv3_format_cigars(pv3_ctx *ctx, uint64_t n, const uint64_t *cig_off, const uint32_t *cig_len, const uint32_t *cig, uint64_t n_cig_words,
	uint64_t *text_off, char *text, uint64_t text_cap, uint64_t *text_used)
{
	if (!ctx) return PV3_ERR_INVALID;
	if (n && !cig_off) return fail(ctx, PV3_ERR_INVALID, "pv3_format_cigars: NULL array");
	OaParams P;
	memset(&P, 0, sizeof P);
	P.n = n; P.cig_off = cig_off; P.cig_len = cig_len; P.cig = cig;
	return format_text(ctx, "pv3_format_cigars", P, n_cig_words, 0, text_off, text, text_cap, text_used);
}

int pv3_get_timing(const pv3_ctx *ctx, pv3_timing *out)
{
	if (!ctx || !out) return PV3_ERR_INVALID;
	*out = ctx->timing;
	return PV3_OK;
}

uint64_t pv3_total_launches(const pv3_ctx *ctx) { return ctx ? ctx->total_launches : 0; }

void *pv3_host_alloc(size_t bytes)
{
	void *p = nullptr;
	if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) return nullptr;
	return p;
}

void pv3_host_free(void *p) { if (p) cudaFreeHost(p); }

} // extern "C"
