// panvc3_b200 — host side of the C-ABI (include/panvc3_b200.h): context, index upload, queries, annotation helpers.
// The batch pipeline (pv3_project, pv3_project_device, pv3_multi_*) is in pv3_pipeline.cu.
//
// The context replaces, for the hot path, what project_task / input_processor own in the reference
// (project-alignments/project_alignments.cc:219-330,485-654): per-task alignment_projector instances and their reusable
// vectors become device buffers owned here and reused across batches; the loaded msa_index and the reference contigs
// (sequence_buffer_store) become a device-resident index, every row and contig in its own allocation.
// There is no CPU fallback: every entry point fails with PV3_ERR_CUDA if the device is unusable.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "pv3_ctx.h"
#include "pv3_index_build.h"

using namespace pv3;

namespace {
std::string g_create_error;
}

namespace pv3 {

void WorkSet::release()
{
	for (DevBuf *b : { &counters, &jobs, &dp_arena, &pool, &deferred, &complex_list, &complex_ce, &big_list, &place, &scratch,
			&deferred2, &plan_rec, &plan_seg0, &plan_nseg, &segs, &seg_out, &seg_ops,
			&out_dst_pos, &out_flags, &out_cig_off, &out_cig_len, &out_arena, &out_rr_off, &out_rr_len, &out_rr, &out_mate }) b->release();
	for (auto &b : in) b.release();
	for (auto &b : un) b.release();
	for (auto &b : st) b.release();
	for (auto &e : ev) if (e) { cudaEventDestroy(e); e = nullptr; }
	for (auto &e : cev) if (e) { cudaEventDestroy(e); e = nullptr; }
}

int fail(pv3_ctx *ctx, int code, const char *fmt, ...)
{
	char buf[512];
	va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
	if (ctx) ctx->error = buf; else g_create_error = buf;
	return code;
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

int check_ready(pv3_ctx *ctx, const char *who)
{
	if (!ctx) return PV3_ERR_INVALID;
	if (!ctx->committed) return fail(ctx, PV3_ERR_STATE, "%s: pv3_index_commit has not been called", who);
	cudaError_t e = cudaSetDevice(ctx->device);
	if (e != cudaSuccess) return fail(ctx, PV3_ERR_CUDA, "%s: cudaSetDevice failed: %s", who, cudaGetErrorString(e));
	return PV3_OK;
}

} // namespace pv3

namespace {

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
	if (const char *e = getenv("PV3_CHUNK_EDGE_LAST")) { long long v = atoll(e); if (v >= 0) ctx->chunk_edge_last = (uint64_t) v; }
	if (const char *e = getenv("PV3_CHUNK_EDGE")) { long long v = atoll(e); if (v >= 0) ctx->chunk_edge = (uint64_t) v; }
	bool ok = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) == cudaSuccess
		&& cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking) == cudaSuccess
		&& cudaStreamCreateWithFlags(&ctx->stream3, cudaStreamNonBlocking) == cudaSuccess;
	ctx->stream = ctx->own_stream;
	for (auto &ws : ctx->sets) {
		for (auto &ev : ws.ev) ok = ok && cudaEventCreate(&ev) == cudaSuccess;
		for (auto &ev : ws.cev) ok = ok && cudaEventCreate(&ev) == cudaSuccess;
		ok = ok && ws.counters.ensure(sizeof(Counters)) == cudaSuccess && ws.place.ensure(64) == cudaSuccess;
	}
	for (auto &ev : ctx->ev_misc) ok = ok && cudaEventCreate(&ev) == cudaSuccess;
	ok = ok && cudaHostAlloc((void **) &ctx->h_counters, sizeof(Counters), cudaHostAllocMapped) == cudaSuccess;
	ok = ok && cudaHostAlloc((void **) &ctx->h_reports, sizeof(ChunkReport) * MAX_CHUNKS, cudaHostAllocMapped) == cudaSuccess;
	ok = ok && ctx->running.ensure(64) == cudaSuccess;
	if (!ok) { pv3_destroy(ctx); return fail(nullptr, PV3_ERR_CUDA, "pv3_create: stream / event / buffer creation failed: %s", cudaGetErrorString(cudaGetLastError())); }
	*out = ctx;
	return PV3_OK;
}

void pv3_destroy(pv3_ctx *ctx)
{
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	if (ctx->stream) cudaStreamSynchronize(ctx->stream);
	if (ctx->own_stream) cudaStreamSynchronize(ctx->own_stream);
	if (ctx->stream2) cudaStreamSynchronize(ctx->stream2);
	if (ctx->stream3) cudaStreamSynchronize(ctx->stream3);
	for (auto &ws : ctx->sets) ws.release();
	for (auto &r : ctx->rows) r.mem.release();
	for (auto &r : ctx->refs) r.mem.release();
	for (DevBuf *b : { &ctx->row_desc, &ctx->ref_desc, &ctx->ex_text, &ctx->ex_words, &ctx->ex_seq, &ctx->ex_status,
			&ctx->q_row, &ctx->q_row2, &ctx->q_pos, &ctx->q_out, &ctx->running })
		b->release();
	for (auto &b : ctx->an) b.release();
	if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
	if (ctx->h_reports) cudaFreeHost(ctx->h_reports);
	for (auto &ev : ctx->ev_misc) if (ev) cudaEventDestroy(ev);
	if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
	if (ctx->stream3) cudaStreamDestroy(ctx->stream3);
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
	ctx->committed = false;   // rows may be added later (lazily registered entries); the next commit uploads the new ones only
	ctx->rows.emplace_back();
	HostRow &row = ctx->rows.back();
	row.n_bits = n_bits;
	row.words.assign(gap_words, gap_words + (n_bits + 63) / 64);
	return (int) ctx->rows.size() - 1;
}

int pv3_index_add_ref(pv3_ctx *ctx, const char *bases, uint64_t n)
{
	if (!ctx || (!bases && n)) return fail(ctx, PV3_ERR_INVALID, "pv3_index_add_ref: bad argument");
	ctx->committed = false;
	int id;
	if (!ctx->free_refs.empty()) { id = ctx->free_refs.back(); ctx->free_refs.pop_back(); }
	else { ctx->refs.emplace_back(); id = (int) ctx->refs.size() - 1; }
	HostRef &ref = ctx->refs[id];
	ref.bases.assign(bases, bases + n);
	ref.n = n; ref.uploaded = false; ref.live = true;
	return id;
}

int pv3_index_remove_ref(pv3_ctx *ctx, int ref_id)
{
	if (!ctx || ref_id < 0 || (size_t) ref_id >= ctx->refs.size() || !ctx->refs[ref_id].live)
		return fail(ctx, PV3_ERR_INVALID, "pv3_index_remove_ref: no such contig");
	cudaSetDevice(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	cudaStreamSynchronize(ctx->own_stream);
	cudaStreamSynchronize(ctx->stream2);
	cudaStreamSynchronize(ctx->stream3);
	HostRef &ref = ctx->refs[ref_id];
	ctx->target_code_bytes -= std::min <uint64_t>(ctx->target_code_bytes, (ref.n + 1) / 2);
	ref.mem.release();
	ref.bases.clear(); ref.bases.shrink_to_fit();
	ref.n = 0; ref.live = false; ref.uploaded = false;
	ctx->free_refs.push_back(ref_id);
	ctx->committed = false;             // the descriptor table must drop the pointer
	return PV3_OK;
}

int pv3_index_commit(pv3_ctx *ctx)
{
	if (!ctx) return PV3_ERR_INVALID;
	if (ctx->committed) return PV3_OK;
	PV3_CUDA(ctx, cudaSetDevice(ctx->device));
	if (ctx->stream) cudaStreamSynchronize(ctx->stream);
	cudaStreamSynchronize(ctx->own_stream);
	cudaStreamSynchronize(ctx->stream2);
	cudaStreamSynchronize(ctx->stream3);

	// host mirrors of the descriptor tables: entries of rows / contigs uploaded by earlier commits stay as they are
	std::vector <RowDesc> &rd = ctx->h_row_desc;
	std::vector <RefDesc> &fd = ctx->h_ref_desc;
	RowDesc const zero_row = {}; RefDesc const zero_ref = {};
	rd.resize(ctx->rows.size() ? ctx->rows.size() : 1, zero_row);
	fd.resize(ctx->refs.size() ? ctx->refs.size() : 1, zero_ref);
	std::vector <uint8_t> host;
	for (size_t r = 0; r < ctx->rows.size(); ++r) {
		HostRow &row = ctx->rows[r];
		if (row.uploaded) continue;
		// Layout of a row's allocation: [words | blk_zeros | sel], 256-byte aligned parts.
		uint64_t n_blocks = (row.n_bits + BLK_BITS - 1) / BLK_BITS; if (!n_blocks) n_blocks = 1;
		uint64_t const n_words_alloc = n_blocks * BLK_WORDS + PAD_WORDS;
		size_t const words_off = 0;
		size_t const blk_off = align_up(words_off + n_words_alloc * 8, 256);
		size_t const sel_off = align_up(blk_off + (n_blocks + 1 + 32) * 8, 256);
		uint64_t zeros = 0;
		for (uint64_t w = 0; w < row.words.size(); ++w) {
			uint64_t v = row.words[w];
			if (w == row.words.size() - 1 && row.n_bits % 64) v |= ~0ull << (row.n_bits % 64);
			zeros += (uint64_t) __builtin_popcountll(~v);
		}
		uint64_t const n_sel = (zeros ? (zeros - 1) / SEL_RATE + 1 : 0) + 1;
		size_t const total = align_up(sel_off + n_sel * 4, 256);
		host.assign(total, 0);
		uint64_t *words = (uint64_t *) (host.data() + words_off);
		uint64_t *blk = (uint64_t *) (host.data() + blk_off);
		uint32_t *sel = (uint32_t *) (host.data() + sel_off);
		for (uint64_t w = 0; w < n_words_alloc; ++w) words[w] = (w < row.words.size()) ? row.words[w] : ~0ull;
		if (row.n_bits % 64) words[row.words.size() - 1] |= ~0ull << (row.n_bits % 64);
		uint64_t z = 0, next_sel = 0;   // next_sel: index of the next sample to write
		for (uint64_t b = 0; b < n_blocks; ++b) {
			blk[b] = z;
			uint64_t zb = 0;
			for (int w = 0; w < BLK_WORDS; ++w) zb += (uint64_t) __builtin_popcountll(~words[b * BLK_WORDS + w]);
			// zero numbers (1-based) in this block: z+1 .. z+zb; sample s covers zero number s*SEL_RATE+1
			while (next_sel + 1 < n_sel && next_sel * (uint64_t) SEL_RATE + 1 <= z + zb) sel[next_sel++] = (uint32_t) b;
			z += zb;
		}
		for (uint64_t b = n_blocks; b < n_blocks + 1 + 32; ++b) blk[b] = z;
		for (uint64_t s = next_sel; s < n_sel; ++s) sel[s] = (uint32_t) (n_blocks - 1);   // sentinel(s)
		PV3_CUDA(ctx, row.mem.ensure(total));
		PV3_CUDA(ctx, cudaMemcpy(row.mem.p, host.data(), total, cudaMemcpyHostToDevice));
		uint8_t *base = row.mem.as <uint8_t>();
		rd[r].words = (const uint64_t *) (base + words_off);
		rd[r].blk_zeros = (const uint64_t *) (base + blk_off);
		rd[r].sel = (const uint32_t *) (base + sel_off);
		rd[r].n_bits = row.n_bits; rd[r].n_zeros = zeros; rd[r].n_blocks = n_blocks; rd[r].n_words_alloc = n_words_alloc;
		ctx->index_bytes += total;
		row.uploaded = true;
		std::vector <uint64_t>().swap(row.words);       // the host copy is not needed again
	}
	for (size_t r = 0; r < ctx->refs.size(); ++r) {
		HostRef &ref = ctx->refs[r];
		if (!ref.live) { fd[r].bases = nullptr; fd[r].codes4 = nullptr; fd[r].n = 0; continue; }
		if (ref.uploaded) continue;
		size_t const code_off = align_up(ref.n + 512, 256);
		size_t const total = align_up(code_off + codes4_bytes(ref.n), 256);
		host.assign(total, 0);
		memcpy(host.data(), ref.bases.data(), ref.n);
		pack_codes4(ref.bases.data(), ref.n, host.data() + code_off);
		PV3_CUDA(ctx, ref.mem.ensure(total));
		PV3_CUDA(ctx, cudaMemcpy(ref.mem.p, host.data(), total, cudaMemcpyHostToDevice));
		fd[r].bases = ref.mem.as <char>();
		fd[r].codes4 = ref.mem.as <uint8_t>() + code_off;
		fd[r].n = ref.n;
		ctx->index_bytes += total;
		ctx->target_code_bytes += (ref.n + 1) / 2;
		ref.uploaded = true;
		std::string().swap(ref.bases);
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

int pv3_project_positions_device(pv3_ctx *ctx, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row, const uint64_t *pos, uint64_t *out_pos)
{
	int st = check_ready(ctx, "pv3_project_positions_device");
	if (st) return st;
	if (!n) return PV3_OK;
	if (!src_row || !dst_row || !pos || !out_pos) return fail(ctx, PV3_ERR_INVALID, "pv3_project_positions_device: NULL argument");
	launch_positions(ctx->row_desc.as <RowDesc>(), (uint32_t) ctx->rows.size(), n, src_row, dst_row, pos, out_pos, ctx->stream);
	PV3_CUDA(ctx, cudaGetLastError());
	ctx->total_launches += 1;
	return PV3_OK;
}

int pv3_project_positions(pv3_ctx *ctx, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row, const uint64_t *pos, uint64_t *out_pos)
{
	int st = check_ready(ctx, "pv3_project_positions");
	if (st) return st;
	if (!n) return PV3_OK;
	if (!src_row || !dst_row || !pos || !out_pos) return fail(ctx, PV3_ERR_INVALID, "pv3_project_positions: NULL argument");
	PV3_CUDA(ctx, ctx->q_row.ensure(n * 4)); PV3_CUDA(ctx, ctx->q_row2.ensure(n * 4));
	PV3_CUDA(ctx, ctx->q_pos.ensure(n * 8)); PV3_CUDA(ctx, ctx->q_out.ensure(n * 8));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->q_row.p, src_row, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->q_row2.p, dst_row, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->q_pos.p, pos, n * 8, cudaMemcpyHostToDevice, ctx->stream));
	launch_positions(ctx->row_desc.as <RowDesc>(), (uint32_t) ctx->rows.size(), n, ctx->q_row.as <uint32_t>(),
		ctx->q_row2.as <uint32_t>(), ctx->q_pos.as <uint64_t>(), ctx->q_out.as <uint64_t>(), ctx->stream);
	PV3_CUDA(ctx, cudaGetLastError());
	ctx->total_launches += 1;
	PV3_CUDA(ctx, cudaMemcpyAsync(out_pos, ctx->q_out.p, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
	PV3_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PV3_OK;
}

int pv3_index_query(pv3_ctx *ctx, int op, uint64_t n, const uint32_t *row, const uint64_t *x, uint64_t *out)
{
	int st = check_ready(ctx, "pv3_index_query");
	if (st) return st;
	if (op < 0 || op > 2) return fail(ctx, PV3_ERR_INVALID, "pv3_index_query: op must be 0, 1 or 2");
	if (!n) return PV3_OK;
	PV3_CUDA(ctx, ctx->q_row.ensure(n * 4)); PV3_CUDA(ctx, ctx->q_pos.ensure(n * 8)); PV3_CUDA(ctx, ctx->q_out.ensure(n * 8));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->q_row.p, row, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	PV3_CUDA(ctx, cudaMemcpyAsync(ctx->q_pos.p, x, n * 8, cudaMemcpyHostToDevice, ctx->stream));
	launch_index_query(ctx->row_desc.as <RowDesc>(), (uint32_t) ctx->rows.size(), op, n, ctx->q_row.as <uint32_t>(),
		ctx->q_pos.as <uint64_t>(), ctx->q_out.as <uint64_t>(), ctx->stream);
	PV3_CUDA(ctx, cudaGetLastError());
	ctx->total_launches += 1;
	PV3_CUDA(ctx, cudaMemcpyAsync(out, ctx->q_out.p, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
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
	PV3_CUDA(ctx, cudaEventRecord(ctx->ev_misc[0], ctx->stream));
	PV3_CUDA(ctx, launch_msa_extract(text, n, gap_words, ungapped, ctx->ex_status.as <unsigned long long>(), ctx->stream, &launches));
	PV3_CUDA(ctx, cudaEventRecord(ctx->ev_misc[1], ctx->stream));
	ctx->total_launches += launches;
	unsigned long long total = 0;
	PV3_CUDA(ctx, cudaMemcpyAsync(&total, ctx->ex_status.as <unsigned long long>() + msa_extract_total_word(n), 8, cudaMemcpyDeviceToHost, ctx->stream));
	PV3_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	cudaEventElapsedTime(&ctx->timing.ms_msa_extract, ctx->ev_misc[0], ctx->ev_misc[1]);
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
	if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) return nullptr;
	return p;
}

void pv3_host_free(void *p) { if (p) cudaFreeHost(p); }

int pv3_host_register(void *p, size_t bytes)
{
	if (!p || !bytes) return PV3_ERR_INVALID;
	return cudaHostRegister(p, bytes, cudaHostRegisterMapped | cudaHostRegisterPortable) == cudaSuccess ? PV3_OK : PV3_ERR_CUDA;
}

int pv3_host_unregister(void *p) { return (p && cudaHostUnregister(p) == cudaSuccess) ? PV3_OK : PV3_ERR_CUDA; }

} // extern "C"
