// panvc3_b200 — the batch pipeline behind pv3_project / pv3_project_device / pv3_multi_project (include/panvc3_b200.h).
//
// One batch (or one chunk of a host batch) is ONE asynchronous sequence on one stream, with no host round trip inside:
//   [H2D copies → unpack_in]  memset counters → rewrite (thread kernel + warp kernel over its deferred list, length read
//   on the device) → DP (thread-per-run and warp-per-run kernels; run count and their split read on the device) → splice
//   → mate positions → [pack_out: ordered output to the caller's buffers + a report in mapped pinned memory]
// The host synchronises once per call. What it cannot know beforehand — arena sizes, DP op space, HBM trace slots — is
// sized from the previous call's reports; a kernel that runs out of room raises a bit in the report and the call is
// repeated with the sizes the report names (inputs are never modified, so a re-run is safe). This replaces the loop of
// libpanvc3/alignment_projector.cc:57-151 (rewrite, find runs, realign each, splice) and the 16-slot task ring around it
// (include/panvc3/spsc_queue.hh, project-alignments/project_alignments.cc:597-631).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pv3_ctx.h"
#include "pv3_segment_core.h"

using namespace pv3;

namespace {

constexpr size_t POOL_BUDGET = (size_t) 8 << 30;   // HBM scratch for traces of runs too large for shared memory

struct Caps { uint64_t arena, rr; };

// Device pointers of one batch / chunk, everything the kernels read and write.
struct DevBatch {
	pv3_batch_in  in;         // device pointers (compact fields unused)
	pv3_batch_out out;        // device pointers: per-record arrays, arena (cig, cig_cap), rr (rr, rr_cap), mate_dst_pos
	uint64_t      arena_off = 0, rr_base = 0;
	uint64_t      expect_deferred = 0;   // host-side guess: records the thread kernel will not take (long reads)
};

int ensure_work(pv3_ctx *ctx, WorkSet &ws, uint64_t n, uint64_t rr_cap)
{
	cudaStream_t const s = ws.stream;
	PV3_CUDA(ctx, ws.jobs.ensure_idle((rr_cap ? rr_cap : 1) * sizeof(Job), s));
	PV3_CUDA(ctx, ws.big_list.ensure_idle((rr_cap ? rr_cap : 1) * 8, s));
	PV3_CUDA(ctx, ws.deferred.ensure_idle(n * 8 + 8, s));
	PV3_CUDA(ctx, ws.complex_list.ensure_idle(n * 8 + 8, s));
	PV3_CUDA(ctx, ws.complex_ce.ensure_idle(n * 16 + 16, s));
	PV3_CUDA(ctx, ws.deferred2.ensure_idle(n * 8 + 8, s));
	PV3_CUDA(ctx, ws.plan_rec.ensure_idle(n * 8 + 8, s));
	PV3_CUDA(ctx, ws.plan_seg0.ensure_idle(n * 8 + 8, s));
	PV3_CUDA(ctx, ws.plan_nseg.ensure_idle(n * 4 + 8, s));
	{
		uint64_t const seg_cap = std::max <uint64_t>(ws.seg_cap_hint, n / 8 + 4096), seg_ops_cap = std::max <uint64_t>(ws.seg_ops_hint, n + 65536);
		PV3_CUDA(ctx, ws.segs.ensure_idle(seg_cap * sizeof(seg::SegDesc), s));
		PV3_CUDA(ctx, ws.seg_out.ensure_idle(seg_cap * sizeof(seg::SegOut), s));
		PV3_CUDA(ctx, ws.seg_ops.ensure_idle(seg_ops_cap * 4, s));
		ws.seg_cap = seg_cap; ws.seg_ops_cap = seg_ops_cap;
	}
	uint64_t const dp_words = std::max <uint64_t>(ws.dp_arena_hint, n / 4 + 4096);
	PV3_CUDA(ctx, ws.dp_arena.ensure_idle((dp_words + 1) * 4, s));
	if (ws.pool_hint) {
		size_t const slot = align_up((size_t) ws.pool_hint, 256);
		uint64_t slots = k3_default_grid_warps(ctx->n_sms);
		uint64_t const fit = POOL_BUDGET / slot ? POOL_BUDGET / slot : 1;
		if (slots > fit) slots = fit / 8 * 8 ? fit / 8 * 8 : 1;
		if (slot > ws.pool_slot || slots > ws.pool_slots || !ws.pool.p) {
			PV3_CUDA(ctx, ws.pool.ensure_idle(slot * slots, s));
			ws.pool_slot = slot; ws.pool_slots = slots;
		}
	}
	return PV3_OK;
}

// Enqueues the kernels of one batch on ws.stream. No synchronisation.
int enqueue_kernels(pv3_ctx *ctx, WorkSet &ws, DevBatch const &B, uint64_t &launches)
{
	cudaStream_t const stream = ws.stream;
	pv3_batch_in const &in = B.in;
	pv3_batch_out const &out = B.out;
	K1Params P;
	memset(&P, 0, sizeof P);
	bool mates_done = false;
	P.n = in.n;
	P.src_row = in.src_row; P.dst_row = in.dst_row; P.dst_ref = in.dst_ref; P.src_pos = in.src_pos;
	P.seq_off = in.seq_off; P.seq = in.seq; P.seq_len = in.seq_len; P.qual_off = in.qual_off; P.qual = in.qual;
	P.cig_off = in.cig_off; P.cig = in.cig;
	P.rows = ctx->row_desc.as <RowDesc>(); P.n_rows = (uint32_t) ctx->rows.size();
	P.refs = ctx->ref_desc.as <RefDesc>(); P.n_refs = (uint32_t) ctx->refs.size();
	P.dst_pos = out.dst_pos; P.flags = out.flags; P.cig_out_off = out.cig_off; P.cig_out_len = out.cig_len;
	P.arena = out.cig; P.arena_cap = out.cig_cap;
	P.rr_off = out.rr_off; P.rr_len = out.rr_len; P.rr = out.rr; P.rr_cap = out.rr_cap;
	P.jobs = ws.jobs.as <Job>();
	P.counters = ws.counters.as <Counters>();
	P.gap_open = ctx->gap_open; P.gap_ext = ctx->gap_ext;
	P.stage = ctx->stage;
	P.deferred = ws.deferred.as <uint64_t>();
	P.complex_list = ws.complex_list.as <uint64_t>();
	P.complex_ce = ws.complex_ce.as <uint64_t>();
	P.plan_rec = ws.plan_rec.as <uint64_t>(); P.plan_seg0 = ws.plan_seg0.as <uint64_t>(); P.plan_nseg = ws.plan_nseg.as <uint32_t>();
	P.segs = ws.segs.p; P.seg_out = ws.seg_out.p; P.seg_cap = ws.seg_cap;
	P.seg_ops = ws.seg_ops.as <uint32_t>(); P.seg_ops_cap = ws.seg_ops_cap;
	P.deferred2 = ws.deferred2.as <uint64_t>();
	P.arena_off = B.arena_off; P.rr_base = B.rr_base;

	PV3_CUDA(ctx, cudaMemsetAsync(ws.counters.p, 0, sizeof(Counters), stream));
	PV3_CUDA(ctx, cudaEventRecord(ws.ev[0], stream));
	if (ctx->kernel_path != 1) {
		// short records: one thread each; whatever exceeds the kernel's fixed quotas comes back in the deferred list, which the
		// warp-per-record kernel takes with a resident grid that reads the list length on the device
		// path 0: the run-walk kernel (v13); path 2: the word-parallel kernel (v14). PV3_K1_KERNEL=word|thread overrides path 0.
		const char *e = getenv("PV3_K1_KERNEL");
		bool const word = ctx->kernel_path == 2 || (ctx->kernel_path == 0 && e && e[0] == 'w');
		if (word) launch_k1_word(P, ctx->n_sms, stream);
		else {
			K1Params PT = P;
			if (in.mate_row && out.mate_dst_pos) { PT.mate_row = in.mate_row; PT.mate_pos = in.mate_pos; PT.mate_dst_pos = out.mate_dst_pos; mates_done = true; }
			ws.k1t_layout = -1; ws.k1t_records = in.n;
			if (k1_thread_layout_is_tunable(ctx->target_code_bytes)) {
				auto &T = ctx->tuner;
				uint32_t const c = T.launches++;
				// two samples of each layout before either is trusted: a context's first launch is cold (page faults of fresh
				// allocations, instruction cache) and once made the staged layout look slower than it is
				if (T.seen[0] < 2 || T.seen[1] < 2) ws.k1t_layout = T.seen[0] <= T.seen[1] ? 0 : 1;
				else {
					int const best = T.ns_per_rec[1] < T.ns_per_rec[0] ? 1 : 0;
					ws.k1t_layout = ((c & 31u) == 31u || T.reprobe) ? 1 - best : best;
				}
				T.reprobe = false;
			}
			launch_k1_thread(PT, ctx->n_sms, ctx->target_code_bytes, stream, ws.k1t_layout);
		}
		PV3_CUDA(ctx, cudaGetLastError());
		PV3_CUDA(ctx, cudaEventRecord(ws.ev[4], stream));
		// what the thread kernel could not take (long reads above all): cut into segments, one thread per segment, reassembled by
		// a warp per record; whatever fails there is left to the warp-per-record kernel. All list lengths are read on the device.
		K1Params P2 = P;
		// Which route the deferred records take is a matter of speed only (both are exact): a handful — short reads beyond a
		// quota — go straight to the warp-per-record kernel (one launch: four launches that each find a dozen records cost
		// 50 us of latency), many — long reads — through the segment kernels. The host's guess: the last batch's count and,
		// where it can see them, the read lengths.
		uint64_t const expected = std::min <uint64_t>(in.n, std::max <uint64_t>(ws.expect_deferred, B.expect_deferred));
		bool const use_segments = !getenv("PV3_NO_SEGMENTS") && (expected >= 256 || getenv("PV3_SEGMENTS_ALWAYS"));
		if (use_segments) {
			launch_k1_segments(P, ctx->n_sms, expected, stream);
			launch_k1_assemble(P, ctx->n_sms, expected, stream);
			PV3_CUDA(ctx, cudaGetLastError());
			P2.rec_list = ws.deferred2.as <uint64_t>();
			P2.n_list_dev = &ws.counters.as <Counters>()->n_deferred2;
			launches += 3;
		} else {
			P2.rec_list = ws.deferred.as <uint64_t>();
			P2.n_list_dev = &ws.counters.as <Counters>()->n_deferred;
		}
		launch_k1_list(P2, ctx->n_sms, use_segments ? expected / 16 + 64 : expected, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		launches += 2;
	} else {
		launch_k1(P, ctx->n_sms, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		PV3_CUDA(ctx, cudaEventRecord(ws.ev[4], stream));
		launches += 1;
	}
	PV3_CUDA(ctx, cudaEventRecord(ws.ev[1], stream));
	if (ctx->stage == 0) {
		K3Params K;
		memset(&K, 0, sizeof K);
		K.k1 = P;
		K.n_jobs = ~0ull;                                      // read on the device
		K.dp_thread_min = ctx->dp_thread_min;
		K.grid_hint = (uint32_t) ctx->n_sms * 8;
		K.dp_arena = ws.dp_arena.as <uint32_t>(); K.dp_arena_cap = ws.dp_arena.cap / 4 - 1;
		K.pool = ws.pool.as <uint8_t>(); K.pool_slot_bytes = ws.pool.p ? ws.pool_slot : 0; K.pool_slots = ws.pool.p ? ws.pool_slots : 0;
		K.big_list = ws.big_list.as <uint64_t>();
		memcpy(K.lut, ctx->lut, sizeof K.lut);
		// small runs (the usual case: a few bases wide): one thread per run once a batch holds enough of them; the rest: one warp
		// per run. Both kernels read the counts the rewrite left on the device and agree on the split.
		launch_k3_thread(K, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		uint64_t grid_warps = k3_default_grid_warps(ctx->n_sms);
		if (ws.pool.p && ws.pool_slots < grid_warps) grid_warps = ws.pool_slots;
		launch_k3(K, ctx->n_sms, grid_warps, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		PV3_CUDA(ctx, cudaEventRecord(ws.ev[2], stream));
		launch_k4(P, ~0ull, K.dp_arena, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		launches += 3;
	} else PV3_CUDA(ctx, cudaEventRecord(ws.ev[2], stream));
	PV3_CUDA(ctx, cudaEventRecord(ws.ev[3], stream));
	if (in.mate_row && out.mate_dst_pos && !mates_done) {
		launch_positions(ctx->row_desc.as <RowDesc>(), (uint32_t) ctx->rows.size(), in.n, in.mate_row, in.dst_row, in.mate_pos, out.mate_dst_pos, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		launches += 1;
	}
	return PV3_OK;
}

// What a finished batch says about the next one's buffers. Returns the overflow bits that need a re-run.
unsigned digest(WorkSet &ws, uint64_t n, unsigned overflow, uint64_t arena_need, uint64_t jobs, uint64_t dp_path, uint64_t dp_max_pool,
	uint64_t seg_need, uint64_t seg_ops_need, uint64_t n_deferred)
{
	if (overflow & 128u) { ws.seg_cap_hint = seg_need + seg_need / 8 + 4096; ws.seg_ops_hint = std::max <uint64_t>(ws.seg_ops_hint, seg_need * 10); }
	if (overflow & 256u) ws.seg_ops_hint = seg_ops_need + seg_ops_need / 8 + 4096;
	if (overflow & 4u) ws.dp_arena_hint = dp_path + dp_path / 4 + 4096;
	else if (dp_path + 4096 > ws.dp_arena_hint) ws.dp_arena_hint = std::max <uint64_t>(ws.dp_arena_hint, dp_path + dp_path / 8 + 1024);
	if (overflow & 8u) ws.pool_hint = dp_max_pool;
	if (overflow & 1u) ws.arena_cap_hint = arena_need + arena_need / 4 + 4096;
	if (overflow & 2u) ws.rr_cap_hint = jobs + jobs / 4 + 1024;
	ws.many_runs = jobs * 16 > n;
	ws.expect_deferred = n_deferred + n_deferred / 4;
	return overflow & (1u | 2u | 4u | 8u | 128u | 256u);
}

void harvest_timing(pv3_ctx *ctx, WorkSet &ws, float *ms /* rewrite, realign, splice, total, rewrite_main */)
{
	float t = 0;
	if (cudaEventElapsedTime(&t, ws.ev[0], ws.ev[1]) == cudaSuccess) ms[0] += t;
	if (cudaEventElapsedTime(&t, ws.ev[1], ws.ev[2]) == cudaSuccess) ms[1] += t;
	if (cudaEventElapsedTime(&t, ws.ev[2], ws.ev[3]) == cudaSuccess) ms[2] += t;
	if (cudaEventElapsedTime(&t, ws.ev[0], ws.ev[3]) == cudaSuccess) ms[3] += t;
	if (cudaEventElapsedTime(&t, ws.ev[0], ws.ev[4]) == cudaSuccess) {
		ms[4] += t;
		if (ws.k1t_layout >= 0 && ws.k1t_records >= 65536) {          // (small launches are latency, not layout)
			auto &T = ctx->tuner;
			float const ns = t * 1e6f / (float) ws.k1t_records;
			float const old = T.ns_per_rec[ws.k1t_layout];
			bool const moved = T.seen[ws.k1t_layout] && (ns > 1.15f * old || ns < 0.85f * old);
			T.ns_per_rec[ws.k1t_layout] = (T.seen[ws.k1t_layout] && !moved) ? 0.5f * (old + ns) : ns;
			if (T.seen[ws.k1t_layout] < 2) ++T.seen[ws.k1t_layout];
			if (moved) T.reprobe = true;
			static bool const trace = getenv("PV3_TRACE") != nullptr;
			if (trace) fprintf(stderr, "[pv3 trace] thread kernel layout %d: %.3f ns per record (staged %.3f, through L1 %.3f)%s\n", ws.k1t_layout, ns,
				T.ns_per_rec[0], T.ns_per_rec[1], moved ? " — moved, probing the other layout" : "");
		}
		ws.k1t_layout = -1;
	}
}

// device alias of pinned (mapped) host memory, or null when the pointer is pageable / foreign
void *mapped_alias(const void *p)
{
	if (!p) return nullptr;
	cudaPointerAttributes a;
	if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return nullptr; }
	if (a.type == cudaMemoryTypeHost && a.devicePointer) return a.devicePointer;
	if (a.type == cudaMemoryTypeManaged && a.devicePointer) return a.devicePointer;
	return nullptr;
}

} // namespace

extern "C" {

int pv3_project_device(pv3_ctx *ctx, const pv3_batch_in *in, pv3_batch_out *out)
{
	int st = check_ready(ctx, "pv3_project_device");
	if (st) return st;
	if (!in || !out) return fail(ctx, PV3_ERR_INVALID, "pv3_project_device: NULL batch");
	out->cig_used = out->rr_used = out->n_flagged = out->n_realigned = out->dp_cells = 0;
	memset(&ctx->timing, 0, sizeof ctx->timing);
	if (!in->n) return PV3_OK;
	WorkSet &ws = ctx->sets[0];
	ws.stream = ctx->stream;
	DevBatch B;
	B.in = *in; B.out = *out;
	uint64_t launches = 0;
	for (int attempt = 0; ; ++attempt) {
		st = ensure_work(ctx, ws, in->n, out->rr_cap);
		if (st) return st;
		st = enqueue_kernels(ctx, ws, B, launches);
		if (!st && small_copy(ctx->h_counters, ws.counters.p, sizeof(Counters), ws.stream) != cudaSuccess) st = fail(ctx, PV3_ERR_CUDA, "counter read-back failed");
		launches += 1;
		cudaError_t const e = cudaStreamSynchronize(ws.stream);
		if (st) return st;
		if (e != cudaSuccess) return fail(ctx, PV3_ERR_CUDA, "pv3_project_device: %s", cudaGetErrorString(e));
		Counters const c = *ctx->h_counters;
		unsigned const again = digest(ws, in->n, c.overflow, c.arena_cursor + c.dp_path, c.job_cursor, c.dp_path, c.dp_max_pool, c.seg_cursor, c.seg_ops_cursor, c.n_deferred);
		out->cig_used = (c.overflow & 1u) ? c.arena_cursor + c.dp_path : c.arena_cursor;
		out->rr_used = c.job_cursor;
		if (again & (1u | 2u)) {
			ctx->total_launches += launches; ctx->timing.launches = launches;
			return fail(ctx, PV3_ERR_CAPACITY, "output capacity too small: need cig_cap >= %llu, rr_cap >= %llu",
				(unsigned long long) (c.arena_cursor + c.dp_path), (unsigned long long) c.job_cursor);
		}
		if (again && attempt < 4) continue;       // DP op space or HBM trace slots were too small: sized now, run again
		if (again) return fail(ctx, PV3_ERR_CUDA, "pv3_project_device: internal buffers could not be sized (overflow 0x%x)", c.overflow);
		out->n_flagged = c.n_flagged; out->n_realigned = c.n_realigned; out->dp_cells = c.dp_cells;
		float ms[5] = { 0, 0, 0, 0, 0 };
		harvest_timing(ctx, ws, ms);
		ctx->timing.ms_rewrite = ms[0]; ctx->timing.ms_realign = ms[1]; ctx->timing.ms_splice = ms[2]; ctx->timing.ms_total = ms[3];
		ctx->timing.ms_rewrite_main = ms[4]; ctx->timing.n_deferred = c.n_deferred; ctx->timing.launches = launches;
		ctx->total_launches += launches;
		return PV3_OK;
	}
}

} // extern "C"

namespace {

struct HostCall {
	pv3_ctx *ctx;
	const pv3_batch_in *in;
	pv3_batch_out *out;
	uint64_t rec_begin, rec_end;        // records of the caller's batch this context handles (pv3_multi shards; whole batch otherwise)
	uint64_t cig_place, rr_place;       // where this context's ordered arenas start in the caller's arrays
	uint64_t cig_room, rr_room;         // and how much room they have
	// state between enqueue and finish
	uint64_t n_chunks = 0, cuts[MAX_CHUNKS + 1];
	bool zero_copy = false, own_streams = false;
	uint64_t launches = 0;
};

// Enqueues everything for the records [rec_begin, rec_end) of the call on the context's two streams. No synchronisation
// unless a buffer has to grow while an earlier chunk is still in flight on its stream, or the caller's memory is pageable.
std::vector <cudaEvent_t> g_timeline;      // PV3_TRACE_TIMELINE: six events per chunk (single-context debugging aid)

void dump_timeline(uint64_t n_chunks)
{
	if (g_timeline.empty()) return;
	static const char *names[6] = { "h2d_begin", "h2d_end", "kernels_end", "pack_begin", "pack_end", "out_end" };
	for (uint64_t c = 0; c < n_chunks; ++c) {
		fprintf(stderr, "[pv3 timeline] chunk %llu:", (unsigned long long) c);
		for (int k = 0; k < 6; ++k) {
			float t = -1;
			if (cudaEventElapsedTime(&t, g_timeline[0], g_timeline[c * 6 + k]) != cudaSuccess) { cudaGetLastError(); t = -1; }
			fprintf(stderr, " %s %.3f", names[k], t);
		}
		fprintf(stderr, "\n");
	}
}

int host_enqueue(HostCall &hc)
{
	pv3_ctx *ctx = hc.ctx;
	const pv3_batch_in *in = hc.in;
	pv3_batch_out *out = hc.out;
	uint64_t const n = hc.rec_end - hc.rec_begin;
	PV3_CUDA(ctx, cudaSetDevice(ctx->device));
	hc.own_streams = (ctx->stream == ctx->own_stream);
	uint64_t n_chunks = hc.own_streams ? (n + ctx->chunk_records / 2) / ctx->chunk_records : 1;
	if (n_chunks < 1) n_chunks = 1;
	if (n_chunks > 64) n_chunks = 64;
	constexpr int NS = pv3_ctx::N_SETS;
	cudaStream_t const streams[NS] = { ctx->stream, hc.own_streams ? ctx->stream2 : ctx->stream, hc.own_streams ? ctx->stream3 : ctx->stream };
	// Chunk cuts (tile-aligned). Only the first chunk's upload and the last chunk's kernels and write-back are not hidden
	// behind other chunks' transfers, so those two chunks are small and the ones between them large.
	{
		uint64_t *cuts = hc.cuts;
		uint64_t const edge = (n_chunks >= 2 && ctx->chunk_edge && n > 4 * ctx->chunk_edge) ? ctx->chunk_edge : 0;
		uint64_t const last = edge ? (ctx->chunk_edge_last ? std::min(ctx->chunk_edge_last, edge) : edge) : 0;     // the drain is the exposed one
		if (edge) {
			uint64_t n_mid = (n - edge - last + ctx->chunk_records / 2) / ctx->chunk_records;
			if (n_mid < 1) n_mid = 1;
			if (n_mid > 62) n_mid = 62;
			n_chunks = n_mid + 2;
			cuts[0] = 0;
			for (uint64_t c = 0; c <= n_mid; ++c) cuts[c + 1] = (edge + (n - edge - last) * c / n_mid) / 128 * 128;
			cuts[n_chunks] = n;
		} else {
			for (uint64_t c = 0; c < n_chunks; ++c) cuts[c] = (n * c / n_chunks) / 128 * 128;
			cuts[n_chunks] = n;
		}
	}
	hc.n_chunks = n_chunks;

	// Where do results go? Pinned (mapped) caller memory: the pack kernel of every chunk writes them in place. Otherwise they
	// are packed into device staging and copied after a per-chunk synchronisation.
	struct Dest { void *host; void *alias; };
	auto dest = [](void *p) { Dest d; d.host = p; d.alias = mapped_alias(p); return d; };
	Dest const d_dst_pos = dest(out->dst_pos), d_dst_pos32 = dest(out->dst_pos32), d_flags = dest(out->flags), d_flags8 = dest(out->flags8),
		d_cig_off = dest(out->cig_off), d_cig_len = dest(out->cig_len), d_cig_len16 = dest(out->cig_len16), d_cig = dest(out->cig),
		d_rr_off = dest(out->rr_off), d_rr_len = dest(out->rr_len), d_rr_len8 = dest(out->rr_len8), d_rr = dest(out->rr),
		d_mate = dest(out->mate_dst_pos), d_mate32 = dest(out->mate_dst_pos32);
	const Dest *all[] = { &d_dst_pos, &d_dst_pos32, &d_flags, &d_flags8, &d_cig_off, &d_cig_len, &d_cig_len16, &d_cig, &d_rr_off, &d_rr_len, &d_rr_len8, &d_rr, &d_mate, &d_mate32 };
	bool zero_copy = true;
	for (auto d : all) if (d->host && !d->alias) zero_copy = false;
	if (getenv("PV3_NO_ZERO_COPY")) zero_copy = false;
	hc.zero_copy = zero_copy;
	bool const compact = in->compact != 0;
	bool const two_bit = compact && in->seq_2bit != 0;
	bool const packed = compact ? (in->seq_bam4 != 0 || two_bit) : in->seq_len != nullptr;
	bool const has_qual = compact ? in->has_qual != nullptr : in->qual_off != nullptr;
	bool const mates = compact ? in->mate_row16 != nullptr : in->mate_row != nullptr;
	const uint8_t *qual_alias = has_qual ? (const uint8_t *) mapped_alias(in->qual) : nullptr;

	static bool const trace_tl = getenv("PV3_TRACE_TIMELINE") != nullptr;      // debugging aid: per-chunk device timeline on stderr
	std::vector <cudaEvent_t> &tl = g_timeline;
	if (trace_tl && tl.empty()) { tl.resize(MAX_CHUNKS * 6); for (auto &e : tl) cudaEventCreate(&e); }
	auto mark = [&](uint64_t c, int k, cudaStream_t st) { if (trace_tl) cudaEventRecord(tl[c * 6 + k], st); };
	PV3_CUDA(ctx, cudaMemsetAsync(ctx->running.p, 0, 16, streams[0]));
	PV3_CUDA(ctx, cudaEventRecord(ctx->ev_misc[2], streams[0]));
	for (uint64_t c = 0; c < n_chunks; ++c) {
		int const k = (int) (c % NS);
		WorkSet &ws = ctx->sets[k];
		ws.stream = streams[k];
		cudaStream_t const stream = ws.stream;
		uint64_t const b = hc.rec_begin + hc.cuts[c], e = hc.rec_begin + hc.cuts[c + 1], m = e - b;
		ChunkReport *report = ctx->h_reports + c;
		memset(report, 0, sizeof *report);
		if (!m) { report->done = 1; continue; }
		uint64_t const s0 = in->seq_off[b], s1 = in->seq_off[e], c0 = in->cig_off[b], c1 = in->cig_off[e];
		uint64_t const q0 = (has_qual && in->qual_off) ? in->qual_off[b] : 0, q1 = (has_qual && in->qual_off) ? in->qual_off[e] : 0;
		if (has_qual && !in->qual_off) return fail(ctx, PV3_ERR_INVALID, "pv3_project: qual_off must be valid on the host when records have QUAL");
		bool const qual_bulk = has_qual && (!qual_alias || ws.many_runs);

		// ---- uploads -------------------------------------------------------------------------------------------------
		struct Copy { int slot; const void *src; size_t bytes; };
		Copy copies[16]; int n_copies = 0;
		auto add = [&](int slot, const void *src, size_t bytes) { copies[n_copies].slot = slot; copies[n_copies].src = src; copies[n_copies].bytes = bytes; ++n_copies; };
		if (!compact) {
			add(0, in->src_row + b, m * 4); add(1, in->dst_row + b, m * 4); add(2, in->dst_ref + b, m * 4); add(3, in->src_pos + b, m * 8);
			add(4, in->seq_off + b, (m + 1) * 8); add(5, in->cig_off + b, (m + 1) * 8);
			if (has_qual) add(6, in->qual_off + b, (m + 1) * 8);
			if (packed) add(7, in->seq_len + b, m * 4);
			if (mates) { add(8, in->mate_row + b, m * 4); add(9, in->mate_pos + b, m * 8); }
		} else {
			if (b % 8) return fail(ctx, PV3_ERR_INVALID, "pv3_project: internal chunk cut not byte-aligned for has_qual");
			add(0, in->src_row16 + b, m * 2); add(1, in->dst_row16 + b, m * 2); add(2, in->dst_ref16 + b, m * 2); add(3, in->src_pos32 + b, m * 4);
			add(4, in->seq_len + b, m * 4); add(5, in->cig_len16 + b, m * 2);
			if (has_qual) add(6, in->has_qual + b / 8, (m + 7) / 8);
			if (mates) { add(8, in->mate_row16 + b, m * 2); add(9, in->mate_pos32 + b, m * 4); }
		}
		uint64_t x0 = 0, x1 = 0;                                      // this chunk's N exceptions (sorted by record)
		if (two_bit) {
			if (in->n_seq_exceptions) {
				x0 = (uint64_t) (std::lower_bound(in->exc_rec, in->exc_rec + in->n_seq_exceptions, (uint32_t) b) - in->exc_rec);
				x1 = (uint64_t) (std::lower_bound(in->exc_rec, in->exc_rec + in->n_seq_exceptions, (uint32_t) e) - in->exc_rec);
			}
			add(13, in->seq + s0, s1 - s0);
			if (x1 > x0) { add(14, in->exc_rec + x0, (x1 - x0) * 4); add(15, in->exc_off + x0, (x1 - x0) * 4); }
			PV3_CUDA(ctx, ws.in[10].ensure_idle(2 * (s1 - s0) + 128, stream));
		} else add(10, in->seq + s0, s1 - s0);
		add(11, in->cig + c0, (c1 - c0) * 4);
		if (qual_bulk) add(12, in->qual + q0, q1 - q0);
		for (int i = 0; i < n_copies; ++i) PV3_CUDA(ctx, ws.in[copies[i].slot].ensure_idle((copies[i].bytes ? copies[i].bytes : 16) + 64, stream));
		PV3_CUDA(ctx, cudaEventRecord(ws.cev[0], stream));
		mark(c, 0, stream);
		for (int i = 0; i < n_copies; ++i)
			if (copies[i].bytes) PV3_CUDA(ctx, cudaMemcpyAsync(ws.in[copies[i].slot].p, copies[i].src, copies[i].bytes, cudaMemcpyHostToDevice, stream));
		PV3_CUDA(ctx, cudaEventRecord(ws.cev[1], stream));
		mark(c, 1, stream);

		// ---- device batch ------------------------------------------------------------------------------------------------
		DevBatch B;
		memset(&B.in, 0, sizeof B.in); memset(&B.out, 0, sizeof B.out);
		B.in.n = m;
		uint64_t const n_tiles = pack_tiles(m);
		PV3_CUDA(ctx, ws.scratch.ensure_idle((3 * n_tiles + 8) * 8 * 2, stream));
		if (!compact) {
			B.in.src_row = ws.in[0].as <uint32_t>(); B.in.dst_row = ws.in[1].as <uint32_t>(); B.in.dst_ref = ws.in[2].as <uint32_t>();
			B.in.src_pos = ws.in[3].as <uint64_t>(); B.in.seq_off = ws.in[4].as <uint64_t>(); B.in.cig_off = ws.in[5].as <uint64_t>();
			B.in.seq = ws.in[10].as <uint8_t>() - s0;                // offsets stay batch-global
			B.in.cig = ws.in[11].as <uint32_t>() - c0;
			if (has_qual) {
				B.in.qual_off = ws.in[6].as <uint64_t>();
				B.in.qual = qual_bulk ? ws.in[12].as <uint8_t>() - q0 : qual_alias;    // in place: the DP kernels read the few windows they need over PCIe
			}
			if (packed) B.in.seq_len = ws.in[7].as <uint32_t>();
			if (mates) { B.in.mate_row = ws.in[8].as <uint32_t>(); B.in.mate_pos = ws.in[9].as <uint64_t>(); }
		} else {
			size_t const sizes[10] = { m * 4, m * 4, m * 4, m * 8, (m + 1) * 8, (m + 1) * 8, (m + 1) * 8, m * 4, m * 8, 0 };
			for (int i = 0; i < 9; ++i) PV3_CUDA(ctx, ws.un[i].ensure_idle(sizes[i] + 64, stream));
			UnpackParams U;
			memset(&U, 0, sizeof U);
			U.n = m;
			U.src_row16 = ws.in[0].as <uint16_t>(); U.dst_row16 = ws.in[1].as <uint16_t>(); U.dst_ref16 = ws.in[2].as <uint16_t>();
			U.src_pos32 = ws.in[3].as <uint32_t>(); U.seq_len = ws.in[4].as <uint32_t>(); U.cig_len16 = ws.in[5].as <uint16_t>();
			U.has_qual = has_qual ? ws.in[6].as <uint8_t>() : nullptr;
			U.mate_row16 = mates ? ws.in[8].as <uint16_t>() : nullptr; U.mate_pos32 = mates ? ws.in[9].as <uint32_t>() : nullptr;
			U.seq_packed = two_bit ? 2 : packed ? 1 : 0;
			U.src_row = ws.un[0].as <uint32_t>(); U.dst_row = ws.un[1].as <uint32_t>(); U.dst_ref = ws.un[2].as <uint32_t>();
			U.src_pos = ws.un[3].as <uint64_t>(); U.seq_off = ws.un[4].as <uint64_t>(); U.cig_off = ws.un[5].as <uint64_t>();
			U.qual_off = has_qual ? ws.un[6].as <uint64_t>() : nullptr;
			U.mate_row = ws.un[7].as <uint32_t>(); U.mate_pos = ws.un[8].as <uint64_t>();
			U.tile_sums = ws.scratch.as <uint64_t>();
			launch_unpack_in(U, stream);
			PV3_CUDA(ctx, cudaGetLastError());
			hc.launches += 3;
			if (two_bit) {
				launch_expand_seq2(ws.in[13].as <uint8_t>(), s1 - s0, ws.in[10].as <uint8_t>(), ws.in[14].as <uint32_t>(), ws.in[15].as <uint32_t>(), x1 - x0,
					b, U.seq_off, stream);
				PV3_CUDA(ctx, cudaGetLastError());
				hc.launches += (s1 > s0 ? 1 : 0) + (x1 > x0 ? 1 : 0);
			}
			B.in.src_row = U.src_row; B.in.dst_row = U.dst_row; B.in.dst_ref = U.dst_ref; B.in.src_pos = U.src_pos;
			B.in.seq_off = U.seq_off; B.in.cig_off = U.cig_off;          // chunk-relative: they start at 0
			B.in.seq = ws.in[10].as <uint8_t>(); B.in.cig = ws.in[11].as <uint32_t>();
			if (has_qual) { B.in.qual_off = U.qual_off; B.in.qual = qual_bulk ? ws.in[12].as <uint8_t>() : qual_alias + q0; }
			if (packed) B.in.seq_len = ws.in[4].as <uint32_t>();
			if (mates) { B.in.mate_row = U.mate_row; B.in.mate_pos = U.mate_pos; }
		}
		uint64_t arena_cap = (c1 - c0) * 3 + m * 8 + 1024, rr_cap = m / 8 + 1024;
		if (arena_cap < ws.arena_cap_hint) arena_cap = ws.arena_cap_hint;
		if (rr_cap < ws.rr_cap_hint) rr_cap = ws.rr_cap_hint;
		ws.arena_cap_hint = arena_cap; ws.rr_cap_hint = rr_cap;
		PV3_CUDA(ctx, ws.out_dst_pos.ensure_idle(m * 8, stream)); PV3_CUDA(ctx, ws.out_flags.ensure_idle(m * 4, stream));
		PV3_CUDA(ctx, ws.out_cig_off.ensure_idle(m * 8, stream)); PV3_CUDA(ctx, ws.out_cig_len.ensure_idle(m * 4, stream));
		PV3_CUDA(ctx, ws.out_rr_off.ensure_idle(m * 8, stream)); PV3_CUDA(ctx, ws.out_rr_len.ensure_idle(m * 4, stream));
		PV3_CUDA(ctx, ws.out_arena.ensure_idle(arena_cap * 4, stream)); PV3_CUDA(ctx, ws.out_rr.ensure_idle(rr_cap * 32, stream));
		if (mates) PV3_CUDA(ctx, ws.out_mate.ensure_idle(m * 8, stream));
		B.out.dst_pos = ws.out_dst_pos.as <uint64_t>(); B.out.flags = ws.out_flags.as <uint32_t>();
		B.out.cig_off = ws.out_cig_off.as <uint64_t>(); B.out.cig_len = ws.out_cig_len.as <uint32_t>();
		B.out.cig = ws.out_arena.as <uint32_t>(); B.out.cig_cap = arena_cap;
		B.out.rr_off = ws.out_rr_off.as <uint64_t>(); B.out.rr_len = ws.out_rr_len.as <uint32_t>();
		B.out.rr = ws.out_rr.as <uint64_t>(); B.out.rr_cap = rr_cap;
		B.out.mate_dst_pos = mates ? ws.out_mate.as <uint64_t>() : nullptr;
		if ((s1 - s0) * (packed ? 2 : 1) > 192 * m) B.expect_deferred = m;      // long reads: none fits the thread kernel
		int st = ensure_work(ctx, ws, m, rr_cap);
		if (st) return st;
		st = enqueue_kernels(ctx, ws, B, hc.launches);
		if (st) return st;

		// ---- ordered output: after the previous chunk has taken its place ---------------------------------------------------
		PackParams K;
		memset(&K, 0, sizeof K);
		K.n = m; K.rec0 = b;
		K.dst_pos = B.out.dst_pos; K.flags = B.out.flags; K.cig_off = B.out.cig_off; K.cig_len = B.out.cig_len; K.arena = B.out.cig; K.arena_off = 0;
		K.rr_off = B.out.rr_off; K.rr_len = B.out.rr_len; K.rr = B.out.rr; K.rr_base = 0;
		K.mate_dst_pos = B.out.mate_dst_pos;
		K.counters = ws.counters.as <Counters>();
		K.running = ctx->running.as <uint64_t>();
		K.place = ws.place.as <uint64_t>();
		K.tile_sums = ws.scratch.as <uint64_t>() + (3 * n_tiles + 8);
		K.report = (ChunkReport *) mapped_alias(report);
		if (!K.report) return fail(ctx, PV3_ERR_CUDA, "pv3_project: report buffer is not mapped");
		K.out_cig_cap = hc.cig_room; K.out_rr_cap = hc.rr_room;
		K.cig_off_bias = hc.cig_place; K.rr_off_bias = hc.rr_place;
		{
			// Ordered output of the chunk in device staging laid out like the caller's arrays: record fields indexed from 0, the arenas
			// from the chunk's place. It leaves either with the coalesced copy kernel (pinned destination: no host round trip) or,
			// after a synchronisation, with plain copies (pageable destination).
			size_t const sizes[12] = { m * 8, m * 4, m * 4, m, m * 8, m * 4, m * 2, m * 8, m * 4, m, m * 8, m * 4 };
			for (int i = 0; i < 12; ++i) PV3_CUDA(ctx, ws.st[i].ensure_idle(sizes[i] + 64, stream));
			K.rec0 = 0;
			K.out_dst_pos = out->dst_pos ? ws.st[0].as <uint64_t>() : nullptr; K.out_dst_pos32 = out->dst_pos32 ? ws.st[1].as <uint32_t>() : nullptr;
			K.out_flags = out->flags ? ws.st[2].as <uint32_t>() : nullptr; K.out_flags8 = out->flags8 ? ws.st[3].as <uint8_t>() : nullptr;
			K.out_cig_off = out->cig_off ? ws.st[4].as <uint64_t>() : nullptr; K.out_cig_len = out->cig_len ? ws.st[5].as <uint32_t>() : nullptr;
			K.out_cig_len16 = out->cig_len16 ? ws.st[6].as <uint16_t>() : nullptr;
			K.out_rr_off = out->rr_off ? ws.st[7].as <uint64_t>() : nullptr; K.out_rr_len = out->rr_len ? ws.st[8].as <uint32_t>() : nullptr;
			K.out_rr_len8 = out->rr_len8 ? ws.st[9].as <uint8_t>() : nullptr;
			K.out_mate_dst_pos = out->mate_dst_pos ? ws.st[10].as <uint64_t>() : nullptr; K.out_mate_dst_pos32 = out->mate_dst_pos32 ? ws.st[11].as <uint32_t>() : nullptr;
			PV3_CUDA(ctx, ws.un[9].ensure_idle(arena_cap * 4 + rr_cap * 32 + 128, stream));
			K.out_cig = ws.un[9].as <uint32_t>(); K.out_rr = (uint64_t *) (ws.un[9].as <uint8_t>() + align_up(arena_cap * 4 + 16, 32));
			K.staged = 1;
			K.cig_dst_word0 = zero_copy ? (((uint64_t) (uintptr_t) d_cig.alias) >> 2) + hc.cig_place : 0;
		}
		mark(c, 2, stream);
		PV3_CUDA(ctx, cudaStreamWaitEvent(stream, c == 0 ? ctx->ev_misc[2] : ctx->sets[(c - 1) % NS].ev[5], 0));
		PV3_CUDA(ctx, cudaEventRecord(ws.cev[2], stream));
		mark(c, 3, stream);
		launch_pack_out(K, stream);
		PV3_CUDA(ctx, cudaGetLastError());
		hc.launches += 4;
		PV3_CUDA(ctx, cudaEventRecord(ws.ev[5], stream));
		mark(c, 4, stream);
		if (zero_copy) {
			StreamOutParams S;
			memset(&S, 0, sizeof S);
			auto seg = [&](void *alias, const void *src, size_t elem) {
				if (!alias || !src) return;
				S.seg[S.n_seg].dst = (uint8_t *) alias + b * elem; S.seg[S.n_seg].src = src; S.seg[S.n_seg].bytes = m * elem; ++S.n_seg;
			};
			seg(d_dst_pos.alias, K.out_dst_pos, 8); seg(d_dst_pos32.alias, K.out_dst_pos32, 4); seg(d_flags.alias, K.out_flags, 4); seg(d_flags8.alias, K.out_flags8, 1);
			seg(d_cig_off.alias, K.out_cig_off, 8); seg(d_cig_len.alias, K.out_cig_len, 4); seg(d_cig_len16.alias, K.out_cig_len16, 2);
			seg(d_rr_off.alias, K.out_rr_off, 8); seg(d_rr_len.alias, K.out_rr_len, 4); seg(d_rr_len8.alias, K.out_rr_len8, 1);
			seg(d_mate.alias, K.out_mate_dst_pos, 8); seg(d_mate32.alias, K.out_mate_dst_pos32, 4);
			S.cig_dst = d_cig.alias ? (uint32_t *) d_cig.alias + hc.cig_place : nullptr; S.cig_src = K.out_cig;
			S.rr_dst = d_rr.alias ? (uint64_t *) d_rr.alias + 4 * hc.rr_place : nullptr; S.rr_src = K.out_rr;
			S.place = K.place;
			launch_stream_out(S, stream);
			PV3_CUDA(ctx, cudaGetLastError());
			hc.launches += 1;
		}
		PV3_CUDA(ctx, cudaEventRecord(ws.cev[3], stream));
		mark(c, 5, stream);

		if (!zero_copy) {
			// pageable destination: wait for the chunk, then plain copies of exactly what it produced
			PV3_CUDA(ctx, cudaStreamSynchronize(stream));
			if (report->overflow) continue;                      // the call is repeated with larger buffers
			uint64_t const cb = report->cig_base, rb = report->rr_base;
			struct Back { void *dst; const void *src; size_t bytes; };
			Back const backs[] = {
				{ out->dst_pos ? out->dst_pos + b : nullptr, K.out_dst_pos, m * 8 }, { out->dst_pos32 ? out->dst_pos32 + b : nullptr, K.out_dst_pos32, m * 4 },
				{ out->flags ? out->flags + b : nullptr, K.out_flags, m * 4 }, { out->flags8 ? out->flags8 + b : nullptr, K.out_flags8, m },
				{ out->cig_off ? out->cig_off + b : nullptr, K.out_cig_off, m * 8 }, { out->cig_len ? out->cig_len + b : nullptr, K.out_cig_len, m * 4 },
				{ out->cig_len16 ? out->cig_len16 + b : nullptr, K.out_cig_len16, m * 2 },
				{ out->rr_off ? out->rr_off + b : nullptr, K.out_rr_off, m * 8 }, { out->rr_len ? out->rr_len + b : nullptr, K.out_rr_len, m * 4 },
				{ out->rr_len8 ? out->rr_len8 + b : nullptr, K.out_rr_len8, m },
				{ out->mate_dst_pos ? out->mate_dst_pos + b : nullptr, K.out_mate_dst_pos, m * 8 }, { out->mate_dst_pos32 ? out->mate_dst_pos32 + b : nullptr, K.out_mate_dst_pos32, m * 4 },
				{ out->cig ? out->cig + hc.cig_place + cb : nullptr, K.out_cig + report->cig_pad, report->cig_used * 4 },
				{ out->rr ? out->rr + 4 * (hc.rr_place + rb) : nullptr, K.out_rr, report->rr_used * 32 },
			};
			for (auto const &bk : backs)
				if (bk.dst && bk.bytes) PV3_CUDA(ctx, cudaMemcpyAsync(bk.dst, bk.src, bk.bytes, cudaMemcpyDeviceToHost, stream));
		}
	}
	return PV3_OK;
}

// After the streams have drained: reads the chunk reports; fills the caller's totals. Returns PV3_OK, a capacity error for
// the caller, or 1 when the call has to be repeated with the buffer sizes the reports asked for.
int host_finish(HostCall &hc, uint64_t &cig_used, uint64_t &rr_used)
{
	pv3_ctx *ctx = hc.ctx;
	pv3_batch_out *out = hc.out;
	bool again = false, caller_short = false, narrow = false;
	cig_used = rr_used = 0;
	uint64_t flagged = 0, realigned = 0, cells = 0, n_deferred = 0;
	for (uint64_t c = 0; c < hc.n_chunks; ++c) {
		ChunkReport const &r = ctx->h_reports[c];
		WorkSet &ws = ctx->sets[c % pv3_ctx::N_SETS];
		uint64_t const m = hc.cuts[c + 1] - hc.cuts[c];
		if (!m) continue;
		if (!r.done) return fail(ctx, PV3_ERR_CUDA, "pv3_project: chunk %llu did not report", (unsigned long long) c);
		if (digest(ws, m, r.overflow, r.arena_need, r.jobs, r.dp_path, r.dp_max_pool, r.seg_need, r.seg_ops_need, r.n_deferred)) again = true;
		if (r.overflow & (16u | 32u)) caller_short = true;
		if (r.overflow & 64u) narrow = true;
		cig_used += r.cig_used; rr_used += r.rr_used;
		flagged += r.n_flagged; realigned += r.n_realigned; cells += r.dp_cells; n_deferred += r.n_deferred;
	}
	if (again) return 1;
	out->n_flagged += flagged; out->n_realigned += realigned; out->dp_cells += cells;
	ctx->timing.n_deferred = n_deferred;
	if (caller_short) {
		// the running totals stop at the first chunk that did not fit: sum what every chunk needs
		return fail(ctx, PV3_ERR_CAPACITY, "host output capacity too small: need cig_cap >= %llu, rr_cap >= %llu",
			(unsigned long long) (hc.cig_place + cig_used), (unsigned long long) (hc.rr_place + rr_used));
	}
	if (narrow) return fail(ctx, PV3_ERR_CAPACITY, "a value does not fit the narrow output field chosen (dst_pos32 / flags8 / cig_len16 / rr_len8): use the wide field");
	return PV3_OK;
}

void host_timing(HostCall &hc)
{
	pv3_ctx *ctx = hc.ctx;
	float ms[5] = { 0, 0, 0, 0, 0 }, h2d = 0, d2h = 0, t = 0;
	// events are reused by the chunks of a set: only the last chunk of every set is still on record
	for (int k = 0; k < pv3_ctx::N_SETS && (uint64_t) k < hc.n_chunks; ++k) {
		WorkSet &ws = ctx->sets[k];
		harvest_timing(ctx, ws, ms);
		if (cudaEventElapsedTime(&t, ws.cev[0], ws.cev[1]) == cudaSuccess) h2d += t;
		if (cudaEventElapsedTime(&t, ws.cev[2], ws.cev[3]) == cudaSuccess) d2h += t;
	}
	ctx->timing.ms_rewrite = ms[0]; ctx->timing.ms_realign = ms[1]; ctx->timing.ms_splice = ms[2]; ctx->timing.ms_total = ms[3];
	ctx->timing.ms_rewrite_main = ms[4]; ctx->timing.ms_h2d = h2d; ctx->timing.ms_d2h = d2h;
	ctx->timing.launches = hc.launches;
	ctx->total_launches += hc.launches;
}

int validate_host_batch(pv3_ctx *ctx, const pv3_batch_in *in, pv3_batch_out *out, const char *who)
{
	if (!in || !out) return fail(ctx, PV3_ERR_INVALID, "%s: NULL batch", who);
	if (!in->n) return PV3_OK;
	if (!in->seq_off || !in->cig_off) return fail(ctx, PV3_ERR_INVALID, "%s: seq_off and cig_off must be valid on the host", who);
	if (in->compact) {
		if (!in->src_row16 || !in->dst_row16 || !in->dst_ref16 || !in->src_pos32 || !in->seq_len || !in->cig_len16)
			return fail(ctx, PV3_ERR_INVALID, "%s: compact encoding needs src_row16, dst_row16, dst_ref16, src_pos32, seq_len, cig_len16", who);
		if (ctx->rows.size() > 65536 || ctx->refs.size() > 65536) return fail(ctx, PV3_ERR_INVALID, "%s: compact encoding holds 16-bit ids", who);
		if (in->seq_2bit && in->seq_bam4) return fail(ctx, PV3_ERR_INVALID, "%s: seq_2bit and seq_bam4 exclude each other", who);
		if (in->seq_2bit && in->n_seq_exceptions && (!in->exc_rec || !in->exc_off)) return fail(ctx, PV3_ERR_INVALID, "%s: exc_rec and exc_off are required with n_seq_exceptions", who);
		if (in->seq_2bit && in->n > 0xffffffffull) return fail(ctx, PV3_ERR_INVALID, "%s: two-bit SEQ addresses records with 32 bits", who);
	} else if (in->seq_2bit) {
		return fail(ctx, PV3_ERR_INVALID, "%s: seq_2bit needs the compact encoding", who);
	}
	if (!in->compact && (!in->src_row || !in->dst_row || !in->dst_ref || !in->src_pos)) return fail(ctx, PV3_ERR_INVALID, "%s: NULL input array", who);
	if (!(out->dst_pos || out->dst_pos32) || !(out->cig_len || out->cig_len16) || !out->cig)
		return fail(ctx, PV3_ERR_INVALID, "%s: dst_pos, cig_len (either width) and cig are required outputs", who);
	return PV3_OK;
}

} // namespace

extern "C" {

int pv3_project(pv3_ctx *ctx, const pv3_batch_in *in, pv3_batch_out *out)
{
	// Host pipeline (replaces the 16-slot SPSC ring of 4-record tasks, include/panvc3/spsc_queue.hh): the batch is cut
	// into chunks; chunk i+1's H2D copy runs on the other stream while chunk i's kernels run and chunk i-1's results are
	// written back, so PCIe transfers in both directions overlap the compute.
	int st = check_ready(ctx, "pv3_project");
	if (st) return st;
	st = validate_host_batch(ctx, in, out, "pv3_project");
	if (st) return st;
	out->cig_used = out->rr_used = out->n_flagged = out->n_realigned = out->dp_cells = 0;
	memset(&ctx->timing, 0, sizeof ctx->timing);
	if (!in->n) return PV3_OK;
	for (int attempt = 0; attempt < 6; ++attempt) {
		HostCall hc;
		hc.ctx = ctx; hc.in = in; hc.out = out; hc.rec_begin = 0; hc.rec_end = in->n;
		hc.cig_place = hc.rr_place = 0; hc.cig_room = out->cig_cap; hc.rr_room = out->rr ? out->rr_cap : 0;
		static bool const trace = getenv("PV3_TRACE") != nullptr;
		auto const t0 = std::chrono::steady_clock::now();
		st = host_enqueue(hc);
		auto const t1 = std::chrono::steady_clock::now();
		// whatever happened, nothing of this call may still be in flight when it returns: the caller owns the buffers again
		cudaError_t e0 = cudaStreamSynchronize(ctx->stream), e1 = cudaStreamSynchronize(ctx->stream2);
		{ cudaError_t const e2 = cudaStreamSynchronize(ctx->stream3); if (e1 == cudaSuccess) e1 = e2; }
		if (trace) {
			auto const t2 = std::chrono::steady_clock::now();
			dump_timeline(hc.n_chunks);
			fprintf(stderr, "[pv3 trace] pv3_project: %llu records, %llu launches; host enqueue %.3f ms, wait %.3f ms\n", (unsigned long long) in->n,
				(unsigned long long) hc.launches, std::chrono::duration <double, std::milli>(t1 - t0).count(), std::chrono::duration <double, std::milli>(t2 - t1).count());
		}
		if (st) return st;
		if (e0 != cudaSuccess || e1 != cudaSuccess) return fail(ctx, PV3_ERR_CUDA, "pv3_project: %s", cudaGetErrorString(e0 != cudaSuccess ? e0 : e1));
		uint64_t cig_used = 0, rr_used = 0;
		st = host_finish(hc, cig_used, rr_used);
		out->cig_used = cig_used; out->rr_used = rr_used;
		host_timing(hc);
		if (st == 1) { out->n_flagged = out->n_realigned = out->dp_cells = 0; continue; }
		return st;
	}
	return fail(ctx, PV3_ERR_CUDA, "pv3_project: internal buffers could not be sized");
}

} // extern "C"

// ---- several GPUs behind one submitter -------------------------------------------------------------------------------

struct pv3_multi {
	std::vector <pv3_ctx *> ctx;
	std::string error;
};

extern "C" {

int pv3_multi_create(pv3_multi **out, int n_devices, const int *devices)
{
	if (!out || n_devices < 1) return PV3_ERR_INVALID;
	*out = nullptr;
	pv3_multi *m = new pv3_multi;
	for (int i = 0; i < n_devices; ++i) {
		pv3_ctx *c = nullptr;
		int const st = pv3_create(&c, devices ? devices[i] : i);
		if (st) { for (auto x : m->ctx) pv3_destroy(x); delete m; return st; }
		m->ctx.push_back(c);
	}
	*out = m;
	return PV3_OK;
}

void pv3_multi_destroy(pv3_multi *m)
{
	if (!m) return;
	for (auto c : m->ctx) pv3_destroy(c);
	delete m;
}

int pv3_multi_devices(const pv3_multi *m) { return m ? (int) m->ctx.size() : 0; }
pv3_ctx *pv3_multi_ctx(pv3_multi *m, int i) { return (m && i >= 0 && (size_t) i < m->ctx.size()) ? m->ctx[i] : nullptr; }
const char *pv3_multi_last_error(const pv3_multi *m) { return m ? m->error.c_str() : ""; }

// Shard g = records [n g / G, n (g + 1) / G) on device g. Every shard's chunks are enqueued from this thread (asynchronous),
// then all devices are drained: one synchronisation per device and call. The ordered arenas of shard g start at
// cig_cap g / G (rr likewise); cig_off / rr_off, which must be given, are absolute.
int pv3_multi_project(pv3_multi *m, const pv3_batch_in *in, pv3_batch_out *out)
{
	if (!m || m->ctx.empty()) return PV3_ERR_INVALID;
	size_t const G = m->ctx.size();
	auto failm = [&](int code, pv3_ctx *c) { m->error = pv3_last_error(c); return code; };
	for (auto c : m->ctx) { int st = check_ready(c, "pv3_multi_project"); if (st) return failm(st, c); }
	int st = validate_host_batch(m->ctx[0], in, out, "pv3_multi_project");
	if (st) return failm(st, m->ctx[0]);
	if (G > 1 && in->n && (!out->cig_off || (out->rr && !out->rr_off))) { fail(m->ctx[0], PV3_ERR_INVALID, "pv3_multi_project: cig_off and rr_off are required with more than one device"); return failm(PV3_ERR_INVALID, m->ctx[0]); }
	out->cig_used = out->rr_used = out->n_flagged = out->n_realigned = out->dp_cells = 0;
	if (!in->n) return PV3_OK;
	for (int attempt = 0; attempt < 6; ++attempt) {
		std::vector <HostCall> hc(G);
		int first_err = PV3_OK; pv3_ctx *err_ctx = nullptr;
		for (size_t g = 0; g < G; ++g) {
			pv3_ctx *c = m->ctx[g];
			memset(&c->timing, 0, sizeof c->timing);
			hc[g].ctx = c; hc[g].in = in; hc[g].out = out;
			hc[g].rec_begin = (in->n * g / G) / 128 * 128; hc[g].rec_end = g + 1 == G ? in->n : (in->n * (g + 1) / G) / 128 * 128;
			hc[g].cig_place = out->cig_cap * g / G; hc[g].cig_room = out->cig_cap * (g + 1) / G - hc[g].cig_place;
			hc[g].rr_place = out->rr ? out->rr_cap * g / G : 0; hc[g].rr_room = out->rr ? out->rr_cap * (g + 1) / G - hc[g].rr_place : 0;
			if (hc[g].rec_end > hc[g].rec_begin && !first_err) { first_err = host_enqueue(hc[g]); if (first_err) err_ctx = c; }
		}
		for (size_t g = 0; g < G; ++g) {
			pv3_ctx *c = m->ctx[g];
			cudaSetDevice(c->device);
			cudaError_t e0 = cudaStreamSynchronize(c->stream), e1 = cudaStreamSynchronize(c->stream2);
			{ cudaError_t const e2 = cudaStreamSynchronize(c->stream3); if (e1 == cudaSuccess) e1 = e2; }
			if ((e0 != cudaSuccess || e1 != cudaSuccess) && !first_err) { first_err = fail(c, PV3_ERR_CUDA, "pv3_multi_project: %s", cudaGetErrorString(e0 != cudaSuccess ? e0 : e1)); err_ctx = c; }
		}
		if (first_err) return failm(first_err, err_ctx);
		bool again = false;
		uint64_t cig_need = 0, rr_need = 0;
		int cap_err = PV3_OK; pv3_ctx *cap_ctx = nullptr;
		out->cig_used = out->rr_used = out->n_flagged = out->n_realigned = out->dp_cells = 0;
		for (size_t g = 0; g < G; ++g) {
			if (hc[g].rec_end <= hc[g].rec_begin) continue;
			uint64_t cu = 0, ru = 0;
			st = host_finish(hc[g], cu, ru);
			host_timing(hc[g]);
			if (st == 1) again = true;
			else if (st) { cap_err = st; cap_ctx = hc[g].ctx; }
			cig_need = std::max <uint64_t>(cig_need, cu * G + G); rr_need = std::max <uint64_t>(rr_need, ru * G + G);
			out->cig_used = std::max <uint64_t>(out->cig_used, hc[g].cig_place + cu); out->rr_used = std::max <uint64_t>(out->rr_used, hc[g].rr_place + ru);
		}
		if (again) continue;
		if (cap_err == PV3_ERR_CAPACITY) { out->cig_used = cig_need; out->rr_used = rr_need; }   // every shard must fit its G-th of the arenas
		if (cap_err) return failm(cap_err, cap_ctx);
		return PV3_OK;
	}
	return failm(fail(m->ctx[0], PV3_ERR_CUDA, "pv3_multi_project: internal buffers could not be sized"), m->ctx[0]);
}

} // extern "C"
