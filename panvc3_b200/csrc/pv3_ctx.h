// panvc3_b200 — the context behind the C-ABI (internal; shared by pv3_api.cu and pv3_pipeline.cu).
#pragma once

#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include "pv3_kernels.h"

namespace pv3 {

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
	// for buffers that work in flight on `stream` may still use: the stream is drained before the buffer is replaced
	cudaError_t ensure_idle(size_t bytes, cudaStream_t stream)
	{
		if (bytes <= cap) return cudaSuccess;
		cudaError_t e = cudaStreamSynchronize(stream);
		return e != cudaSuccess ? e : ensure(bytes);
	}
	void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
	template <typename T> T *as() const { return (T *) p; }
};

struct HostRow { std::vector <uint64_t> words; uint64_t n_bits = 0; DevBuf mem; bool uploaded = false; };
struct HostRef { std::string bases; uint64_t n = 0; DevBuf mem; bool uploaded = false; bool live = true; };

// Device working set of one stream: everything a batch (or a chunk of the host pipeline) needs besides the index, so that two
// chunks can be in flight on two streams.
struct WorkSet {
	cudaStream_t stream = nullptr;
	cudaEvent_t  ev[6] = {};            // 0 kernels start, 1 rewrite end, 2 realign end, 3 splice end, 4 main rewrite kernel end, 5 pack end
	cudaEvent_t  cev[4] = {};           // H2D begin / end, write-back begin / end
	DevBuf   counters, jobs, dp_arena, pool, deferred, complex_list, complex_ce, big_list, place, scratch;
	DevBuf   deferred2, plan_rec, plan_seg0, plan_nseg, segs, seg_out, seg_ops;       // long records as segments
	uint64_t seg_cap = 0, seg_ops_cap = 0, seg_cap_hint = 0, seg_ops_hint = 0;
	uint64_t pool_slot = 0, pool_slots = 0;
	// chunk inputs (as uploaded), unpacked inputs (compact encoding), per-record outputs, arenas, staged ordered output
	DevBuf   in[16], un[10], out_dst_pos, out_flags, out_cig_off, out_cig_len, out_arena, out_rr_off, out_rr_len, out_rr, out_mate;
	DevBuf   st[12];
	uint64_t arena_cap_hint = 0, rr_cap_hint = 0, dp_arena_hint = 0, pool_hint = 0;
	bool     qual_bulk = false;         // the last chunk here realigned many records: upload QUAL in full instead of reading it in place
	int      k1t_layout = -1;           // layout of this set's last thread-kernel launch (-1: not tunable) and its records
	uint64_t k1t_records = 0;
	bool     many_runs = false;
	uint64_t expect_deferred = 0;       // the last batch's deferred records (+ 25 %): sizes the grids of the deferred-record kernels
	void release();
};

} // namespace pv3

struct pv3_ctx {
	int          device = 0;
	int          n_sms = 148;
	cudaStream_t stream = nullptr;       // the stream all work is issued on
	cudaStream_t own_stream = nullptr;
	cudaStream_t stream2 = nullptr, stream3 = nullptr;      // further streams of the chunked host pipeline
	// Work sets of the host pipeline, used round robin by the chunks of a call: with two, the upload of chunk c + 2 waited for
	// chunk c's write-back (same stream) and the copy engine idled once uploads were short (two-bit SEQ); three keep it busy.
	static constexpr int N_SETS = 3;
	pv3::WorkSet sets[N_SETS];
	cudaEvent_t  ev_misc[4] = {};
	std::string  error;

	std::vector <pv3::HostRow> rows;
	std::vector <pv3::HostRef> refs;
	std::vector <int> free_refs;         // contig ids released by pv3_index_remove_ref
	bool         committed = false;
	pv3::DevBuf  row_desc, ref_desc;
	std::vector <pv3::RowDesc> h_row_desc;
	std::vector <pv3::RefDesc> h_ref_desc;
	uint64_t     index_bytes = 0;
	// Thread kernel layout for a large target, chosen by measurement: ns per record of the two layouts (moving average of the
	// kernel's event-timed launches), the other one probed again every 32nd launch and whenever a layout's time moves by more
	// than 15 % (the input changed, e.g. from scattered to coordinate-sorted records).
	struct LayoutTuner { float ns_per_rec[2] = { 0, 0 }; uint32_t seen[2] = { 0, 0 }; uint32_t launches = 0; bool reprobe = false; } tuner;
	uint64_t     target_code_bytes = 0;   // packed target codes of all references (chooses the thread kernel's layout)

	int32_t      gap_open = 0, gap_ext = -1;   // project-alignments/cmdline.ggo defaults
	int32_t      lut[84];
	int          stage = 0;
	uint64_t     chunk_records = 250000;     // records per chunk of the host pipeline (PV3_CHUNK_RECORDS overrides)
	uint64_t     chunk_edge_last = 0;        // records of the last chunk when it should differ from the first (PV3_CHUNK_EDGE_LAST; 0 = chunk_edge)
	uint64_t     chunk_edge = 125000;        // records of the first and the last chunk: a short pipeline fill and drain (PV3_CHUNK_EDGE; 0 = equal chunks)
	uint64_t     dp_thread_min = 16384;      // fewer small runs than this: the warp-per-run kernel takes them all
	int          kernel_path = 0;            // 0: thread-per-record run-walk kernel + warp-per-record for deferred records; 1: warp-per-record only;
	                                         // 2: thread-per-record word-parallel kernel + warp-per-record for deferred records

	pv3::DevBuf  ex_text, ex_words, ex_seq, ex_status;   // pv3_msa_row_extract
	pv3::DevBuf  an[16];                                 // pv3_format_oa / pv3_format_cigars
	pv3::DevBuf  q_row, q_row2, q_pos, q_out;            // position / index queries
	pv3::DevBuf  running;                                // [2] ordered-output totals of the chunks of a call
	pv3::Counters    *h_counters = nullptr;              // pinned
	pv3::ChunkReport *h_reports = nullptr;               // pinned, mapped: one per chunk of a host call
	void        *multi_owner = nullptr;

	pv3_timing   timing = {};
	uint64_t     total_launches = 0;
};

namespace pv3 {

constexpr int MAX_CHUNKS = 66;

int fail(pv3_ctx *ctx, int code, const char *fmt, ...);

#define PV3_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
	return pv3::fail((ctx), PV3_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); } while (0)

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

cudaError_t small_copy(void *dst, const void *src, size_t bytes, cudaStream_t stream);
int check_ready(pv3_ctx *ctx, const char *who);

} // namespace pv3
