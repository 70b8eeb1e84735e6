// panvc3_b200 — indel-run scan shared by both rewrite kernels.
#pragma once

#include "pv3_device.cuh"
#include "pv3_kernels.h"

namespace pv3 {

// indel_run_checker::find_next_range_for_realigning over the rewritten ops (I S H P D = X only).
// Scalar scan; op k is read from ops[k * stride]. Used warp-uniformly by the warp-per-record kernel (all lanes read the
// same address, i_write = lane 0) and privately by the thread-per-record kernel. WRITE = false counts runs and sums the
// DP work; WRITE = true stores the Job and the reported ranges of each run (only where i_write).
template <bool WRITE>
__device__ __forceinline__ uint32_t scan_runs(const uint32_t *ops, uint32_t stride, uint32_t n_ops, uint64_t dst_pos, uint64_t rec, uint64_t nq, uint64_t ref_n,
	Job *jobs, uint64_t *rr, uint32_t &flags, unsigned long long &cells, unsigned long long &path,
	unsigned long long &pool, unsigned long long &max_pool, bool affine, bool i_write,
	uint32_t *n_big = nullptr)
{
	uint64_t ref = dst_pos, qry = 0, run_ref = dst_pos, run_qry = 0;
	uint32_t run_k = 0, type = 0, n_runs = 0;
	for (uint32_t k = 0; k <= n_ops; ++k) {
		uint64_t const ref0 = ref, qry0 = qry;
		uint32_t const type0 = type;
		if (k < n_ops) {
			uint32_t const v = ops[(size_t) k * stride], n = v >> 4, op = v & 0xfu;
			if (!type) { run_ref = ref; run_qry = qry; run_k = k; }       // indel_run_checker.cc:44-49
			if (op == OP_I) { qry += n; type |= 1u; continue; }            // :60-63
			if (op == OP_D) { ref += n; type |= 2u; continue; }            // :65-68
			if (op == OP_S) qry += n;                                      // :70-75
			else if (op == OP_EQ || op == OP_X || op == OP_M) { ref += n; qry += n; }  // :84-92
			else if (op == OP_N) ref += n;                                 // :77-82 (never produced by the rewrite)
			type = 0;                                                      // H, P: :54-58
		}
		if (type0 == 3u) {                                                 // report_run :104-108 / final run :111-118
			uint64_t const rlen = ref0 - run_ref, qlen = qry0 - run_qry;
			if (run_ref + rlen > ref_n) flags |= PV3_FLAG_REF_OOB;
			if (run_qry + qlen > nq) flags |= PV3_FLAG_QUERY_OOB;
			if (rlen > 0x7fffffffull || qlen > 0x7fffffffull) flags |= PV3_FLAG_REF_OOB;
			if (WRITE) {
				if (i_write) {
					Job j;
					j.rec = rec; j.ref_loc = run_ref; j.ref_len = (uint32_t) rlen; j.qry_loc = (uint32_t) run_qry;
					j.qry_len = (uint32_t) qlen; j.op_begin = run_k; j.op_end = k; j.out_len = 0; j.out_off = 0;
					jobs[n_runs] = j;
					// alignment_projector.cc:69-73 — location is the live m_ref_pos, already advanced over the
					// run-terminating op; the length is the true window length.
					rr[4 * n_runs + 0] = ref; rr[4 * n_runs + 1] = rlen; rr[4 * n_runs + 2] = run_qry; rr[4 * n_runs + 3] = qlen;
				}
			} else {
				cells += rlen * qlen;
				path += rlen + qlen;
				unsigned long long const need = dp_pool_bytes((uint32_t) rlen, (uint32_t) qlen, affine);
				pool += need;
				if (need > max_pool) max_pool = need;
				if (n_big && !dp_job_is_small((uint32_t) rlen, (uint32_t) qlen)) ++*n_big;
			}
			++n_runs;
			type = (k < n_ops) ? type : 0;
		}
		if (k == n_ops) break;
	}
	return n_runs;
}

} // namespace pv3
