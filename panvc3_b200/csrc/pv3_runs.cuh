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

// The same scan by a whole warp, 32 ops per step (the warp-per-record kernel and the long-read assembly: a 10-kb read has
// hundreds of ops, and the scalar scan above — executed by 32 lanes in lockstep, twice — was 80 % of the assembly kernel).
// Lane l looks at op base + l (the position past the last op counts as a non-indel op). Positions before each op come from
// two warp scans; a run ends at a non-indel op whose predecessor is an I or a D, starts right after the last non-indel op
// before it (bit tricks on the ballot of indel ops, or the state carried over from the previous 32 ops), and is reported when
// the stretch holds both an I and a D. Runs keep their order: a run's index is the number of runs ending before it.
// All lanes must call it; results (return value, flags and the sums) are uniform.
template <bool WRITE>
__device__ __forceinline__ uint32_t scan_runs_warp(const uint32_t *ops, uint32_t n_ops, uint64_t dst_pos, uint64_t rec, uint64_t nq, uint64_t ref_n,
	Job *jobs, uint64_t *rr, uint32_t &flags, unsigned long long &cells, unsigned long long &path,
	unsigned long long &pool, unsigned long long &max_pool, bool affine, uint32_t *n_big = nullptr)
{
	unsigned const lane = lane_id();
	unsigned const below = (1u << lane) - 1u;
	uint64_t ref_c = dst_pos, qry_c = 0;                  // positions before the first op of the current 32
	bool open = false;                                    // a stretch of I / D ops runs into the current 32 ...
	uint32_t open_type = 0, open_k = 0;                   // ... holding these kinds, begun at this op ...
	uint64_t open_ref = 0, open_qry = 0;                  // ... and these positions
	uint32_t n_runs = 0, my_flags = 0, my_big = 0;
	unsigned long long my_cells = 0, my_path = 0, my_pool = 0, my_max = 0;
	for (uint32_t base = 0; base <= n_ops; base += 32) {
		uint32_t const k = base + lane;
		bool const real = k < n_ops, item = k <= n_ops;
		uint32_t const v = real ? ops[k] : 0u, n = v >> 4, op = v & 0xfu;
		bool const is_i = real && op == OP_I, is_d = real && op == OP_D, indel = is_i || is_d;
		uint64_t dr = 0, dq = 0;
		if (real) {
			if (op == OP_I || op == OP_S) dq = n;                                   // indel_run_checker.cc:60-63, :70-75
			else if (op == OP_D || op == OP_N) dr = n;                              // :65-68, :77-82
			else if (op == OP_EQ || op == OP_X || op == OP_M) { dr = n; dq = n; }   // :84-92   (H, P: nothing)
		}
		uint64_t ir = dr, iq = dq;
		#pragma unroll
		for (int d = 1; d < 32; d <<= 1) {
			uint64_t const a = shfl_up_u64(ir, d), b = shfl_up_u64(iq, d);
			if ((int) lane >= d) { ir += a; iq += b; }
		}
		uint64_t const ref_before = ref_c + ir - dr, qry_before = qry_c + iq - dq;
		unsigned const m_i = __ballot_sync(FULL, is_i), m_d = __ballot_sync(FULL, is_d), m_ind = m_i | m_d;
		bool const prev_indel = lane ? ((m_ind >> (lane - 1)) & 1u) != 0 : open;
		bool const is_end = item && !indel && prev_indel;
		unsigned const zeros_below = ~m_ind & below;
		int const a = zeros_below ? 32 - __clz((int) zeros_below) : 0;     // first op of the stretch inside these 32
		bool const from_carry = !zeros_below && open;
		unsigned const range = below & ~((1u << a) - 1u);                    // ops [a, lane)
		uint32_t const type = ((m_i & range) ? 1u : 0u) | ((m_d & range) ? 2u : 0u) | (from_carry ? open_type : 0u);
		uint64_t s_ref = shfl_u64(ref_before, a), s_qry = shfl_u64(qry_before, a);
		uint32_t s_k = base + (uint32_t) a;
		if (from_carry) { s_ref = open_ref; s_qry = open_qry; s_k = open_k; }
		bool const is_run = is_end && type == 3u;                            // report_run :104-108 / final run :111-118
		unsigned const m_run = __ballot_sync(FULL, is_run);
		if (is_run) {
			uint64_t const rlen = ref_before - s_ref, qlen = qry_before - s_qry;
			if (s_ref + rlen > ref_n) my_flags |= PV3_FLAG_REF_OOB;
			if (s_qry + qlen > nq) my_flags |= PV3_FLAG_QUERY_OOB;
			if (rlen > 0x7fffffffull || qlen > 0x7fffffffull) my_flags |= PV3_FLAG_REF_OOB;
			if (WRITE) {
				uint32_t const at = n_runs + __popc(m_run & below);
				Job j;
				j.rec = rec; j.ref_loc = s_ref; j.ref_len = (uint32_t) rlen; j.qry_loc = (uint32_t) s_qry;
				j.qry_len = (uint32_t) qlen; j.op_begin = s_k; j.op_end = k; j.out_len = 0; j.out_off = 0;
				jobs[at] = j;
				// alignment_projector.cc:69-73 — location is the live m_ref_pos, already advanced over the run-terminating op (F10)
				rr[4 * at + 0] = ref_before + dr; rr[4 * at + 1] = rlen; rr[4 * at + 2] = s_qry; rr[4 * at + 3] = qlen;
			} else {
				my_cells += rlen * qlen;
				my_path += rlen + qlen;
				unsigned long long const need = dp_pool_bytes((uint32_t) rlen, (uint32_t) qlen, affine);
				my_pool += need;
				if (need > my_max) my_max = need;
				if (!dp_job_is_small((uint32_t) rlen, (uint32_t) qlen)) ++my_big;
			}
		}
		n_runs += __popc(m_run);
		// what runs into the next 32 ops
		bool const new_open = (m_ind >> 31) != 0;
		if (new_open) {
			unsigned const z = ~m_ind;
			int const a2 = z ? 32 - __clz((int) z) : 0;
			bool const cont = !z && open;
			unsigned const range2 = ~((1u << a2) - 1u);
			uint32_t const t2 = ((m_i & range2) ? 1u : 0u) | ((m_d & range2) ? 2u : 0u) | (cont ? open_type : 0u);
			uint64_t const r2 = shfl_u64(ref_before, a2), q2 = shfl_u64(qry_before, a2);
			if (!cont) { open_ref = r2; open_qry = q2; open_k = base + (uint32_t) a2; }
			open_type = t2;
		}
		open = new_open;
		ref_c += shfl_u64(ir, 31); qry_c += shfl_u64(iq, 31);
	}
	flags |= __reduce_or_sync(FULL, my_flags);
	if (!WRITE && n_runs) {
		#pragma unroll
		for (int d = 16; d; d >>= 1) {
			my_cells += shfl_xor_u64(my_cells, d); my_path += shfl_xor_u64(my_path, d); my_pool += shfl_xor_u64(my_pool, d);
			unsigned long long const o = shfl_xor_u64(my_max, d);
			if (o > my_max) my_max = o;
		}
		cells += my_cells; path += my_path; pool += my_pool;
		if (my_max > max_pool) max_pool = my_max;
		if (n_big) *n_big += __reduce_add_sync(FULL, my_big);
	}
	return n_runs;
}

} // namespace pv3
