// panvc3_b200 — the two per-record computations of `panvc3 recalculate-mapq` (sm_100a), SURVEY §8f N4.
//
//   rescore_kernel   cigar_alignment_scorer::calculate_score   recalculate-mapq/main.cc:190-283  (with fill_penalties, :181-186)
//   mapq_kernel      bowtie2_v2_score_calculator::calculate_mapq  recalculate-mapq/main.cc:455-565  (tables :378-430)
//
// Scores are doubles in the reference (alignment_score_type, :45). The alignment score is a sum of int32 penalties, exact in
// either type; the MAPQ formula is evaluated in f64 with the reference's operation order and WITHOUT contraction (the
// reference's `-0.6 + (-0.6 * read_length)` is a multiply and an add on baseline x86-64): __dmul_rn / __dadd_rn / __ddiv_rn.
// Where the reference aborts or has undefined behaviour the record is flagged and its output is unspecified.
#include "pv3_device.cuh"
#include "pv3_kernels.h"

namespace pv3 {

__global__ void __launch_bounds__(128)
rescore_kernel(RescoreParams P)
{
	uint64_t const i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= P.n) return;
	uint64_t const q0 = P.seq_off[i], nq = P.seq_off[i + 1] - q0;
	uint64_t const l0 = P.qual_off[i], nl = P.qual_off[i + 1] - l0;
	uint64_t const n0 = P.refn_off ? P.refn_off[i] : 0, nn = P.refn_off ? P.refn_off[i + 1] - n0 : 0;
	uint64_t query_pos = 0, ref_pos = 0;
	bool prev_gap = false;                                   // prev_op is 'I' or 'D' (:222, :231)
	long long penalty = 0;
	uint32_t flags = 0;
	for (uint64_t c = P.cig_off[i]; c < P.cig_off[i + 1] && !flags; ++c) {
		uint32_t const op = __ldg(P.cig + c), count = op >> 4;
		switch (op & 15u) {
			case OP_H: case OP_P: break;                                                      // :212-215
			case OP_N: ref_pos += count; break;                                               // :217-220
			case OP_I: case OP_D:                                                             // :222-238
				if (!prev_gap) penalty += P.gap_open;
				penalty += (long long) (uint32_t) (count * (uint32_t) P.gap_ext);                // uint32 * int32 is unsigned arithmetic in the reference
				if ((op & 15u) == OP_I) query_pos += count; else ref_pos += count;
				prev_gap = true;
				break;
			case OP_S: query_pos += count; prev_gap = false; break;                           // :240-243
			case OP_EQ:                                                                       // :245-254
				for (uint32_t k = 0; k < count; ++k) {
					uint64_t const qp = query_pos + k;
					if (qp >= nq) { flags = PV3_FLAG_QUERY_OOB; break; }
					if (__ldg(P.seq + q0 + qp) == 3) {
						if (qp >= nl) { flags = PV3_FLAG_BAD_QUAL; break; }
						penalty += P.n_penalties[__ldg(P.qual + l0 + qp)];
					}
				}
				query_pos += count; ref_pos += count; prev_gap = false;
				break;
			case OP_X:                                                                        // :256-272
				for (uint32_t k = 0; k < count; ++k) {
					uint64_t const qp = query_pos + k, rp = ref_pos + k;
					if (qp >= nq) { flags = PV3_FLAG_QUERY_OOB; break; }
					if (qp >= nl) { flags = PV3_FLAG_BAD_QUAL; break; }
					if (rp / 8 >= nn) { flags = PV3_FLAG_REF_OOB; break; }                       // libbio_always_assert_lt (:264)
					uint32_t const ql = __ldg(P.qual + l0 + qp);
					bool const is_n = __ldg(P.seq + q0 + qp) == 3 || ((1u << (rp % 8)) & __ldg(P.refn + n0 + rp / 8));
					penalty += is_n ? P.n_penalties[ql] : P.mismatch_penalties[ql];
				}
				query_pos += count; ref_pos += count; prev_gap = false;
				break;
			default: flags = PV3_FLAG_BAD_OP; break;                                          // 'M' and the rest: libbio_fail (:274-277)
		}
	}
	P.score[i] = -(double) penalty;
	P.flags[i] = flags;
}

namespace {

struct UniqueEntry { double threshold; uint8_t mapq; };
struct NonUniqueEntry { double diff_threshold, score_threshold; uint8_t mapq; };

// main.cc:378-386
__constant__ UniqueEntry const unique_scores[7] = { {0.0, 0}, {0.3, 3}, {0.4, 8}, {0.5, 23}, {0.6, 24}, {0.7, 40}, {0.8, 42} };
// main.cc:397-428
__constant__ NonUniqueEntry const non_unique_scores[30] = {
	{0.0, 0.0, 2}, {0.0, 0.67, 6},
	{0.1, 0.0, 0}, {0.1, 0.67, 7}, {0.1, 0.88, 12}, {0.1, 1.0, 30},
	{0.2, 0.0, 0}, {0.2, 0.67, 11}, {0.2, 0.88, 17}, {0.2, 1.0, 31},
	{0.3, 0.0, 3}, {0.3, 0.67, 15}, {0.3, 0.88, 18}, {0.3, 1.0, 32},
	{0.4, 0.0, 4}, {0.4, 0.68, 14}, {0.4, 0.84, 21}, {0.4, 1.0, 34},
	{0.5, 0.0, 5}, {0.5, 0.68, 16}, {0.5, 0.84, 25}, {0.5, 1.0, 35},
	{0.6, 0.0, 22}, {0.6, 1.0, 36},
	{0.7, 0.0, 26}, {0.7, 1.0, 37},
	{0.8, 0.0, 27}, {0.8, 1.0, 38},
	{0.9, 0.0, 33}, {0.9, 1.0, 39}
};

__device__ __forceinline__ double read_min_score(uint32_t len)     // :455-460
{
	if (!len) return 0.0;
	return __dadd_rn(-0.6, __dmul_rn(-0.6, (double) len));
}

} // namespace

__global__ void __launch_bounds__(256)
mapq_kernel(uint64_t n, const uint32_t *read_len, const uint32_t *other_len, const double *score_in, const double *next_in,
	uint8_t *mapq, uint32_t *flags)
{
	uint64_t const i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	double const DMIN = -1.7976931348623157e308;                                           // ALIGNMENT_SCORE_MIN (:50)
	double const score = score_in[i], next_ = next_in[i];
	double const min_score = __dadd_rn(read_min_score(read_len[i]), read_min_score(other_len[i]));   // :487
	double const max_score = 0.0;                                                            // :474-477, :488
	double const range0 = __dsub_rn(max_score, min_score);
	double const score_range = 1.0 < range0 ? range0 : 1.0;                                  // std::max(1, max - min) (:489)
	uint32_t f = 0;
	uint8_t q = 0;
	if (score < min_score) q = 0;                                                            // :492-493
	else {
		double const next_score = (min_score < next_) ? next_ : min_score;                    // :496
		double const nsq = __ddiv_rn(__dsub_rn(score, min_score), score_range);              // :499-500
		double const diff_next = __dsub_rn(score, next_score);                               // :503
		double const dnq = __ddiv_rn(diff_next, score_range);                                // :504
		if (DMIN == next_score) {                                                            // :510-521
			int k = 0;
			while (k < 7 && !(nsq < unique_scores[k].threshold)) ++k;                          // upper_bound
			if (k == 0) f = PV3_FLAG_TRACE_UB; else q = unique_scores[k - 1].mapq;
		} else if (diff_next != 0.0) {                                                       // :522-548
			int it = 0;
			while (it < 30 && !(dnq < non_unique_scores[it].diff_threshold)) ++it;             // upper_bound by the first component
			if (it == 0) f = PV3_FLAG_TRACE_UB;                                                 // (it - 1) before begin: reference UB
			else {
				double const thr = non_unique_scores[it - 1].diff_threshold;
				int k = 0;                                                                      // upper_bound of (thr, nsq) in [0, it), lexicographic
				while (k < it) {
					NonUniqueEntry const e = non_unique_scores[k];
					bool const less = thr < e.diff_threshold || (!(e.diff_threshold < thr) && nsq < e.score_threshold);
					if (less) break;
					++k;
				}
				if (k == 0) f = PV3_FLAG_TRACE_UB; else q = non_unique_scores[k - 1].mapq;
			}
		} else q = (nsq >= 0.67) ? 1 : 0;                                                    // :560-563
	}
	mapq[i] = q;
	flags[i] = f;
}

void launch_rescore(const RescoreParams &P, cudaStream_t stream)
{
	if (P.n) rescore_kernel<<<(unsigned) ((P.n + 127) / 128), 128, 0, stream>>>(P);
}

void launch_mapq(uint64_t n, const uint32_t *read_len, const uint32_t *other_len, const double *score, const double *next_score,
	uint8_t *mapq, uint32_t *flags, cudaStream_t stream)
{
	if (n) mapq_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, stream>>>(n, read_len, other_len, score, next_score, mapq, flags);
}

} // namespace pv3
