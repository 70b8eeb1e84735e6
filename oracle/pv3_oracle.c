/*
 * TEST INFRASTRUCTURE ONLY — plain-C restatement ("port") of the reference's project_alignments hot path.
 *
 * Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the CHECKER for the
 * CUDA path. The product (panvc3_b200/) never includes, links or loads it. Parity status: PINNED — every function
 * here is diffed by tests/test_oracle_vs_ref.py against the reference's own sources compiled in place
 * (oracle/_ref/libpv3_ref.so, see oracle/Makefile) and against the reference's fixtures (tests/golden/).
 *
 * Reference lines restated (paths under /root/reference):
 *   rank0/select0/access       include/panvc3/msa_index.hh:239-258 over sdsl rrr_vector<15>
 *                              (lib/sdsl-lite/include/sdsl/rrr_vector_15.hpp:460-559,664-701) — results are a pure
 *                              function of the bit sequence, so a plain bit vector + counts restates them
 *   rewrite_cigar              include/panvc3/rewrite_cigar.hh:19-157
 *   cigar_buffer / collapse    libpanvc3/cigar.cc:13-95
 *   indel_run_checker          libpanvc3/indel_run_checker.cc:13-121
 *   alignment_projector        libpanvc3/alignment_projector.cc:30-154
 *   align_global               include/panvc3/align.hh:23-113 and the three seqan3 algorithms it dispatches to
 *                              
 */
#define _GNU_SOURCE
#include "pv3_oracle_api.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

enum { OP_M = 0, OP_I = 1, OP_D = 2, OP_N = 3, OP_S = 4, OP_H = 5, OP_P = 6, OP_EQ = 7, OP_X = 8 };

/* ------------------------------------------------------------------ bit rows */

typedef struct {
	uint64_t *words;
	uint64_t  n_bits;
	uint64_t  n_words;
	uint64_t *zeros_before; /* per 64-bit word, zeros in [0, 64*w) counting only bits < n_bits */
	uint64_t  n_zeros;
} row_t;

typedef struct {
	char    *bases;
	uint64_t n;
} ref_t;

struct pv3o_index {
	row_t *rows; uint64_t n_rows, cap_rows;
	ref_t *refs; uint64_t n_refs, cap_refs;
};

const char *pv3o_kind(void) { return "port"; }

pv3o_index *pv3o_index_create(void) { return (pv3o_index *) calloc(1, sizeof(pv3o_index)); }

void pv3o_index_destroy(pv3o_index *idx)
{
	if (!idx) return;
	for (uint64_t i = 0; i < idx->n_rows; ++i) { free(idx->rows[i].words); free(idx->rows[i].zeros_before); }
	for (uint64_t i = 0; i < idx->n_refs; ++i) free(idx->refs[i].bases);
	free(idx->rows); free(idx->refs); free(idx);
}

int64_t pv3o_index_add_row(pv3o_index *idx, const uint64_t *gap_words, uint64_t n_bits)
{
	if (idx->n_rows == idx->cap_rows) {
		idx->cap_rows = idx->cap_rows ? 2 * idx->cap_rows : 8;
		idx->rows = (row_t *) realloc(idx->rows, idx->cap_rows * sizeof(row_t));
	}
	row_t *r = &idx->rows[idx->n_rows];
	r->n_bits = n_bits;
	r->n_words = (n_bits + 63) / 64;
	r->words = (uint64_t *) malloc((r->n_words + 1) * 8);
	r->zeros_before = (uint64_t *) malloc((r->n_words + 1) * 8);
	memcpy(r->words, gap_words, r->n_words * 8);
	/* Bits at and after n_bits count as gaps (ones) so that they are never selected. */
	if (n_bits % 64) r->words[r->n_words - 1] |= ~UINT64_C(0) << (n_bits % 64);
	r->words[r->n_words] = ~UINT64_C(0);
	uint64_t z = 0;
	for (uint64_t w = 0; w < r->n_words; ++w) { r->zeros_before[w] = z; z += (uint64_t) __builtin_popcountll(~r->words[w]); }
	r->zeros_before[r->n_words] = z;
	r->n_zeros = z;
	return (int64_t) idx->n_rows++;
}

int64_t pv3o_index_add_ref(pv3o_index *idx, const char *bases, uint64_t n)
{
	if (idx->n_refs == idx->cap_refs) {
		idx->cap_refs = idx->cap_refs ? 2 * idx->cap_refs : 8;
		idx->refs = (ref_t *) realloc(idx->refs, idx->cap_refs * sizeof(ref_t));
	}
	ref_t *r = &idx->refs[idx->n_refs];
	r->bases = (char *) malloc(n ? n : 1);
	memcpy(r->bases, bases, n);
	r->n = n;
	return (int64_t) idx->n_refs++;
}

/* rank0(pos) = zeros in [0,pos)   (msa_index.hh:247-251; rank_support_rrr<0>, rrr_vector_15.hpp:527-559) */
static uint64_t row_rank0(const row_t *r, uint64_t pos)
{
	if (pos >= r->n_bits) return r->n_zeros;
	uint64_t w = pos / 64, b = pos % 64;
	uint64_t z = r->zeros_before[w];
	if (b) z += (uint64_t) __builtin_popcountll(~r->words[w] & ((UINT64_C(1) << b) - 1));
	return z;
}

/* select0(k) = position of the k-th zero (k>=1); size() when there is none (rrr_vector_15.hpp:664-667) */
static uint64_t row_select0(const row_t *r, uint64_t k)
{
	if (k == 0 || k > r->n_zeros) return r->n_bits;
	uint64_t lo = 0, hi = r->n_words; /* last w with zeros_before[w] < k */
	while (hi - lo > 1) {
		uint64_t mid = lo + (hi - lo) / 2;
		if (r->zeros_before[mid] < k) lo = mid; else hi = mid;
	}
	uint64_t need = k - r->zeros_before[lo];
	uint64_t inv = ~r->words[lo];
	for (uint64_t i = 1; i < need; ++i) inv &= inv - 1;
	return lo * 64 + (uint64_t) __builtin_ctzll(inv);
}

static int row_access(const row_t *r, uint64_t pos) { return (int) ((r->words[pos / 64] >> (pos % 64)) & 1); }

uint64_t pv3o_rank0(const pv3o_index *idx, uint32_t row, uint64_t pos) { return row_rank0(&idx->rows[row], pos); }
uint64_t pv3o_select0(const pv3o_index *idx, uint32_t row, uint64_t k) { return row_select0(&idx->rows[row], k); }
int pv3o_access(const pv3o_index *idx, uint32_t row, uint64_t pos) { return row_access(&idx->rows[row], pos); }

/* msa_index.hh:253-258 */
uint64_t pv3o_project_position(const pv3o_index *idx, uint32_t src_row, uint32_t dst_row, uint64_t pos)
{
	return row_rank0(&idx->rows[dst_row], row_select0(&idx->rows[src_row], pos + 1));
}

/* ------------------------------------------------------------------ cigar_buffer (cigar.cc:54-95) */

typedef struct {
	uint32_t *ops; uint64_t n, cap;
	uint32_t cur_op; uint64_t cur_count; int is_first;
} cigbuf_t;

static void cb_append(cigbuf_t *b, uint32_t op, uint64_t count)
{
	if (b->n == b->cap) { b->cap = b->cap ? 2 * b->cap : 64; b->ops = (uint32_t *) realloc(b->ops, b->cap * 4); }
	b->ops[b->n++] = (uint32_t) (count << 4) | op; /* count is uint32 in the reference (cigar.hh:21) and wraps alike */
}
static void cb_clear(cigbuf_t *b) { b->n = 0; b->is_first = 1; b->cur_count = 0; }
static void cb_push(cigbuf_t *b, uint32_t op, uint64_t count)
{
	count &= 0xffffffffu;
	if (!count) return;
	if (b->is_first) { b->is_first = 0; b->cur_op = op; b->cur_count = count; }
	else if (b->cur_op == op) b->cur_count = (b->cur_count + count) & 0xffffffffu;
	else { if (b->cur_count) cb_append(b, b->cur_op, b->cur_count); b->cur_op = op; b->cur_count = count; }
}
static void cb_finish(cigbuf_t *b) { if (!b->is_first && b->cur_count) cb_append(b, b->cur_op, b->cur_count); }

/* ------------------------------------------------------------------ rewrite_cigar (rewrite_cigar.hh:19-157) */

static const char dna5_chars[5] = { 'A', 'C', 'G', 'N', 'T' }; /* dna5::to_char, dna5.hpp:95-97 */

/* dna5::assign_char: case-insensitive, U->T, everything else -> N (dna5.hpp:99-110) */
static uint8_t dna5_rank_of_char(char c)
{
	switch (c) {
		case 'A': case 'a': return 0;
		case 'C': case 'c': return 1;
		case 'G': case 'g': return 2;
		case 'T': case 't': case 'U': case 'u': return 4;
		default: return 3;
	}
}

static uint64_t rewrite_cigar(uint64_t src_pos, const uint32_t *cig, uint64_t n_cig,
	const row_t *src, const row_t *dst, const uint8_t *query, uint64_t n_query,
	const ref_t *ref, cigbuf_t *out, uint32_t *flags)
{
	uint64_t query_pos = 0;
	uint64_t aln_pos = row_select0(src, src_pos + 1);            /* :51; == size() past the last non-gap, not an error yet */
	uint64_t prev_excess = row_rank0(dst, aln_pos);              /* :52 */
	uint64_t const retval = prev_excess;                         /* :56 */

	cb_clear(out);
	for (uint64_t k = 0; k < n_cig; ++k) {
		uint64_t const count = cig[k] >> 4;
		uint32_t const op = cig[k] & 0xf;
		switch (op) {
			case OP_I: case OP_S:                                /* :68-72 */
				cb_push(out, op, count); query_pos += count; break;
			case OP_H: case OP_P:                                /* :74-77 */
				cb_push(out, op, count); break;
			case OP_M: case OP_EQ: case OP_X:                    /* :79-117 */
				for (uint64_t i = 0; i < count; ++i) {
					uint64_t const c = row_select0(src, 1 + src_pos);
					if (c >= src->n_bits) { *flags |= PV3O_FLAG_POS_OOB; return retval; }
					uint64_t const excess = row_rank0(dst, c);
					if (prev_excess < excess) cb_push(out, OP_D, excess - prev_excess);
					if (row_access(dst, c)) { cb_push(out, OP_I, 1); prev_excess = excess; }
					else {
						if (query_pos >= n_query) { *flags |= PV3O_FLAG_QUERY_OOB; return retval; }
						if (excess >= ref->n) { *flags |= PV3O_FLAG_REF_OOB; return retval; }
						char const qc = query[query_pos] < 5 ? dna5_chars[query[query_pos]] : 'N';
						cb_push(out, qc == ref->bases[excess] ? OP_EQ : OP_X, 1); /* raw byte compare, SURVEY F7 */
						prev_excess = excess + 1;
					}
					++query_pos; ++src_pos;
				}
				break;
			case OP_D: case OP_N:                                /* :119-147 */
				for (uint64_t i = 0; i < count; ++i) {
					uint64_t const c = row_select0(src, 1 + src_pos);
					if (c >= src->n_bits) { *flags |= PV3O_FLAG_POS_OOB; return retval; }
					uint64_t const excess = row_rank0(dst, c);
					if (prev_excess < excess) cb_push(out, OP_D, excess - prev_excess);
					if (!row_access(dst, c)) { cb_push(out, OP_D, 1); prev_excess = excess + 1; }
					else prev_excess = excess;
					++src_pos;
				}
				break;
			default:                                             /* :149-151 libbio_fail */
				*flags |= PV3O_FLAG_BAD_OP; return retval;
		}
	}
	cb_finish(out);                                              /* :155 */
	return retval;
}

/* ------------------------------------------------------------------ indel_run_checker (indel_run_checker.cc:31-121) */

typedef struct {
	uint64_t op_begin, op_end;
	uint64_t ref_loc, ref_len, qry_loc, qry_len; /* the DP window (m_ref_range / m_query_range) */
	uint64_t ref_pos_after;                      /* live m_ref_pos when the run is reported */
} run_t;

typedef struct { run_t *runs; uint64_t n, cap; } runvec_t;

static void rv_push(runvec_t *v, run_t r)
{
	if (v->n == v->cap) { v->cap = v->cap ? 2 * v->cap : 8; v->runs = (run_t *) realloc(v->runs, v->cap * sizeof(run_t)); }
	v->runs[v->n++] = r;
}

/* Returns 0, or -1 on an op the reference would libbio_fail on. final_pos: {query_pos, ref_pos}. */
static int find_runs(const uint32_t *ops, uint64_t n_ops, uint64_t ref_pos0, runvec_t *out, uint64_t *final_pos)
{
	uint64_t ref = ref_pos0, qry = 0, run_ref = ref_pos0, run_qry = 0, run_k = 0;
	unsigned type = 0;
	out->n = 0;
	for (uint64_t k = 0; k < n_ops; ++k) {
		uint64_t const ref0 = ref, qry0 = qry;
		unsigned const type0 = type;
		uint64_t const n = ops[k] >> 4;
		if (!type) { run_ref = ref; run_qry = qry; run_k = k; }      /* :44-49 */
		switch (ops[k] & 0xf) {
			case OP_I: qry += n; type |= 1; continue;                /* :60-63 */
			case OP_D: ref += n; type |= 2; continue;                /* :65-68 */
			case OP_H: case OP_P: type = 0; break;                   /* :54-58 */
			case OP_S: qry += n; type = 0; break;                    /* :70-75 */
			case OP_N: ref += n; type = 0; break;                    /* :77-82 */
			case OP_M: case OP_EQ: case OP_X: ref += n; qry += n; type = 0; break; /* :84-92 */
			default: return -1;                                      /* :94-98 */
		}
		if (type0 == 3) {                                            /* report_run :104-108 */
			run_t r = { run_k, k, run_ref, ref0 - run_ref, run_qry, qry0 - run_qry, ref };
			rv_push(out, r);
		}
	}
	if (type == 3) {                                                 /* :111-118 */
		run_t r = { run_k, n_ops, run_ref, ref - run_ref, run_qry, qry - run_qry, ref };
		rv_push(out, r);
	}
	if (final_pos) { final_pos[0] = qry; final_pos[1] = ref; }
	return 0;
}

int64_t pv3o_indel_runs(const uint32_t *ops, uint64_t n_ops, uint64_t ref_pos, uint64_t *out, uint64_t cap, uint64_t *final_pos)
{
	runvec_t rv = { 0, 0, 0 };
	if (find_runs(ops, n_ops, ref_pos, &rv, final_pos) < 0) { free(rv.runs); return -3; }
	if (rv.n > cap) { free(rv.runs); return -1; }
	for (uint64_t i = 0; i < rv.n; ++i) {
		uint64_t *d = out + 6 * i;
		d[0] = rv.runs[i].op_begin; d[1] = rv.runs[i].op_end;
		d[2] = rv.runs[i].ref_loc; d[3] = rv.runs[i].ref_len;
		d[4] = rv.runs[i].qry_loc; d[5] = rv.runs[i].qry_len;
	}
	int64_t n = (int64_t) rv.n;
	free(rv.runs);
	return n;
}

/* ------------------------------------------------------------------ align_global (align.hh:66-113) */

/* Quality LUT: base_quality_scorer::score (align.hh:28-37) truncated toward zero to int32 because seqan3's score_t is
 * int32_t when no align_cfg::score_type is given. Reference base always has the max phred42 letter (41)
 * (alignment_projector.cc:108). Expression order kept as in the reference. */
static void quality_scores(int qq, int32_t *match, int32_t *mismatch)
{
	double const lhs_e = 1.0 / pow(10.0, 41 / 10.0);
	double const rhs_e = 1.0 / pow(10.0, qq / 10.0);
	double const combined_e = lhs_e + rhs_e - lhs_e / 3.0 * rhs_e * 4.0;
	*match = (int32_t) (2.0 + log2(1.0 - combined_e));
	*mismatch = (int32_t) (2.0 - log2(3.0) + log2(combined_e));
}

void pv3o_quality_lut(int32_t *lut84)
{
	for (int q = 0; q < 42; ++q) quality_scores(q, &lut84[2 * q], &lut84[2 * q + 1]);
}

enum { STEP_DIAG = 0, STEP_UP = 1, STEP_LEFT = 2 };

/* Emits the reversed path through cigar_buffer as align.hh:94-110 does: Left ⇒ D, Up ⇒ I, Diag ⇒ =/X. */
static void emit_path(const uint8_t *steps, uint64_t n_steps, const uint8_t *r, const uint8_t *q, cigbuf_t *out)
{
	uint64_t i = 0, j = 0;
	cb_clear(out);
	for (uint64_t s = n_steps; s-- > 0;) {
		switch (steps[s]) {
			case STEP_LEFT: cb_push(out, OP_D, 1); ++j; break;
			case STEP_UP:   cb_push(out, OP_I, 1); ++i; break;
			default:        cb_push(out, r[j] == q[i] ? OP_EQ : OP_X, 1); ++i; ++j; break;
		}
	}
	cb_finish(out);
}

/* Mode A6: qual empty && gap_open == 0 ⇒ seqan3 edit distance (alignment_configurator.hpp:323-349), unit costs,
 * gap_extension ignored, traceback priority Left > Up > Diagonal (edit_distance_trace_matrix_full.hpp:273-286).
 * SURVEY Appendix C.1. */
static int32_t align_edit(const uint8_t *r, uint64_t n, const uint8_t *q, uint64_t m, cigbuf_t *out)
{
	uint64_t const w = n + 1;
	int32_t *D = (int32_t *) malloc((m + 1) * w * sizeof(int32_t));
	for (uint64_t j = 0; j <= n; ++j) D[j] = (int32_t) j;
	for (uint64_t i = 1; i <= m; ++i) {
		D[i * w] = (int32_t) i;
		for (uint64_t j = 1; j <= n; ++j) {
			int32_t best = D[(i - 1) * w + j - 1] + (r[j - 1] != q[i - 1]);
			int32_t const up = D[(i - 1) * w + j] + 1, left = D[i * w + j - 1] + 1;
			if (up < best) best = up;
			if (left < best) best = left;
			D[i * w + j] = best;
		}
	}
	uint8_t *steps = (uint8_t *) malloc(n + m + 1);
	uint64_t ns = 0, i = m, j = n;
	while (i > 0 || j > 0) {
		if (j > 0 && (i == 0 || D[i * w + j] == D[i * w + j - 1] + 1)) { steps[ns++] = STEP_LEFT; --j; }
		else if (i > 0 && (j == 0 || D[i * w + j] == D[(i - 1) * w + j] + 1)) { steps[ns++] = STEP_UP; --i; }
		else { steps[ns++] = STEP_DIAG; --i; --j; }
	}
	int32_t const score = -D[m * w + n];
	emit_path(steps, ns, r, q, out);
	free(steps); free(D);
	return score;
}

/* Modes A7/A8: seqan3 alignment_algorithm + affine_gap_policy (affine_gap_policy.hpp:87-131), column-major Gotoh with a
 * 5-bit trace and the stateful traceback of trace_iterator_base.hpp:150-179,254-274. SURVEY Appendix C.2.
 * lut == NULL ⇒ nucleotide_scoring_scheme 0/-1 (A7); else lut[2*qual+{0:match,1:mismatch}] (A8).
 * Returns 0, or -2 when the traceback would leave the matrix (reference UB, only with gap_open < 0). */
enum { T_D = 1, T_UO = 2, T_U = 4, T_LO = 8, T_L = 16 };

static int trace_pick(uint8_t t) { return (t & T_D) ? STEP_DIAG : (t & (T_U | T_UO)) ? STEP_UP : (t & (T_L | T_LO)) ? STEP_LEFT : -1; }

static int align_affine(const uint8_t *r, uint64_t n, const uint8_t *q, uint64_t m, const uint8_t *qual, const int32_t *lut,
	int32_t gap_open, int32_t gap_ext, cigbuf_t *out, int32_t *score)
{
	int32_t const gext = gap_ext, gopen = gap_open + gap_ext;    /* affine_gap_policy.hpp:183-190 */
	uint64_t const h = m + 1;
	uint8_t *T = (uint8_t *) malloc((n + 1) * h);                /* column-major: T[j*h + i] */
	int32_t *cur = (int32_t *) malloc(h * sizeof(int32_t));
	int32_t *left = (int32_t *) malloc(h * sizeof(int32_t));
	uint8_t *tleft = (uint8_t *) malloc(h);

	/* column 0 (affine_gap_init_policy.hpp:84-216) */
	int32_t up = gopen; uint8_t tup = T_UO;
	cur[0] = 0; T[0] = 0; left[0] = gopen; tleft[0] = T_LO;
	for (uint64_t i = 1; i <= m; ++i) {
		cur[i] = up; T[i] = tup; up += gext; tup = T_U; left[i] = cur[i] + gopen; tleft[i] = T_LO;
	}
	for (uint64_t j = 1; j <= n; ++j) {
		int32_t diag = cur[0];
		cur[0] = left[0]; T[j * h] = tleft[0]; up = cur[0] + gopen; tup = T_UO; left[0] += gext; tleft[0] = T_L;
		for (uint64_t i = 1; i <= m; ++i) {
			int32_t s;
			if (lut) s = lut[2 * qual[i - 1] + (r[j - 1] == q[i - 1] ? 0 : 1)];
			else s = (r[j - 1] == q[i - 1]) ? 0 : -1;
			int32_t tmp = diag + s; diag = cur[i];
			uint8_t t;
			if (tmp < up) { t = tup; tmp = up; } else t = (uint8_t) (T_D | tup);
			if (tmp < left[i]) { t = tleft[i]; tmp = left[i]; } else t |= tleft[i];
			cur[i] = tmp; T[j * h + i] = t;
			tmp += gopen; up += gext;
			int32_t const wl = left[i] + gext;
			if (up < tmp) { tup = T_UO; up = tmp; } else tup = T_U;
			if (wl < tmp) { tleft[i] = T_LO; left[i] = tmp; } else { tleft[i] = T_L; left[i] = wl; }
		}
	}
	if (score) *score = cur[m];

	uint8_t *steps = (uint8_t *) malloc(n + m + 2);
	uint64_t ns = 0, i = m, j = n;
	int status = 0;
	int dir = trace_pick(T[j * h + i]);
	while (T[j * h + i] != 0) {
		uint8_t const old = T[j * h + i];
		if ((dir == STEP_UP && i == 0) || (dir == STEP_LEFT && j == 0) || (dir == STEP_DIAG && (i == 0 || j == 0)) || dir < 0
			|| ns >= n + m + 1) { status = -2; break; }
		steps[ns++] = (uint8_t) dir;
		if (dir == STEP_UP) { --i; if (old & T_UO) dir = trace_pick(T[j * h + i]); }
		else if (dir == STEP_LEFT) { --j; if (old & T_LO) dir = trace_pick(T[j * h + i]); }
		else { --i; --j; dir = trace_pick(T[j * h + i]); }
	}
	if (!status && (i != 0 || j != 0)) status = -2;
	if (!status) emit_path(steps, ns, r, q, out);
	free(steps); free(tleft); free(left); free(cur); free(T);
	return status;
}

/* Mode routing: qual empty && go==0 → A6; qual empty && go!=0 → A7; qual present → A8. */
static int align_global(const uint8_t *r, uint64_t n, const uint8_t *q, uint64_t m, const uint8_t *qual,
	int32_t gap_open, int32_t gap_ext, const int32_t *lut, cigbuf_t *out, int32_t *score)
{
	if (!qual && gap_open == 0) { int32_t s = align_edit(r, n, q, m, out); if (score) *score = s; return 0; }
	return align_affine(r, n, q, m, qual, qual ? lut : NULL, gap_open, gap_ext, out, score);
}

int64_t pv3o_align_global(const uint8_t *ref_ranks, uint64_t n, const uint8_t *qry_ranks, uint64_t m,
	const uint8_t *qual, int32_t gap_open, int32_t gap_ext, uint32_t *out_ops, uint64_t cap, int32_t *score)
{
	int32_t lut[84];
	pv3o_quality_lut(lut);
	cigbuf_t buf; memset(&buf, 0, sizeof buf);
	int const st = align_global(ref_ranks, n, qry_ranks, m, qual, gap_open, gap_ext, lut, &buf, score);
	int64_t ret;
	if (st < 0) ret = st;
	else if (buf.n > cap) ret = -1;
	else { memcpy(out_ops, buf.ops, buf.n * 4); ret = (int64_t) buf.n; }
	free(buf.ops);
	return ret;
}

/* ------------------------------------------------------------------ alignment_projector::project_alignment
 * (alignment_projector.cc:30-154) and the batch loop */

typedef struct {
	uint64_t  dst_pos;
	uint32_t  flags;
	uint32_t *ops; uint64_t n_ops;
	uint64_t *rr;  uint64_t n_rr;
} result_t;

typedef struct {
	const pv3o_index *idx; uint64_t begin, end;
	const uint32_t *src_row, *dst_row, *dst_ref; const uint64_t *src_pos;
	const uint64_t *seq_off; const uint8_t *seq;
	const uint64_t *qual_off; const uint8_t *qual;
	const uint64_t *cig_off; const uint32_t *cig;
	int32_t go, ge; int stage;
	const int32_t *lut;
	result_t *results;
} shard_t;

static void vec_append(uint32_t **v, uint64_t *n, uint64_t *cap, const uint32_t *src, uint64_t k)
{
	if (*n + k > *cap) { while (*n + k > *cap) *cap = *cap ? 2 * *cap : 64; *v = (uint32_t *) realloc(*v, *cap * 4); }
	memcpy(*v + *n, src, k * 4); *n += k;
}

static void *shard_main(void *arg)
{
	shard_t *sh = (shard_t *) arg;
	cigbuf_t rewritten, realigned; memset(&rewritten, 0, sizeof rewritten); memset(&realigned, 0, sizeof realigned);
	runvec_t runs = { 0, 0, 0 };
	uint32_t *fin = NULL; uint64_t n_fin = 0, cap_fin = 0;
	uint8_t *rbuf = NULL; uint64_t cap_rbuf = 0;

	for (uint64_t rec = sh->begin; rec < sh->end; ++rec) {
		result_t *res = &sh->results[rec];
		memset(res, 0, sizeof *res);
		const row_t *src = &sh->idx->rows[sh->src_row[rec]], *dst = &sh->idx->rows[sh->dst_row[rec]];
		const ref_t *ref = &sh->idx->refs[sh->dst_ref[rec]];
		const uint8_t *query = sh->seq + sh->seq_off[rec];
		uint64_t const n_query = sh->seq_off[rec + 1] - sh->seq_off[rec];
		const uint8_t *qual = NULL;
		if (sh->qual_off && sh->qual_off[rec + 1] > sh->qual_off[rec]) {
			qual = sh->qual + sh->qual_off[rec];
			if (sh->qual_off[rec + 1] - sh->qual_off[rec] != n_query) res->flags |= PV3O_FLAG_BAD_QUAL;
		}
		const uint32_t *cig = sh->cig + sh->cig_off[rec];
		uint64_t const n_cig = sh->cig_off[rec + 1] - sh->cig_off[rec];

		res->dst_pos = rewrite_cigar(sh->src_pos[rec], cig, n_cig, src, dst, query, n_query, ref, &rewritten, &res->flags);
		if (res->flags) continue;

		if (sh->stage == 1) {
			res->ops = (uint32_t *) malloc((rewritten.n ? rewritten.n : 1) * 4);
			memcpy(res->ops, rewritten.ops, rewritten.n * 4); res->n_ops = rewritten.n;
			continue;
		}

		find_runs(rewritten.ops, rewritten.n, res->dst_pos, &runs, NULL);  /* ops are I S H P D = X only: cannot fail */
		n_fin = 0;
		uint64_t op_begin = 0;
		if (runs.n) res->rr = (uint64_t *) malloc(runs.n * 4 * 8);
		for (uint64_t k = 0; k < runs.n && !res->flags; ++k) {
			run_t const *run = &runs.runs[k];
			vec_append(&fin, &n_fin, &cap_fin, rewritten.ops + op_begin, run->op_begin - op_begin);  /* :61-66 */
			op_begin = run->op_end;
			uint64_t *rr = res->rr + 4 * res->n_rr++;
			rr[0] = run->ref_pos_after; rr[1] = run->ref_len;      /* :69-72 — shifted location, SURVEY F10 */
			rr[2] = run->qry_loc; rr[3] = run->qry_len;            /* :73 */
			if (run->ref_loc + run->ref_len > ref->n) { res->flags |= PV3O_FLAG_REF_OOB; break; }
			if (run->qry_loc + run->qry_len > n_query) { res->flags |= PV3O_FLAG_QUERY_OOB; break; }
			if (run->ref_len > cap_rbuf) { cap_rbuf = 2 * run->ref_len; rbuf = (uint8_t *) realloc(rbuf, cap_rbuf); }
			for (uint64_t t = 0; t < run->ref_len; ++t) rbuf[t] = dna5_rank_of_char(ref->bases[run->ref_loc + t]); /* :78-82 */
			int const st = align_global(rbuf, run->ref_len, query + run->qry_loc, run->qry_len,
				qual ? qual + run->qry_loc : NULL, sh->go, sh->ge, sh->lut, &realigned, NULL);
			if (st < 0) { res->flags |= PV3O_FLAG_TRACE_UB; break; }
			vec_append(&fin, &n_fin, &cap_fin, realigned.ops, realigned.n);                            /* :140-141 */
		}
		if (res->flags) continue;
		vec_append(&fin, &n_fin, &cap_fin, rewritten.ops + op_begin, rewritten.n - op_begin);         /* :145-148 */

		/* collapse_cigar_operations (cigar.cc:13-51); an empty vector is UB there */
		if (!n_fin) { res->flags |= PV3O_FLAG_EMPTY_CIGAR; continue; }
		uint64_t w = 0;
		for (uint64_t k = 1; k < n_fin; ++k) {
			if ((fin[w] & 0xf) == (fin[k] & 0xf)) fin[w] = (((fin[w] >> 4) + (fin[k] >> 4)) << 4) | (fin[w] & 0xf);
			else fin[++w] = fin[k];
		}
		res->n_ops = w + 1;
		res->ops = (uint32_t *) malloc(res->n_ops * 4);
		memcpy(res->ops, fin, res->n_ops * 4);
	}
	free(rewritten.ops); free(realigned.ops); free(runs.runs); free(fin); free(rbuf);
	return NULL;
}

int pv3o_project_batch(const pv3o_index *idx, uint64_t n,
	const uint32_t *src_row, const uint32_t *dst_row, const uint32_t *dst_ref, const uint64_t *src_pos,
	const uint64_t *seq_off, const uint8_t *seq,
	const uint64_t *qual_off, const uint8_t *qual,
	const uint64_t *cig_off, const uint32_t *cig,
	int32_t gap_open, int32_t gap_ext, int n_threads, int stage,
	uint64_t *dst_pos,
	uint64_t *out_cig_off, uint32_t *out_cig, uint64_t out_cig_cap,
	uint64_t *rr_off, uint64_t *rr, uint64_t rr_cap,
	uint32_t *flags)
{
	int32_t lut[84];
	pv3o_quality_lut(lut);
	result_t *results = (result_t *) calloc(n ? n : 1, sizeof(result_t));
	if (n_threads < 1) n_threads = 1;
	if ((uint64_t) n_threads > n) n_threads = n ? (int) n : 1;
	shard_t *shards = (shard_t *) calloc((size_t) n_threads, sizeof(shard_t));
	pthread_t *tids = (pthread_t *) calloc((size_t) n_threads, sizeof(pthread_t));
	for (int t = 0; t < n_threads; ++t) {
		shard_t s = { idx, n * (uint64_t) t / (uint64_t) n_threads, n * (uint64_t) (t + 1) / (uint64_t) n_threads,
			src_row, dst_row, dst_ref, src_pos, seq_off, seq, qual_off, qual, cig_off, cig, gap_open, gap_ext, stage, lut, results };
		shards[t] = s;
		if (n_threads == 1) shard_main(&shards[t]);
		else pthread_create(&tids[t], NULL, shard_main, &shards[t]);
	}
	if (n_threads > 1) for (int t = 0; t < n_threads; ++t) pthread_join(tids[t], NULL);

	uint64_t co = 0, ro = 0; int status = 0;
	for (uint64_t i = 0; i < n; ++i) {
		result_t *res = &results[i];
		dst_pos[i] = res->dst_pos;
		if (flags) flags[i] = res->flags;
		out_cig_off[i] = co; rr_off[i] = ro;
		if (res->flags) { res->n_ops = 0; res->n_rr = 0; }
		if (co + res->n_ops > out_cig_cap || ro + res->n_rr > rr_cap) status = -1;
		else {
			if (res->n_ops) memcpy(out_cig + co, res->ops, res->n_ops * 4);
			if (res->n_rr) memcpy(rr + 4 * ro, res->rr, res->n_rr * 32);
			co += res->n_ops; ro += res->n_rr;
		}
		free(res->ops); free(res->rr);
	}
	out_cig_off[n] = co; rr_off[n] = ro;
	free(results); free(shards); free(tids);
	return status;
}

/* ---- index-msa: gap bits + unaligned sequence of one MSA row (index-msa/index_vector_builder.cc:28-50) ---------------- */
void pv3o_msa_row_extract(const char *text, uint64_t n, uint64_t *gap_words, char *ungapped, uint64_t *n_ungapped)
{
	uint64_t non_gap = 0;
	memset(gap_words, 0, ((n + 63) / 64) * 8);
	for (uint64_t i = 0; i < n; ++i) {
		if ('-' == text[i]) gap_words[i >> 6] |= 1ull << (i & 63);     /* m_bv[m_pos + i] = 1 */
		else { if (ungapped) ungapped[non_gap] = text[i]; ++non_gap; } /* std::cout << cc */
	}
	if (n_ungapped) *n_ungapped = non_gap;
}

/* ---- OA entry / CIGAR text of a record (project-alignments/project_alignments.cc:803-845) -----------------------------
 * PINNED: the code restated here sits inside project_task::process of the CLI, which cannot be built in this container,
 * but `make ref` cuts those statements (project_alignments.cc:762-845) out of the file where it lies and compiles them as a
 * function body (oracle/ref_driver_annotate.cc); tests/test_oracle.py compares this restatement with them. One difference
 * of type: POS is the int32 of seqan3's reference_position() in the reference, a uint64 here (equal for every valid POS).
 * Returns the bytes written (or needed when cap is too small: nothing is written past cap). ref_id == NULL ⇒ CIGAR text only. */
#include <stdio.h>
uint64_t pv3o_format_oa(uint64_t n, const uint32_t *ref_id, const uint64_t *pos, const uint16_t *flag, const uint8_t *mapq,
	const int32_t *nm, const uint8_t *has_nm, const uint64_t *cig_off, const uint32_t *cig_len, const uint32_t *cig,
	uint32_t n_names, const uint64_t *name_off, const char *names, uint64_t *text_off, char *text, uint64_t cap)
{
	uint64_t at = 0;
	char buf[64];
	#define PUT(ptr, len) do { uint64_t l_ = (len); if (at + l_ <= cap) memcpy(text + at, (ptr), l_); at += l_; } while (0)
	for (uint64_t i = 0; i < n; ++i) {
		text_off[i] = at;
		if (ref_id) {
			if (ref_id[i] < n_names) PUT(names + name_off[ref_id[i]], name_off[ref_id[i] + 1] - name_off[ref_id[i]]);   /* :812 */
			PUT(",", 1);
			PUT(buf, (uint64_t) snprintf(buf, sizeof buf, "%llu,", (unsigned long long) pos[i]));                          /* :815 */
			PUT((flag[i] & 0x10) ? "-," : "+,", 2);                                                                     /* :819 */
		}
		uint64_t const c0 = cig_off[i], c1 = cig_len ? c0 + cig_len[i] : cig_off[i + 1];
		for (uint64_t c = c0; c < c1; ++c)                                                                              /* :822-826 */
			PUT(buf, (uint64_t) snprintf(buf, sizeof buf, "%u%c", cig[c] >> 4, "MIDNSHP=X???????"[cig[c] & 15]));
		if (ref_id) {
			PUT(buf, (uint64_t) snprintf(buf, sizeof buf, ",%d,", (int) mapq[i]));                                        /* :827-830 */
			if (has_nm && has_nm[i]) PUT(buf, (uint64_t) snprintf(buf, sizeof buf, "%d", (int) nm[i]));                   /* :834-835 */
			PUT(";", 1);                                                                                                /* :838 */
		}
	}
	#undef PUT
	text_off[n] = at;
	return at;
}

/* ---- recalculate-mapq: alignment rescoring and the Bowtie 2 (v2) MAPQ formula -----------------------------------------
 * PINNED: both functions live in the first part of recalculate-mapq/main.cc; the CLI's translation unit is not buildable
 * here, but `make ref` cuts that part out of the file where it lies and compiles it (oracle/ref_driver_mapq.cc), and
 * tests/test_oracle.py::test_rescore_and_mapq_port_match_the_reference_code compares this restatement with it. */
#include <float.h>

static void fill_penalties(int32_t *penalties, int32_t min, int32_t max)               /* main.cc:181-186 */
{
	for (size_t i = 0; i < 256; ++i) penalties[i] = (int32_t) ((i < 40 ? i : 40) / (double) 40 * (max - min) + min);
}

/* cigar_alignment_scorer::calculate_score, main.cc:190-283. flags[i] != 0 where the reference aborts or reads out of range
 * (1 bad op, 4 SEQ too short, 8 X base beyond the XN bitmap, 0x40 QUAL too short). */
void pv3o_rescore_alignments(const int32_t *scoring5, uint64_t n, const uint64_t *seq_off, const uint8_t *seq,
	const uint64_t *qual_off, const uint8_t *qual, const uint64_t *cig_off, const uint32_t *cig,
	const uint64_t *refn_off, const uint8_t *refn, double *score, uint32_t *flags)
{
	int32_t mm[256], np[256];
	fill_penalties(mm, scoring5[0], scoring5[1]);
	fill_penalties(np, scoring5[2], scoring5[2]);
	int32_t const gap_open = scoring5[3], gap_ext = scoring5[4];
	for (uint64_t i = 0; i < n; ++i) {
		const uint8_t *q = seq + seq_off[i], *ql = qual + qual_off[i];
		uint64_t const nq = seq_off[i + 1] - seq_off[i], nl = qual_off[i + 1] - qual_off[i];
		const uint8_t *rn = refn_off ? refn + refn_off[i] : NULL;
		uint64_t const nn = refn_off ? refn_off[i + 1] - refn_off[i] : 0;
		uint64_t query_pos = 0, ref_pos = 0;
		char prev_op = 0;
		double retval = 0;
		uint32_t f = 0;
		for (uint64_t c = cig_off[i]; c < cig_off[i + 1] && !f; ++c) {
			uint32_t const count = cig[c] >> 4;
			char const op = "MIDNSHP=X???????"[cig[c] & 15];
			switch (op) {
				case 'H': case 'P': break;
				case 'N': ref_pos += count; break;
				case 'I':
					if (!('I' == prev_op || 'D' == prev_op)) retval -= gap_open;
					retval -= count * gap_ext;                  /* uint32_t * int32_t: unsigned, as in the reference */
					query_pos += count; prev_op = op; break;
				case 'D':
					if (!('I' == prev_op || 'D' == prev_op)) retval -= gap_open;
					retval -= count * gap_ext;
					ref_pos += count; prev_op = op; break;
				case 'S': query_pos += count; prev_op = op; break;
				case '=':
					for (uint64_t k = 0; k < count; ++k) {
						if (query_pos + k >= nq) { f = 4; break; }
						if (3 == q[query_pos + k]) {
							if (query_pos + k >= nl) { f = 0x40; break; }
							retval -= np[ql[query_pos + k]];
						}
					}
					query_pos += count; ref_pos += count; prev_op = op; break;
				case 'X':
					for (uint64_t k = 0; k < count; ++k) {
						uint64_t const qp = query_pos + k, rp = ref_pos + k;
						if (qp >= nq) { f = 4; break; }
						if (qp >= nl) { f = 0x40; break; }
						if (rp / 8 >= nn) { f = 8; break; }
						if (3 == q[qp] || ((0x1 << (rp % 8)) & rn[rp / 8])) retval -= np[ql[qp]];
						else retval -= mm[ql[qp]];
					}
					query_pos += count; ref_pos += count; prev_op = op; break;
				default: f = 1; break;
			}
		}
		score[i] = retval; flags[i] = f;
	}
}

static double read_min_score(uint32_t read_length) { if (!read_length) return 0; return -0.6 + (-0.6 * read_length); }   /* :455-460 */

typedef struct { double diff, score; uint8_t mapq; } nu_entry;
static const struct { double thr; uint8_t mapq; } unique_scores[7] = { {0.0, 0}, {0.3, 3}, {0.4, 8}, {0.5, 23}, {0.6, 24}, {0.7, 40}, {0.8, 42} };
static const nu_entry non_unique_scores[30] = {
	{0.0, 0.0, 2}, {0.0, 0.67, 6}, {0.1, 0.0, 0}, {0.1, 0.67, 7}, {0.1, 0.88, 12}, {0.1, 1.0, 30}, {0.2, 0.0, 0}, {0.2, 0.67, 11},
	{0.2, 0.88, 17}, {0.2, 1.0, 31}, {0.3, 0.0, 3}, {0.3, 0.67, 15}, {0.3, 0.88, 18}, {0.3, 1.0, 32}, {0.4, 0.0, 4}, {0.4, 0.68, 14},
	{0.4, 0.84, 21}, {0.4, 1.0, 34}, {0.5, 0.0, 5}, {0.5, 0.68, 16}, {0.5, 0.84, 25}, {0.5, 1.0, 35}, {0.6, 0.0, 22}, {0.6, 1.0, 36},
	{0.7, 0.0, 26}, {0.7, 1.0, 37}, {0.8, 0.0, 27}, {0.8, 1.0, 38}, {0.9, 0.0, 33}, {0.9, 1.0, 39} };

/* bowtie2_v2_score_calculator::calculate_mapq, main.cc:479-565. flags 0x10 where the reference would index before its table. */
void pv3o_calculate_mapq(uint64_t n, const uint32_t *read_len, const uint32_t *other_len, const double *score_in, const double *next_in,
	uint8_t *mapq, uint32_t *flags)
{
	for (uint64_t i = 0; i < n; ++i) {
		double const score = score_in[i];
		double const min_score = read_min_score(read_len[i]) + read_min_score(other_len[i]);
		double const max_score = 0;
		double const range0 = max_score - min_score;
		double const score_range = 1.0 < range0 ? range0 : 1.0;
		mapq[i] = 0; flags[i] = 0;
		if (score < min_score) continue;
		double const next_score = min_score < next_in[i] ? next_in[i] : min_score;
		double const nsq = (score - min_score) / score_range;
		double const diff_next = score - next_score;
		double const dnq = diff_next / score_range;
		if (-DBL_MAX == next_score) {
			int k = 0;
			while (k < 7 && !(nsq < unique_scores[k].thr)) ++k;
			if (!k) flags[i] = 0x10; else mapq[i] = unique_scores[k - 1].mapq;
		} else if (diff_next) {
			int it = 0;
			while (it < 30 && !(dnq < non_unique_scores[it].diff)) ++it;
			if (!it) { flags[i] = 0x10; continue; }
			double const thr = non_unique_scores[it - 1].diff;
			int k = 0;
			while (k < it && !(thr < non_unique_scores[k].diff || (!(non_unique_scores[k].diff < thr) && nsq < non_unique_scores[k].score))) ++k;
			if (!k) flags[i] = 0x10; else mapq[i] = non_unique_scores[k - 1].mapq;
		} else mapq[i] = nsq >= 0.67 ? 1 : 0;
	}
}
