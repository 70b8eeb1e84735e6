/*
 * TEST INFRASTRUCTURE ONLY — flat C interface shared by the two CPU checkers:
 *   oracle/_ref/libpv3_ref.so     the reference's own libpanvc3 sources compiled in place (ref_driver.cc)
 *   oracle/libpv3_oracle.so       the plain-C restatement (pv3_oracle.c)
 * Both export exactly these symbols, so tests can run either behind the same ctypes binding.
 * Nothing under panvc3_b200/ (the product) may include, link or load this.
 *
 * Encodings (same as the product C-ABI, include/panvc3_b200.h):
 *   seq   dna5 ranks   A=0 C=1 G=2 N=3 T=4        (seqan3 dna5.hpp:95-110)
 *   qual  phred42 rank 0..41; empty range ⇒ integer mode (alignment_projector.cc:76)
 *   cig   BAM packing  len<<4 | op,  op order M I D N S H P = X  (0..8)
 *   gap rows: plain 64-bit words, LSB first, bit=1 ⇔ gap ('-')  (msa_index.hh:22-32)
 */
#ifndef PV3_ORACLE_API_H
#define PV3_ORACLE_API_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pv3o_index pv3o_index;

/* per-record flags shared with the product (include/panvc3_b200.h) */
#define PV3O_FLAG_BAD_OP        0x1u   /* CIGAR op outside MIDNSHP=X  (rewrite_cigar.hh:150 libbio_fail) */
#define PV3O_FLAG_POS_OOB       0x2u   /* select0 past the last non-gap (rrr_vector_15.hpp:666) */
#define PV3O_FLAG_QUERY_OOB     0x4u   /* query_pos >= |query| (rewrite_cigar.hh:102) */
#define PV3O_FLAG_REF_OOB       0x8u   /* excess >= |dst_seq| (rewrite_cigar.hh:103) or DP window past contig */
#define PV3O_FLAG_TRACE_UB      0x10u  /* affine traceback walks off row/col 0 (reference UB, gap_open<0) */
#define PV3O_FLAG_EMPTY_CIGAR   0x20u  /* collapse_cigar_operations on empty vector (cigar.cc:16-18) */
#define PV3O_FLAG_BAD_QUAL      0x40u  /* |QUAL| != |SEQ| for a non-empty QUAL */

const char *pv3o_kind(void); /* "reference" or "port" */

pv3o_index *pv3o_index_create(void);
void        pv3o_index_destroy(pv3o_index *);
int64_t     pv3o_index_add_row(pv3o_index *, const uint64_t *gap_words, uint64_t n_bits);
int64_t     pv3o_index_add_ref(pv3o_index *, const char *bases, uint64_t n);

uint64_t pv3o_rank0(const pv3o_index *, uint32_t row, uint64_t pos);   /* zeros in [0,pos) */
uint64_t pv3o_select0(const pv3o_index *, uint32_t row, uint64_t k);   /* position of k-th zero, k>=1; n_bits if none */
int      pv3o_access(const pv3o_index *, uint32_t row, uint64_t pos);
uint64_t pv3o_project_position(const pv3o_index *, uint32_t src_row, uint32_t dst_row, uint64_t pos);

/*
 * stage 0: full alignment_projector::project_alignment; stage 1: rewrite_cigar only (no realignment).
 * Outputs are CSR in record order. rr holds 4 u64 per realigned run:
 *   reported_ref_location, ref_length, query_location, query_length  (alignment_projector.cc:69-73).
 * Returns 0, or -1 if an output capacity was too small.
 * The reference build aborts where the reference aborts; the port sets flags instead.
 */
int pv3o_project_batch(const pv3o_index *, uint64_t n,
	const uint32_t *src_row, const uint32_t *dst_row, const uint32_t *dst_ref, const uint64_t *src_pos,
	const uint64_t *seq_off, const uint8_t *seq,
	const uint64_t *qual_off, const uint8_t *qual,      /* qual_off may be NULL ⇒ no record has QUAL */
	const uint64_t *cig_off, const uint32_t *cig,
	int32_t gap_open, int32_t gap_ext, int n_threads, int stage,
	uint64_t *dst_pos,
	uint64_t *out_cig_off /* n+1 */, uint32_t *out_cig, uint64_t out_cig_cap,
	uint64_t *rr_off /* n+1 */, uint64_t *rr, uint64_t rr_cap /* in runs */,
	uint32_t *flags);

/*
 * One global alignment as panvc3::align_global (align.hh:66-113): ref = DP columns, qry = DP rows.
 * qual == NULL ⇒ align_global<false>; else align_global<true, dna5, phred42> with the reference at max quality.
 * Returns the number of ops written (BAM packed, I D = X only), or -1 on overflow, -2 on trace UB (port only).
 */
int64_t pv3o_align_global(const uint8_t *ref_ranks, uint64_t n, const uint8_t *qry_ranks, uint64_t m,
	const uint8_t *qual, int32_t gap_open, int32_t gap_ext,
	uint32_t *out_ops, uint64_t cap, int32_t *score);

/* indel_run_checker alone (indel_run_checker.cc:31-121): ops BAM packed; out 6 u64 per run:
 * op_begin, op_end, ref_loc, ref_len, qry_loc, qry_len (true DP window, not the shifted report).
 * final[0]=query_position(), final[1]=reference_position() after the scan. Returns run count. */
int64_t pv3o_indel_runs(const uint32_t *ops, uint64_t n_ops, uint64_t ref_pos, uint64_t *out, uint64_t cap, uint64_t *final_pos);

/* The quality LUT (align.hh:28-37 truncated to int32): lut[2*q+0]=match, lut[2*q+1]=mismatch, q=0..41 */
void pv3o_quality_lut(int32_t *lut84);

/* index-msa, one MSA row: bit i = 1 iff text[i] == '-' (index_vector_builder.cc:28-50), every other byte is copied to the
 * unaligned sequence in order (:40-43). gap_words: (n + 63) / 64 words; ungapped: capacity n, may be NULL. */
/* OA entry of every record (project_alignments.cc:803-845), or with ref_id == NULL the SAM text of every CIGAR (port only).
 * Reference build: the reference's own statements, ref_driver_annotate.cc (also pv3o_filter_tags, :774-800). Returns the bytes needed; text_off has n + 1 entries. */
uint64_t pv3o_format_oa(uint64_t n, const uint32_t *ref_id, const uint64_t *pos, const uint16_t *flag, const uint8_t *mapq,
	const int32_t *nm, const uint8_t *has_nm, const uint64_t *cig_off, const uint32_t *cig_len, const uint32_t *cig,
	uint32_t n_names, const uint64_t *name_off, const char *names, uint64_t *text_off, char *text, uint64_t cap);

/* recalculate-mapq (recalculate-mapq/main.cc:181-283 and :455-565). Port: pv3_oracle.c; reference: ref_driver_mapq.cc (the
 * reference's own code, excerpted at build time). In the reference build flags are 0xffffffff where the reference aborted.
 * scoring5 = min/max mismatch, N, gap opening, gap extension penalty. */
void pv3o_rescore_alignments(const int32_t *scoring5, uint64_t n, const uint64_t *seq_off, const uint8_t *seq,
	const uint64_t *qual_off, const uint8_t *qual, const uint64_t *cig_off, const uint32_t *cig,
	const uint64_t *refn_off, const uint8_t *refn, double *score, uint32_t *flags);
void pv3o_calculate_mapq(uint64_t n, const uint32_t *read_len, const uint32_t *other_len, const double *score, const double *next_score,
	uint8_t *mapq, uint32_t *flags);

void pv3o_msa_row_extract(const char *text, uint64_t n, uint64_t *gap_words, char *ungapped, uint64_t *n_ungapped);

#ifdef __cplusplus
}
#endif
#endif
