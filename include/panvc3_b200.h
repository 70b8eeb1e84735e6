/*
 * panvc3_b200 — C-ABI of the B200-native project_alignments hot path.
 *
 * This is the drop-in boundary for the per-record work of PanVC 3's project_alignments
 * (reference: libpanvc3/alignment_projector.cc:30-154, called from
 * project-alignments/project_alignments.cc:721-732 and, for mates, :900).  One call processes a whole SoA batch of
 * records on one GPU: projection of the start position through the MSA gap vectors, CIGAR rewrite, indel-run
 * detection, realignment of the flagged runs and splice/collapse — bit-exact with the reference CPU code.
 *
 * Plain C, plain pointers and sizes; no CUDA or torch types.  All functions return PV3_OK (0) or a negative
 * pv3_status; pv3_last_error() gives the message.  THREADING: a context is single-submitter — calls on one context must
 * not overlap (no internal locking; its index tables, parameters, work buffers and streams are unsynchronised state).
 * Give every submitting thread its own context (the index is small next to 180 GB of HBM), or serialise the calls. A
 * context owns its CUDA streams and all device memory and is bound to one device (one context per GPU / per process
 * rank); pv3_multi below drives several contexts — one per GPU — from one thread with results in submission order.
 *
 * Encodings (what the reference's in-memory types already hold, so the adapter can memcpy):
 *   seq   one byte per base, seqan3::dna5 rank  A=0 C=1 G=2 N=3 T=4     (lib/seqan3/.../nucleotide/dna5.hpp:95-110)
 *   qual  one byte per base, seqan3::phred42 rank 0..41; an empty range selects the integer DP
 *         (alignment_projector.cc:76,97)
 *   cig   uint32  len << 4 | op,  op order M I D N S H P = X  (BAM)
 *   gap rows  uint64 words, LSB first, bit = 1 where the aligned row has '-'   (msa_index.hh:22-32,58-65)
 *   reference contigs  raw FASTA bytes, no case folding (rewrite_cigar.hh:104-110 compares raw bytes)
 */
#ifndef PANVC3_B200_H
#define PANVC3_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PV3_ABI_VERSION 6

typedef struct pv3_ctx pv3_ctx;

typedef enum {
	PV3_OK = 0,
	PV3_ERR_INVALID = -1,   /* bad argument */
	PV3_ERR_CUDA = -2,      /* CUDA runtime failure (no device, OOM, launch error) — never falls back to the CPU */
	PV3_ERR_CAPACITY = -3,  /* an output buffer was too small; required sizes are in pv3_batch_out.{cig,rr}_used */
	PV3_ERR_STATE = -4      /* call order (e.g. project before index_commit) */
} pv3_status;

/* Per-record flags: the record was rejected where the reference would abort or has undefined behaviour.
 * Outputs of a flagged record are unspecified (cig_len = 0, rr_len = 0). */
#define PV3_FLAG_BAD_OP      0x1u   /* CIGAR op outside MIDNSHP=X            (rewrite_cigar.hh:149-151 libbio_fail) */
#define PV3_FLAG_POS_OOB     0x2u   /* reference-consuming base past the last non-gap of the source row
                                       (select0 returns size(), rrr_vector_15.hpp:666, then OOB access) */
#define PV3_FLAG_QUERY_OOB   0x4u   /* query_pos >= |SEQ|                    (rewrite_cigar.hh:102 assert) */
#define PV3_FLAG_REF_OOB     0x8u   /* excess >= |dst contig|                (rewrite_cigar.hh:103 assert) */
#define PV3_FLAG_TRACE_UB    0x10u  /* affine traceback leaves the matrix — reference UB, only with gap_open < 0
                                       (affine_gap_policy.hpp:96-101) */
#define PV3_FLAG_EMPTY_CIGAR 0x20u  /* final CIGAR empty: collapse_cigar_operations UB (cigar.cc:16-18) */
#define PV3_FLAG_BAD_QUAL    0x40u  /* non-empty QUAL whose length differs from SEQ */
#define PV3_FLAG_BAD_ID      0x80u  /* row or contig id not in the index (the CLI exits on failed lookups,
                                       project_alignments.cc:113-114) */

/* Input batch, structure of arrays. Replaces the reference's per-record project_task slots
 * (project_alignments.cc:219-330) and the arguments of alignment_projector::project_alignment
 * (include/panvc3/alignment_projector.hh:58-68). */
typedef struct {
	uint64_t        n;         /* records */
	const uint32_t *src_row;   /* [n] row id (pv3_index_add_row) of the sequence the read was aligned to  — src_seq_entry */
	const uint32_t *dst_row;   /* [n] row id of the projection target (e.g. REF)                          — dst_seq_entry */
	const uint32_t *dst_ref;   /* [n] contig id (pv3_index_add_ref) of the target's unaligned bases       — ref_seq */
	const uint64_t *src_pos;   /* [n] 0-based position in the source sequence                             — src_pos */
	const uint64_t *seq_off;   /* [n+1] */
	const uint8_t  *seq;       /* dna5 ranks                                                              — query_seq */
	const uint64_t *qual_off;  /* [n+1], or NULL when no record has QUAL                                  — base_qualities */
	const uint8_t  *qual;      /* phred42 ranks */
	const uint64_t *cig_off;   /* [n+1] */
	const uint32_t *cig;       /* BAM-packed ops                                                          — cigar_seq */
	const uint32_t *seq_len;   /* NULL (default): seq as above. Non-NULL: SEQ as a BAM file stores it — 4-bit codes "=ACMGRSVTWYHKDBN",
	                              two bases per byte, first base in the high nibble, every record starting on a byte; seq_off
	                              then counts BYTES and seq_len[i] is the record's number of bases. Codes other than A C G T are
	                              read as N, as seqan3 does when it reads a BAM SEQ into dna5 (no unpacking on
	                              the host, half the bytes over PCIe). */
	/* -- ABI 5 (all optional; zero-initialise the struct) ------------------------------------------------------------- */
	const uint32_t *mate_row;  /* [n] row id of the sequence the MATE was aligned to (RNEXT); its position is projected to
	                              dst_row with sequence_entry::project_position (project_alignments.cc:880-906). NULL: no mates */
	const uint64_t *mate_pos;  /* [n] */
	/* Compact encoding (pv3_project only, compact != 0): the arrays below are what crosses PCIe instead of src_row,
	 * dst_row, dst_ref, src_pos, seq_off, cig_off, qual_off, mate_row, mate_pos (u16 ids, u32 positions, lengths instead of
	 * offsets: 16 B per record instead of 52). seq / cig / qual hold the records back to back in record order; seq_len
	 * counts bases (SEQ is BAM 4-bit packed iff seq_bam4); QUAL, where present (bit i of has_qual), is as long as SEQ.
	 * seq_off, cig_off and qual_off must still be valid ON THE HOST (the call cuts the batch into chunks with them; a
	 * packer has them as running totals anyway) but are not uploaded. */
	int             compact;
	int             seq_bam4;
	const uint16_t *src_row16, *dst_row16, *dst_ref16, *mate_row16;
	const uint32_t *src_pos32, *mate_pos32;
	const uint16_t *cig_len16;
	const uint8_t  *has_qual;  /* bit i & 7 of byte i / 8; NULL: no record has QUAL */
	/* -- ABI 6 (optional) ---------------------------------------------------------------------------------------------
	 * Two-bit SEQ (pv3_project / pv3_multi_project with compact != 0, instead of seq_bam4): seq holds A C G T as 0 1 2 3,
	 * FOUR bases per byte, base 4 k + i of a record in bits 2 i .. 2 i + 1 of its byte k, every record starting on a byte;
	 * seq_off (host) counts these bytes. Bases that are not A C G T (dna5's N) are listed apart: exception j says that base
	 * exc_off[j] of record exc_rec[j] is N (its two bits are ignored); entries sorted by record. SEQ is 3/4 of what a
	 * 150-bp record sends over PCIe in the compact encoding, so this is 60 instead of 97 bytes per record. The device
	 * expands it to the 4-bit layout the kernels read (one pass, 0.03 ms per 10^6 records). */
	int             seq_2bit;
	uint64_t        n_seq_exceptions;
	const uint32_t *exc_rec, *exc_off;
} pv3_batch_in;

/* Output batch. CIGARs and realigned ranges live in arenas and are addressed per record by (off, len); arena order
 * is not record order. Mirrors the accessors alignment(), realigned_reference_ranges(), realigned_query_ranges()
 * (alignment_projector.hh:70-72) and project_alignment's return value. */
typedef struct {
	uint64_t *dst_pos;   /* [n] projected position */
	uint32_t *flags;     /* [n] PV3_FLAG_* */
	uint64_t *cig_off;   /* [n] */
	uint32_t *cig_len;   /* [n] */
	uint32_t *cig;       /* [cig_cap] arena */
	uint64_t  cig_cap;
	uint64_t *rr_off;    /* [n] */
	uint32_t *rr_len;    /* [n] realigned runs of the record */
	uint64_t *rr;        /* [4*rr_cap]: ref_location (as the reference reports it, alignment_projector.cc:69-72), ref_length,
	                        query_location, query_length */
	uint64_t  rr_cap;
	/* filled by the call */
	uint64_t  cig_used;        /* arena words used (or required, with PV3_ERR_CAPACITY) */
	uint64_t  rr_used;         /* runs written (or required) */
	uint64_t  n_flagged;       /* records with flags != 0 */
	uint64_t  n_realigned;     /* records with at least one realigned run */
	uint64_t  dp_cells;        /* sum over runs of ref_length * query_length (GCUPS numerator) */
	/* -- ABI 5 (all optional; zero-initialise the struct) ------------------------------------------------------------- */
	uint64_t *mate_dst_pos;    /* [n] projected mate positions (pv3_batch_in.mate_row set) */
	/* pv3_project writes the arenas in RECORD ORDER: cig_off[i + 1] == cig_off[i] + cig_len[i], likewise rr. Offsets are
	 * therefore redundant: cig_off, rr_off, flags, rr_len may be NULL (not written), and each of the narrow arrays below,
	 * when non-NULL, is written INSTEAD of its wide twin (39 B per 150-bp record over PCIe instead of 67). A value that
	 * does not fit its narrow field fails the call with PV3_ERR_CAPACITY (retry with the wide field).
	 * pv3_project_device keeps arena order (one atomic allocation per tile) and ignores these. */
	uint32_t *dst_pos32;
	uint8_t  *flags8;
	uint16_t *cig_len16;
	uint8_t  *rr_len8;
	uint32_t *mate_dst_pos32;
} pv3_batch_out;

/* Device-side durations of the last pv3_project or pv3_project_device call, from CUDA events on the context's stream. */
typedef struct {
	float    ms_rewrite;     /* projection + CIGAR rewrite + run detection kernel */
	float    ms_realign;     /* DP + traceback kernel (0 when no run was found) */
	float    ms_splice;      /* splice + collapse kernel */
	float    ms_total;       /* first kernel start to last kernel end */
	float    ms_h2d, ms_d2h; /* host path only */
	uint64_t launches;       /* kernels launched by the call */
	uint64_t n_deferred;     /* records the thread-per-record rewrite kernel handed to the warp-per-record kernel */
	float    ms_rewrite_main;/* the thread-per-record rewrite kernel alone (ms_rewrite also holds the deferred pass) */
	float    ms_msa_extract; /* last pv3_msa_row_extract[_device]: the extraction kernel alone */
} pv3_timing;

/* -- lifetime ------------------------------------------------------------------------------------------------------ */
int         pv3_abi_version(void);
int         pv3_create(pv3_ctx **out, int device);          /* fails with PV3_ERR_CUDA when no usable GPU exists */
void        pv3_destroy(pv3_ctx *);
const char *pv3_last_error(const pv3_ctx *);                 /* ctx may be NULL: error of the last failed pv3_create */

/* Runs all work of the context on the caller's CUDA stream (a cudaStream_t passed as void*; NULL restores the context's
 * own stream), so that callers can bracket calls with their own events or order them after their own kernels. */
int         pv3_set_stream(pv3_ctx *, void *cuda_stream);

/* -- index (replaces msa_index::sequence_entry over sdsl::rrr_vector<15>, include/panvc3/msa_index.hh:47-74) --------- */
int pv3_index_add_row(pv3_ctx *, const uint64_t *gap_words, uint64_t n_bits);  /* returns row id >= 0 */
int pv3_index_add_ref(pv3_ctx *, const char *bases, uint64_t n);               /* returns contig id >= 0 */
int pv3_index_commit(pv3_ctx *);                /* builds rank/select directories and uploads what was added since the last
                                                   commit (every row and contig has its own device allocation; the host copies
                                                   are dropped). Rows/contigs may still be added afterwards, ids are stable */
int pv3_index_remove_ref(pv3_ctx *, int ref_id);/* frees a contig's device memory; the id is reused by a later pv3_index_add_ref
                                                   (sequence_buffer_store recycles contig buffers the same way,
                                                   project_alignments.cc:330-380). Records must no longer refer to it */

/* -- parameters (project_alignments -g / -e; defaults 0 / -1, project-alignments/cmdline.ggo) ------------------------ */
int pv3_set_params(pv3_ctx *, int32_t gap_open, int32_t gap_ext);
int pv3_set_quality_lut(pv3_ctx *, const int32_t *lut84);   /* optional override of the Malde LUT (align.hh:28-37) */
int pv3_get_quality_lut(const pv3_ctx *, int32_t *lut84);

/* -- the hot path -------------------------------------------------------------------------------------------------- */
/* Host buffers in and out; H2D copies, kernels and the write-back all happen inside the call, which synchronises ONCE,
 * at its end. Use pinned memory (pv3_host_alloc / pv3_host_register): results are then written straight into the
 * caller's arrays by the last kernel of every chunk and QUAL is read in place, only where a run is realigned. Pageable
 * memory works (staged copies, one synchronisation per chunk, QUAL uploaded in full) but is slower. */
int pv3_project(pv3_ctx *, const pv3_batch_in *in, pv3_batch_out *out);
/* Every pointer inside *in and *out is a DEVICE pointer on the context's device (the structs themselves are host). */
int pv3_project_device(pv3_ctx *, const pv3_batch_in *in, pv3_batch_out *out);
/* msa_index::sequence_entry::project_position (msa_index.hh:253-258) for n independent positions (mates,
 * project_alignments.cc:880-906). Host buffers. */
int pv3_project_positions(pv3_ctx *, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row, const uint64_t *pos,
                          uint64_t *out_pos);
/* The same with device pointers, asynchronous on the context's stream. */
int pv3_project_positions_device(pv3_ctx *, uint64_t n, const uint32_t *src_row, const uint32_t *dst_row, const uint64_t *pos,
                                 uint64_t *out_pos);
/* Raw index queries for parity tests: op 0 = rank0(row,x) zeros in [0,x); 1 = select0(row,x) (x >= 1); 2 = access. */
int pv3_index_query(pv3_ctx *, int op, uint64_t n, const uint32_t *row, const uint64_t *x, uint64_t *out);

/* Parity probe: stage 1 stops after rewrite_cigar() (include/panvc3/rewrite_cigar.hh:19) — no run detection, no
 * realignment; stage 0 (default) is the full project_alignment. */
int pv3_set_stage(pv3_ctx *, int stage);
/* Parity probe: indel runs with small windows are realigned one thread per run once a batch holds at least `n` of them
 * (default 16384), otherwise one warp per run like the large windows; 0 sends every small run to the thread-per-run kernel.
 * Results are identical. */
int pv3_set_dp_thread_min(pv3_ctx *, uint64_t n);
/* Parity probe: 0 (default) = thread-per-record rewrite kernel (run walk), records beyond its quotas go to the
 * warp-per-record kernel; 1 = warp-per-record kernel for every record; 2 = thread-per-record word-parallel kernel, records
 * beyond its quotas go to the warp-per-record kernel. Results are identical. */
int pv3_set_kernel_path(pv3_ctx *, int path);

int pv3_get_timing(const pv3_ctx *, pv3_timing *out);
uint64_t pv3_total_launches(const pv3_ctx *);

/* -- pinned host memory for batches --------------------------------------------------------------------------------- */
void *pv3_host_alloc(size_t bytes);
void  pv3_host_free(void *);
/* Pins (page-locks and maps) memory the caller allocated itself — e.g. a POSIX shared-memory segment into which several
 * ranks write their shards of one ordered output. */
int   pv3_host_register(void *p, size_t bytes);
int   pv3_host_unregister(void *p);

/* -- several GPUs behind one submitter --------------------------------------------------------------------------------
 * One context per device, the index replicated; pv3_multi_project cuts the batch into contiguous shards (record order is
 * kept by construction: shard g holds records [n g / G, n (g + 1) / G)), runs every shard's asynchronous pipeline on its
 * device from the calling thread and synchronises once. Output is ordered across the whole batch as for pv3_project.
 * The reference's contract: one submitter, records out in input order (project_alignments.cc:958-970). */
typedef struct pv3_multi pv3_multi;
int  pv3_multi_create(pv3_multi **out, int n_devices, const int *devices /* NULL: 0..n-1 */);
void pv3_multi_destroy(pv3_multi *);
int  pv3_multi_devices(const pv3_multi *);
pv3_ctx *pv3_multi_ctx(pv3_multi *, int i);     /* context of device i: add rows / contigs / params to EVERY context */
int  pv3_multi_project(pv3_multi *, const pv3_batch_in *in, pv3_batch_out *out);
const char *pv3_multi_last_error(const pv3_multi *);

/* -- record annotation text ------------------------------------------------------------------------------
 * The OA entry project_task::process appends to every projected record (project-alignments/project_alignments.cc:803-845):
 *     RNAME ',' POS ',' ('-' if FLAG & 0x10 else '+') ',' CIGAR ',' MAPQ ',' [NM] ';'
 * with POS printed as the reference holds it (seqan3's 0-based reference_position), the ORIGINAL CIGAR as decimal count +
 * op char, and NM only when the record had the tag. One string per record into a text arena: record i is
 * text[text_off[i] .. text_off[i+1]), no terminators. Host buffers. PV3_ERR_CAPACITY (with *text_used = bytes needed) when
 * text_cap is too small; calling with text_cap = 0 is the way to size the buffer. */
typedef struct {
	uint64_t        n;
	const uint32_t *ref_id;    /* [n] RNAME of the source record as an index into the name table (ref_ids[ref_id], :812) */
	const uint64_t *pos;       /* [n] src_pos (:815) */
	const uint16_t *flag;      /* [n] SAM FLAG; only on_reverse_strand (0x10) is read (:819) */
	const uint8_t  *mapq;      /* [n] (:830) */
	const int32_t  *nm;        /* [n] original NM value (:835) */
	const uint8_t  *has_nm;    /* [n] 0 = the record had no NM tag; NULL = no record has one */
	const uint64_t *cig_off;   /* [n+1] */
	const uint32_t *cig;       /* BAM-packed ops of the ORIGINAL alignment (the same arrays as pv3_batch_in) */
	uint32_t        n_names;
	const uint64_t *name_off;  /* [n_names+1] */
	const char     *names;     /* header reference names, concatenated */
} pv3_oa_in;
int pv3_format_oa(pv3_ctx *, const pv3_oa_in *in, uint64_t *text_off /* [n+1] */, char *text, uint64_t text_cap,
                  uint64_t *text_used);
/* SAM text of n CIGARs (count + op char per op; an empty CIGAR gives an empty string, the caller prints '*'). Record i is
 * cig[cig_off[i] .. +cig_len[i]) — the (off, len) addressing of pv3_batch_out — or, with cig_len == NULL,
 * cig[cig_off[i] .. cig_off[i+1]). n_cig_words = size of the cig array. */
int pv3_format_cigars(pv3_ctx *, uint64_t n, const uint64_t *cig_off, const uint32_t *cig_len, const uint32_t *cig,
                      uint64_t n_cig_words, uint64_t *text_off /* [n+1] */, char *text, uint64_t text_cap, uint64_t *text_used);

/* -- recalculate-mapq: the two per-record computations ------------------------------------------------------
 * pv3_rescore_alignments: cigar_alignment_scorer::calculate_score (recalculate-mapq/main.cc:190-283) — the alignment score
 * of every record from its CIGAR (= and X, not M), base qualities, N bases of the query and the N positions of the reference
 * (the XN tag, a bit per aligned reference position); penalties per quality rank as fill_penalties builds them (:181-186).
 * score[i] is what the reference assigns to AS (a double holding an integer). Flags: PV3_FLAG_BAD_OP (an M or unknown op:
 * libbio_fail :274-277), PV3_FLAG_REF_OOB (an X base beyond the XN bitmap: libbio_always_assert_lt :264),
 * PV3_FLAG_QUERY_OOB / PV3_FLAG_BAD_QUAL (SEQ or QUAL shorter than the CIGAR consumes: out-of-range reads in the reference).
 * Host buffers; refn_off == NULL when no record has the XN tag. */
typedef struct {
	int32_t min_mismatch_penalty, max_mismatch_penalty, n_penalty, gap_opening_penalty, gap_extension_penalty;   /* cmdline.ggo: 2 6 1 5 3 */
} pv3_scoring;
int pv3_rescore_alignments(pv3_ctx *, const pv3_scoring *, uint64_t n, const uint64_t *seq_off, const uint8_t *seq,
                           const uint64_t *qual_off, const uint8_t *qual, const uint64_t *cig_off, const uint32_t *cig,
                           const uint64_t *refn_off, const uint8_t *refn, double *score /* [n] */, uint32_t *flags /* [n] */);
/* pv3_calculate_mapq: bowtie2_v2_score_calculator::calculate_mapq (main.cc:455-565, tables :378-430) for n independent
 * (read length, mate length or 0, score, next best score or -DBL_MAX) tuples, in f64 with the reference's operation order
 * and no fused multiply-add. PV3_FLAG_TRACE_UB marks tuples for which the reference indexes before its table (next > score). */
int pv3_calculate_mapq(pv3_ctx *, uint64_t n, const uint32_t *read_len, const uint32_t *other_len, const double *score,
                       const double *next_score, uint8_t *mapq /* [n] */, uint32_t *flags /* [n] */);

/* -- index-msa on the device --------------------------------------------------------------------------------------------
 * The per-sequence work of `panvc3 index-msa` (index-msa/index_vector_builder.cc:28-50, driven from
 * input_processor.cc:155-175): bit i of gap_words = 1 ⇔ text[i] == '-', and every other byte is appended to the unaligned
 * sequence (what the tool prints as FASTA) — exactly the two inputs pv3_index_add_row / pv3_index_add_ref take, and, with
 * pv3_index_file_write, the index file. gap_words holds (n + 63) / 64 words; ungapped has capacity n and may be NULL (no
 * sequence wanted); *n_ungapped receives the number of non-gap bytes. Needs no committed index. The _device variant
 * takes device pointers (text 16-byte aligned and readable to the next multiple of 16). */
int pv3_msa_row_extract(pv3_ctx *, const char *text, uint64_t n, uint64_t *gap_words, char *ungapped, uint64_t *n_ungapped);
int pv3_msa_row_extract_device(pv3_ctx *, const char *text, uint64_t n, uint64_t *gap_words, char *ungapped,
                               uint64_t *n_ungapped /* host */);

/* No counterpart in the reference. pv3_msa_row_extract_device for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): needs no context and
 * allocates nothing — the kernels' scratch (tile counts and offsets) lives in a fixed array in device memory (768 KiB, held by
 * every GPU that loads the library until exit), which limits a call to 2^30 columns (PV3_ERR_CAPACITY beyond; message through
 * pv3_last_error(NULL)). All pointers are device pointers on `device`, aligned as for pv3_msa_row_extract_device;
 * n_ungapped (8 bytes, may be NULL) is a DEVICE pointer too. Asynchronous on cuda_stream (a cudaStream_t; NULL = the default
 * stream), no synchronisation. All calls on a device share the scratch: calls that overlap on the GPU (different streams
 * without ordering) race. */
int pv3_msa_row_extract_noalloc(int device, void *cuda_stream, const char *text, uint64_t n, uint64_t *gap_words, char *ungapped,
                                uint64_t *n_ungapped /* device */);

/* -- MSA index file (host only, no GPU needed) ---------------------------------------------------------------------------
 * The file `panvc3 index-msa` writes and `project-alignments --msa-index` reads: panvc3::msa_index through
 * cereal::PortableBinary{Output,Input}Archive (include/panvc3/msa_index.hh:40-192; written at
 * index-msa/input_processor.cc:186-200, loaded at project-alignments/project_alignments.cc:1457-1465). Every gap vector is
 * an sdsl::rrr_vector<15> (lib/sdsl-lite/include/sdsl/rrr_vector_15.hpp:197-273, 389-407); pv3_index_file_open expands
 * them to the plain words pv3_index_add_row takes (bit i of word i/64 = 1 ⇔ column i is a gap), and pv3_index_file_write
 * produces the same bytes the reference would for the same rows. Entries come back in file order (chr ids sorted, seq ids
 * sorted inside each). Errors: PV3_ERR_INVALID with pv3_index_file_error() (not thread-safe, like pv3_last_error(NULL)). */
typedef struct pv3_index_file pv3_index_file;
int pv3_index_file_open(const char *path, pv3_index_file **out);
uint64_t pv3_index_file_entries(const pv3_index_file *);
/* Pointers stay valid until pv3_index_file_close; any out argument may be NULL. */
int pv3_index_file_entry(const pv3_index_file *, uint64_t i, const char **chr_id, const char **seq_id,
                         const uint64_t **gap_words, uint64_t *n_bits);
void pv3_index_file_close(pv3_index_file *);
int pv3_index_file_write(const char *path, uint64_t n_entries, const char *const *chr_ids, const char *const *seq_ids,
                         const uint64_t *const *gap_words, const uint64_t *n_bits);
const char *pv3_index_file_error(void);

#ifdef __cplusplus
}
#endif
#endif
