/* bwamem_b200.h -- C ABI of the B200-native BWA-MEM alignment core (libbwamem_b200.so).
 *
 * Plain C: pointers and sizes only, no CUDA/torch types.  The struct layouts below are this
 * fork's layouts (sunwookim028/bwa-mem-gpu, bwa 0.7.17-r1198), re-declared here so that the
 * reference's host code (bwa mem CLI, bwtindex, SAM formatting) can link against this library
 * without its own bwt.c / ksw.c / bwamem.c seeding-chaining-extension code.  Each entry point cites
 * the reference interface it replaces (file:line under the reference tree).
 *
 * If the reference's own headers are already included (bwa.h / bwt.h / ksw.h), define
 * BWAMEM_B200_NO_TYPES before including this file to skip the type re-declarations.
 */
#ifndef BWAMEM_B200_H_
#define BWAMEM_B200_H_

#include <stdint.h>
#include <stddef.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef BWAMEM_B200_NO_TYPES
/* ---- FM-index (bwt.h:46-66) */
typedef uint64_t bwtint_t;
typedef struct {
	bwtint_t primary;          /* S^{-1}(0) */
	bwtint_t L2[5];            /* cumulative base counts */
	bwtint_t seq_len;          /* forward + reverse-complement text length */
	bwtint_t bwt_size;         /* number of uint32 words in bwt[] */
	uint32_t *bwt;             /* 64-byte blocks: 4 x uint64 Occ checkpoint + 128 bases 2-bit BWT */
	uint32_t cnt_table[256];
	int sa_intv;
	bwtint_t n_sa;
	bwtint_t *sa;
} bwt_t;
typedef struct { bwtint_t x[3], info; } bwtintv_t;
typedef struct { int n, m; bwtintv_t *a; } bwtintv_v;

/* ---- reference sequence store (bntseq.h:41-64) */
typedef struct {
	int64_t offset;
	int32_t len;
	int32_t n_ambs;
	uint32_t gi;
	int32_t is_alt;
	char *name, *anno;
} bntann1_t;
typedef struct { int64_t offset; int32_t len; char amb; } bntamb1_t;
typedef struct {
	int64_t l_pac;
	int32_t n_seqs;
	uint32_t seed;
	bntann1_t *anns;
	int32_t n_holes;
	bntamb1_t *ambs;
	FILE *fp_pac;
} bntseq_t;

/* ---- options and results (bwa.h:86-171) */
typedef struct {
	uint64_t max_mem_intv;
	int a, b;
	int o_del, e_del;
	int o_ins, e_ins;
	int pen_unpaired;
	int pen_clip5, pen_clip3;
	int w;
	int zdrop;
	int T;
	int flag;
	int min_seed_len;
	int min_chain_weight;
	int max_chain_extend;
	float split_factor;
	int split_width;
	int max_occ;
	int max_chain_gap;
	int n_threads;
	int chunk_size;
	float mask_level;
	float drop_ratio;
	float XA_drop_ratio;
	float mask_level_redun;
	float mapQ_coef_len;
	int mapQ_coef_fac;
	int max_ins;
	int max_matesw;
	int max_XA_hits, max_XA_hits_alt;
	int8_t mat[25];
} mem_opt_t;

typedef struct {
	int64_t rb, re;
	uint64_t hash;
	float frac_rep;
	int qb, qe;
	int rid;
	int score;
	int truesc;
	int sub;
	int alt_sc;
	int csub;
	int sub_n;
	int w;
	int seedcov;
	int secondary;
	int secondary_all;
	int seedlen0;
	int n_comp:30, is_alt:2;
} mem_alnreg_t;
typedef struct { int n, m; mem_alnreg_t *a; } mem_alnreg_v;

typedef struct { int low, high; int failed; double avg, std; } mem_pestat_t;

typedef struct {
	int l_seq, id;
	char *name, *comment, *seq, *qual, *sam;
	int8_t l_name, l_comment;
	int16_t l_qual;
} bseq1_t;

#define MEM_F_PE        0x2
/* ksw.h:11-19 */
struct _kswq_t;
typedef struct _kswq_t kswq_t;
typedef struct { int score; int te, qe; int score2, te2; int tb, qb; } kswr_t;
#endif /* BWAMEM_B200_NO_TYPES */

/* =====================================================================================
 * Context and device index loader
 * replaces: transferIndex / newProcess (cuda/streams.cu:8-104,164) -- blocking cudaMalloc+cudaMemcpy
 * of bwt/sa/bns/pac.  Here: one upload per GPU, the 64-byte Occ blocks verbatim (128-bit aligned),
 * the sampled SA expanded once on the device into a full SA (HBM is large), contig table and pac.
 * ===================================================================================== */
typedef struct b200_ctx b200_ctx_t;

/* Error returns of the calls whose non-negative result is a COUNT and whose small negative results mean "-(needed)":
 * b200_align_batch, b200_batch_download, b200_align_sam_se_batch.  Anything <= B200_E_CAPACITY is an error (message in
 * b200_last_error()), never a count. */
#define B200_E_FAIL      (INT64_MIN + 1)   /* CUDA error, missing index, out of memory ... */
#define B200_E_CAPACITY  (INT64_MIN + 2)   /* an internal capacity was exceeded and re-running with the reported need did not help */
#define B200_IS_ERROR(rc) ((int64_t)(rc) <= B200_E_CAPACITY)

b200_ctx_t *b200_ctx_create(int device);
void        b200_ctx_destroy(b200_ctx_t *ctx);
const char *b200_last_error(void);
/* tunables: "chunk_reads" (reads per pipeline chunk), "smem_lcap" (shared-memory list entries per read in the
 * seeding kernel; longer lists take the general kernel), "cap_intv" (interval slots per read; 0 = auto; a chunk that
 * needs more is re-run with the reported need unless "cap_retry" is 0, in which case the call fails), "kmer_k" (depth of the
 * k-mer interval table, before the index upload; -1 = by index size, 9..15), "seed_stats" (1: the seeding kernels count
 * bwt_extend calls / two-block calls for b200_batch_counters; default 0) */
int         b200_ctx_tune(b200_ctx_t *ctx, const char *key, int64_t value);
/* flags for b200_index_upload */
#define B200_IDX_FULL_SA   0x1   /* expand sa[] (every 32nd) into a full suffix array on the device: 4 bytes per row, 5 when seq_len >= 2^32 */
int b200_index_upload(b200_ctx_t *ctx, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac, int flags);
/* same, from flat arrays (no reference structs needed) */
int b200_index_upload_raw(b200_ctx_t *ctx, const uint32_t *bwt, uint64_t bwt_size, uint64_t primary,
                          const uint64_t L2[5], uint64_t seq_len, const uint64_t *sa, uint64_t n_sa, int sa_intv,
                          const uint8_t *pac, int64_t l_pac, int n_seqs, const int64_t *ann_offset,
                          const int32_t *ann_len, const int32_t *ann_is_alt, int flags);

/* One index, several contexts.  dst on src's device: dst BORROWS src's index arrays (no copy; src must outlive dst) -- two
 * contexts per device let two host threads of b200_reads_to_sam keep one GPU busy.  dst on another device: the arrays are
 * copied device to device (peer copy; NVLink where present) instead of being uploaded or rebuilt: the fan-out of one index
 * to the GPUs of a box (replaces the per-process transferIndex of cuda/streams.cu:8-104). */
int b200_index_replicate(b200_ctx_t *dst, b200_ctx_t *src);

/* On-disk device-ready index (csrc/index_io.cu): everything the context's index holds in HBM -- the reference-layout arrays
 * of bwa_idx_load (bwa.c:384-434) AND the derived occ32 sectors and k-mer interval table, with_full_sa != 0 also the full
 * suffix array (8 bytes per row on 64-bit indexes: 50 GB at 3.1 Gbp; without it b200_index_load with B200_IDX_FULL_SA
 * re-expands the sampled SA on the device) -- written as it lies, and read back with large copies instead of the re-layout
 * of every start.  The file is specific to this library version (magic "B200IDX6"). */
int b200_index_save(b200_ctx_t *ctx, const char *path, int with_full_sa);
int b200_index_load(b200_ctx_t *ctx, const char *path, int flags);

/* Host-side index construction for in-memory genomes: the same arrays `bwa index` writes
 * (bwa_idx_build, bwtindex.c:253-323: pac, BWT+Occ blocks, SA sampled every 32 rows), built with an
 * own SA-IS suffix sorter.  fwd = nt4 codes 0..3 of the concatenated contigs.
 * info: [0]=primary [1..5]=L2 [6]=seq_len [7]=bwt_size [8]=n_sa [9]=sa_intv [10]=l_pac */
typedef struct b200_hostidx b200_hostidx_t;
b200_hostidx_t *b200_hostidx_build(const uint8_t *fwd, int64_t l_pac);
void            b200_hostidx_free(b200_hostidx_t *h);
void            b200_hostidx_info(const b200_hostidx_t *h, uint64_t *info /* 11 */);
const uint32_t *b200_hostidx_bwt(const b200_hostidx_t *h);
const uint64_t *b200_hostidx_sa(const b200_hostidx_t *h);
const uint8_t  *b200_hostidx_pac(const b200_hostidx_t *h);
int b200_index_upload_host(b200_ctx_t *ctx, const b200_hostidx_t *h, int n_seqs, const int64_t *ann_offset,
                           const int32_t *ann_len, const int32_t *ann_is_alt, int flags);

/* Device-side index construction (csrc/index_dev.cu): the arrays of bwa_idx_build (bwtindex.c:253-323; BWT of
 * forward + reverse complement bwtindex.c:63-126, Occ checkpoints :149-172, sampled SA bwt.c:62-84) computed on the
 * GPU from the 2-bit forward reference `pac` (bntseq.c:230 layout, l_pac/4+1 bytes) and left resident, ready for
 * alignment -- seconds instead of hours for a 3.1 Gbp text.  Suffixes are sorted by their first 32 bases in key-range
 * groups; suffixes colliding on all 32 bases are ordered on the host, at most "idx_tie_cap" (tunable) per group, else
 * the call fails (repeat-saturated texts: use b200_hostidx_build or `bwa index`).
 * b200_index_info: [0]=primary [1..5]=L2 [6]=seq_len [7]=bwt_size [8]=n_sa [9]=sa_intv [10]=l_pac
 *                  [11]=sort groups [12]=suffixes ordered on the host (0 for an uploaded index).
 * b200_index_download copies the reference-layout arrays (bwt_t.bwt: bwt_size words; bwt_t.sa: n_sa values) of the
 * resident index to the host, e.g. to fill a bwt_t for the host side of mem_process_seqs; either pointer may be NULL. */
int b200_index_build_device(b200_ctx_t *ctx, const uint8_t *pac, int64_t l_pac, int n_seqs, const int64_t *ann_offset,
                            const int32_t *ann_len, const int32_t *ann_is_alt, int flags);
int b200_index_info(b200_ctx_t *ctx, uint64_t *info /* 13 */);
int b200_index_download(b200_ctx_t *ctx, uint32_t *bwt, uint64_t *sa);

/* mem_reg2aln (bwamem.c:1099-1169; bwa_gen_cigar2 bwa.c:261-347, ksw_global2 ksw.c:504-606) for every region of a batch:
 * the fields of mem_aln_t that do not need floating point (mapq = mem_approx_mapq_se stays with the caller).
 * cigar words are bwa's (len<<4 | op, op in "MIDSH"); MD strings are not 0-terminated (l_md characters each). */
typedef struct {
	int64_t pos;                 /* 0-based on contig rid; -1 = unmapped */
	int32_t rid, flag, is_rev, is_alt;
	int32_t n_cigar, NM, score, sub, alt_sc, l_md;
	int64_t cigar_off, md_off;   /* into the cigar / MD pools of the call */
} b200_aln_t;
/* regs / cnt: the regions of each read in order (as b200_align_batch returns them, after the caller's primary marking).
 * Returns the number of regions; -2 when a pool is too small (totals[0..1] = CIGAR words / MD bytes needed). */
int64_t b200_reg2aln_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, const uint8_t *seqs, const int64_t *off,
                           const mem_alnreg_t *regs, const int32_t *cnt, b200_aln_t *out_aln, int64_t cap_cigar,
                           uint32_t *out_cigar, int64_t cap_md, char *out_md, int64_t totals[2]);

/* The single-end half of worker2 (bwamem.c:1195-1205) up to, not including, the SAM text: for every read
 * mem_mark_primary_se (bwamem.c:529-565, tie-break hash_64(n_processed + read index + i)), mem_reorder_primary5
 * (:988-1010, with MEM_F_PRIMARY5), then mem_reg2sam's walk (:1024-1042) -- which regions become SAM records, their
 * supplementary flag, MAPQ (mem_approx_mapq_se :962-986, capped by the first record's unless MEM_F_KEEP_SUPP_MAPQ) --
 * mem_gen_alt's choice of XA entries (bwamem_extra.c:116-170) and mem_reg2aln for every region that reaches the SAM.
 * regs is updated in place to the reference's order and fields (hash, sub, sub_n, alt_sc, secondary, secondary_all).
 * out_sel[g] / out_aln[g] describe region g of the updated array:
 *   kind   bit 0: a SAM record of its own, in this order; bit 1: an XA entry of the record of region xa_of of the
 *          same read, appended in this order; 0: the region does not reach the SAM (out_aln[g] is empty).
 *   mapq   final MAPQ of the record;  flag  0x800 / 0x10000 for supplementary records (also OR-ed into out_aln.flag).
 * out_aln[g].sub is -1 for secondary records as in mem_reg2sam.  A read none of whose regions has kind bit 0 gets the
 * unmapped record (mem_reg2aln with ar = 0: rid -1, pos -1, flag 0x4), which needs no device work: b200_sam_se_format
 * writes it.  Return value and pools as b200_reg2aln_batch; after -2 regs may already be updated, so call again with
 * the original regions (marking is not idempotent: sub_n accumulates, as in the reference). */
typedef struct { int32_t kind, mapq, xa_of, flag; } b200_sel_t;
int64_t b200_reg2sam_se_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, int64_t n_processed, const uint8_t *seqs,
                              const int64_t *off, mem_alnreg_t *regs, const int32_t *cnt, b200_sel_t *out_sel,
                              b200_aln_t *out_aln, int64_t cap_cigar, uint32_t *out_cigar, int64_t cap_md, char *out_md,
                              int64_t totals[2]);

/* Reads in host memory -> SAM text in host memory for single-end reads, as ONE call: the batch is cut into sub-batches
 * of sub_reads reads (0 = 2^20); each goes through b200_align_batch (worker1; only region counts come back),
 * b200_batch_reg2sam_se (worker2 on the resident regions) and b200_sam_se_format (n_threads host threads), and the text of
 * sub-batch k is written while the device works on sub-batch k+1 (two sets of page-locked record buffers owned by the
 * context).  Arguments as in the three calls; names / comments / sam_off are indexed by read.  Returns the number of
 * bytes written, -(needed) when cap is too small (nothing useful in sam), B200_E_FAIL / B200_E_CAPACITY on error. */
int64_t b200_align_sam_se_batch(b200_ctx_t *ctx, const mem_opt_t *opt, const char *const *ann_names, const char *const *ann_annos,
                                int n_reads, int64_t n_processed, const char *const *names, const char *const *comments,
                                const uint8_t *seqs, const uint8_t *quals, const int64_t *off, const char *rg_id, int n_threads,
                                int sub_reads, int64_t cap, char *sam, int64_t *sam_off);
/* Reads in host memory -> SAM text in host memory on one or several devices, as ONE call, with NO host formatting:
 * the per-GPU shard driver (csrc/reads2sam.cu).  The read set is cut into sub-batches of sub_reads reads (0 = about 2^20,
 * their number a multiple of n_ctx), handed to the contexts' host threads from one counter; each sub-batch runs worker1 and
 * worker2 (bwamem.c:1183-1213) and the SAM formatter (mem_aln2sam, bwamem.c:832-956; csrc/samdev.cu) on its device, and its
 * text is copied straight to its place in `sam` -- input order, offsets fixed by a running prefix over the sub-batches'
 * byte counts.  Every context must hold the same index (replicated FM-index); nothing moves between devices.
 * reads: flat fields as a FASTQ reader produces them (b200_fastq_read); names / comments are not 0-terminated.
 * `sam` should be page-locked (b200_host_alloc) for the copies to overlap the kernels.  sam_off (may be NULL): n_reads + 1
 * offsets of the reads' records.  stats (may be NULL, 8 doubles): [0] wall seconds, [1..3] summed over the device threads:
 * worker1, worker2 + formatter, waiting for placement; [4] sub-batches.
 * Returns the bytes written, -(needed) when cap is too small, B200_E_FAIL / B200_E_CAPACITY on error.  Single-end reads;
 * pes0 is reserved for the paired-end variant (b200_pairs_to_sam). */
typedef struct {
	int64_t n_reads;
	const uint8_t *seq; const int64_t *seq_off;          /* bases (letters or nt4 codes), n_reads + 1 offsets */
	const uint8_t *qual;                                  /* same offsets; NULL = no qualities ('*') */
	const char *name; const int64_t *name_off;            /* n_reads + 1 offsets */
	const char *comment; const int64_t *comment_off;      /* NULL = none; an empty slice = no comment for that read */
} b200_reads_t;
int64_t b200_reads_to_sam(b200_ctx_t *const *ctxs, int n_ctx, const mem_opt_t *opt, const char *const *ann_names,
                          const char *const *ann_annos, const b200_reads_t *reads, int64_t n_processed,
                          const mem_pestat_t *pes0, const char *rg_id, int sub_reads, int64_t cap, char *sam,
                          int64_t *sam_off, double *stats);

/* Input side: FASTQ / FASTA files (plain, gzip, or BGZF = gzip by block) -> batches of reads in the flat layout above, in
 * page-locked memory, parsed by n_threads threads (csrc/fastq.cpp).  Same record grammar as kseq.h, same name trimming
 * (trim_readno, bwa.c:64-68) and the same batch boundaries as bseq_read (bwa.c:191-224): records in file order -- mates
 * alternating when two files are given -- until the batch holds chunk_bases bases on an even record count.
 * b200_fastq_read fills *out with pointers into buffers owned by the reader, valid until the next call or the close;
 * returns the number of reads, 0 at the end of the input, < 0 on error.  Sequences are the file's letters (the aligner
 * converts), qualities as in the file or NULL, comments NULL when no record of the batch has one. */
typedef struct b200_fastq b200_fastq_t;
b200_fastq_t *b200_fastq_open(const char *fn1, const char *fn2 /* NULL: one file */, int n_threads);
int64_t b200_fastq_read(b200_fastq_t *f, int64_t chunk_bases, b200_reads_t *out);
void b200_fastq_close(b200_fastq_t *f);

/* FASTQ files -> SAM text on a file descriptor, the whole data path of `bwa mem` behind the option parsing: a reader thread
 * (b200_fastq_read, batches of chunk_bases bases = the batches bseq_read would make, so mem_pestat sees the reference's
 * batches), the devices (b200_reads_to_sam) and a writer thread (write(2) on out_fd; < 0 = discard) run concurrently, as the
 * three steps of upstream's kt_pipeline do (fastmap.c:305).  copy_comment = the -C option.  Returns the number of reads, or
 * B200_E_FAIL.  stats (may be NULL, 8 doubles): [0] wall seconds, [1] reading / parsing, [2] aligning, [3] writing,
 * [4] batches, [5] SAM bytes. */
int64_t b200_fastq_to_sam(b200_ctx_t *const *ctxs, int n_ctx, const mem_opt_t *opt, const char *const *ann_names,
                          const char *const *ann_annos, const char *fn1, const char *fn2, int64_t chunk_bases,
                          const mem_pestat_t *pes0, const char *rg_id, int n_threads, int copy_comment, int out_fd, double *stats);

/* Page-locked host memory (cudaHostAlloc) for callers that want the copies of the batched calls at full speed. */
void *b200_host_alloc(size_t bytes);
void b200_host_free(void *p);

/* The same on the batch the context has just aligned (b200_align_batch, or b200_batch_upload + b200_batch_run): reads
 * and regions are taken where they lie in device memory -- nothing is uploaded, and the regions need not have been
 * downloaded (b200_align_batch with cap 0 returns -(count) but still fills out_n).  cnt: the per-read region counts
 * that call returned; regs_out (may be NULL): the regions as updated.  The resident regions themselves are left as
 * worker1 produced them, so the call may be repeated (e.g. with larger pools after -2). */
int64_t b200_batch_reg2sam_se(b200_ctx_t *ctx, const mem_opt_t *opt, int64_t n_processed, const int32_t *cnt,
                              mem_alnreg_t *regs_out, b200_sel_t *out_sel, b200_aln_t *out_aln, int64_t cap_cigar,
                              uint32_t *out_cigar, int64_t cap_md, char *out_md, int64_t totals[2]);

/* SAM text of single-end reads from what b200_reg2sam_se_batch returned: mem_aln2sam (bwamem.c:832-956) with add_cigar
 * (:819-830; hard clips for supplementary records unless MEM_F_SOFTCLIP), the unmapped record of mem_reg2sam
 * (:1043-1047), SA and pa tags, the XA / XB strings of mem_gen_alt (bwamem_extra.c:143-161), RG (bwa_rg_id), the FASTQ
 * comment and XR (MEM_F_REF_HDR).  Host code (string formatting by n_threads threads over blocks of reads).
 * ann_names[rid] = bntann1_t.name, ann_annos (may be NULL) = bntann1_t.anno; names[r] = bseq1_t.name; comments (may be
 * NULL, entries may be NULL); seqs as given to the aligner (letters or nt4 codes), quals (may be NULL) with the same
 * offsets.  Returns the number of bytes written to sam (no terminating 0), or -(needed) when cap is too small;
 * sam_off (may be NULL): n_reads + 1 offsets of the reads' records. */
int64_t b200_sam_se_format(const mem_opt_t *opt, const char *const *ann_names, const char *const *ann_annos, int n_reads,
                           const char *const *names, const char *const *comments, const uint8_t *seqs, const uint8_t *quals,
                           const int64_t *off, const int32_t *cnt, const b200_sel_t *sel, const b200_aln_t *aln,
                           const uint32_t *cigar, const char *md, const char *rg_id, int n_threads, int64_t cap, char *sam,
                           int64_t *sam_off);

/* =====================================================================================
 * Batched device entry points (arrays of problems; host buffers in, host buffers out).
 * All return 0 on success, <0 on error (message in b200_last_error()); a capacity overflow
 * returns -(needed) where stated.
 * ===================================================================================== */

/* bwt_occ4 (bwt.c:169) / bwt_2occ4 (bwt.c:189) / bwt_extend (bwt.c:262) / bwt_sa (bwt.c:86) */
int b200_occ4_batch(b200_ctx_t *ctx, int64_t n, const uint64_t *k, uint64_t *cnt /* n x 4 */);
int b200_2occ4_batch(b200_ctx_t *ctx, int64_t n, const uint64_t *k, const uint64_t *l,
                     uint64_t *cntk /* n x 4 */, uint64_t *cntl /* n x 4 */);
int b200_extend_batch(b200_ctx_t *ctx, int64_t n, const uint64_t *ik /* n x 3 */, int is_back,
                      uint64_t *ok /* n x 4 x 3 */);
int b200_sa_batch(b200_ctx_t *ctx, int64_t n, const uint64_t *k, uint64_t *sa, int use_full_sa);

/* bwt_smem1a (bwt.c:289): problem p = (read[p], x[p], min_intv[p]); out_intv is cap entries per problem,
 * sorted by start like the reference; out_ret[p] = the function's return value. */
int b200_smem1a_batch(b200_ctx_t *ctx, int64_t n, const uint8_t *seqs, const int64_t *off, const int32_t *x,
                      const int32_t *min_intv, uint64_t max_intv, int cap, bwtintv_t *out_intv,
                      int32_t *out_n, int32_t *out_ret);

/* mem_collect_intv (bwamem.c:137): all three seeding passes + sort, one result list per read.
 * out_intv: cap_total entries, read r's list at [out_off[r], out_off[r]+out_n[r]). */
int64_t b200_collect_intv_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, const uint8_t *seqs,
                                const int64_t *off, int64_t cap_total, bwtintv_t *out_intv, int32_t *out_n);

/* mem_chain (bwamem.c:258) [stage 0] + mem_chain_flt (bwamem.c:334) [stage 1]
 * out_chn: 8 x int64 per chain: pos, n, rid, w, kept, first, is_alt, frac_rep(float bits)
 * out_seed: 4 x int64 per seed: rbeg, qbeg, len, score.  totals[0..1] = chains, seeds written. */
int b200_chain_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int stage, int n_reads, const uint8_t *seqs,
                     const int64_t *off, int64_t cap_chn, int64_t *out_chn, int32_t *out_nchn,
                     int64_t cap_seed, int64_t *out_seed, int64_t *totals);

/* ksw_extend2 (ksw.c:380) over an array of problems; out: 6 x int32 per problem:
 * score, qle, tle, gtle, gscore, max_off */
int b200_ksw_extend2_batch(b200_ctx_t *ctx, int64_t n, const int32_t *qlen, const int64_t *qoff, const uint8_t *q,
                           const int32_t *tlen, const int64_t *toff, const uint8_t *t, const int8_t *mat,
                           int o_del, int e_del, int o_ins, int e_ins, const int32_t *w, int end_bonus,
                           int zdrop, const int32_t *h0, int32_t *out);

/* ksw_extend2 over an array of problems for callers that may not allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * Same arithmetic and the same 6 x int32 results per problem as b200_ksw_extend2_batch (a variant of its
 * kernel), but: every array argument except `mat` (25 host bytes, copied into the kernel parameters) is DEVICE memory
 * of `device`, owned by the caller; the warp-private H / E rows and the work counter live in one file-scope __device__
 * array of b200_noalloc_scratch_bytes() = 19 398 912 bytes (18.5 MiB) that every GPU loading the library holds until the
 * process exits; the call makes no cudaMalloc / cudaFree / allocating library call and no synchronisation: it enqueues one
 * 256-byte memset (the work counter) and one kernel on `stream` (a cudaStream_t, NULL = the default stream) and returns.
 * Limits: max_qlen (the longest query of the call, stated by the caller) <= B200_NOALLOC_MAX_QLEN and
 * n <= B200_NOALLOC_MAX_PROBLEMS, else -2 and a message in b200_last_error(); a problem whose own qlen exceeds the limit or
 * whose h0 <= 0 (ksw.c:385) gets INT32_MIN in all six outputs.  All calls share the one scratch array: calls that overlap
 * on the GPU (different streams or host threads) race; serialise them. */
#define B200_NOALLOC_MAX_QLEN     1022
#define B200_NOALLOC_MAX_PROBLEMS ((int64_t)1 << 30)
int b200_ksw_extend2_noalloc(int device, int64_t n, int max_qlen, const int32_t *d_qlen, const int64_t *d_qoff,
                             const uint8_t *d_q, const int32_t *d_tlen, const int64_t *d_toff, const uint8_t *d_t,
                             const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins, const int32_t *d_w,
                             int end_bonus, int zdrop, const int32_t *d_h0, int32_t *d_out, void *stream);
int64_t b200_noalloc_scratch_bytes(void);
/* optional set-up (once per device, e.g. before a graph capture starts): looks up the scratch array's device address;
 * b200_ksw_extend2_noalloc otherwise does it on its first call for that device.  Also no counterpart in the reference. */
int b200_noalloc_setup(int device);

/* ksw_align2 (ksw.c:343) over an array of problems (striped u8 / i16 local SW incl. the KSW_XSTART reverse pass);
 * xtra per problem as in ksw.h:33-36; out: 7 x int32 per problem = kswr_t {score, te, qe, score2, te2, tb, qb} */
int b200_ksw_align2_batch(b200_ctx_t *ctx, int64_t n, const int32_t *qlen, const int64_t *qoff, const uint8_t *q,
                          const int32_t *tlen, const int64_t *toff, const uint8_t *t, const int8_t *mat,
                          int o_del, int e_del, int o_ins, int e_ins, const int32_t *xtra, int32_t *out);

/* mem_align1_core (bwamem.c:1061) for a batch = worker1 of mem_process_seqs (bwamem.c:1183):
 * seeding -> chaining -> filtering -> extension -> mem_sort_dedup_patch.  Reads may be ASCII or nt4.
 * out_regs holds cap records; read r's regions are the next out_n[r] records in read order.
 * Returns the total number of regions, or -(needed) if cap is too small (out_n is filled either way), or
 * B200_E_FAIL / B200_E_CAPACITY on error (test with B200_IS_ERROR before negating). */
int64_t b200_align_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, const uint8_t *seqs,
                         const int64_t *off, int64_t cap, mem_alnreg_t *out_regs, int32_t *out_n);

/* Device-resident variant used for kernel-only timing: inputs are uploaded once with
 * b200_batch_upload(), b200_batch_run() launches the whole stage pipeline without host copies and
 * b200_batch_download() fetches the regions.  stage_ms (may be NULL) receives CUDA-event times:
 * [0]=pack [1]=smem [2]=seed/sa [3]=chain [4]=extend [5]=total. */
/* Frees the device memory of the batch pipeline held by the context (resident batch, stage and worker2 / SAM buffers);
 * the index stays, buffers grow again on the next call.  0 on success. */
int b200_ctx_trim(b200_ctx_t *ctx);
int b200_batch_upload(b200_ctx_t *ctx, int n_reads, const uint8_t *seqs, const int64_t *off);
int b200_batch_run(b200_ctx_t *ctx, const mem_opt_t *opt, float *stage_ms);
int64_t b200_batch_download(b200_ctx_t *ctx, int64_t cap, mem_alnreg_t *out_regs, int32_t *out_n);
/* Paired-end stage on the resident batch (reads interleaved, mates at 2i / 2i+1): the mate-rescue block of
 * mem_sam_pe (bwamem_pair.c:291-302 around mem_matesw, :137-206, ksw_align2 + mem_sort_dedup_patch) for every pair,
 * given the batch's insert-size statistics pes[4] (mem_pestat, bwamem_pair.c:72, computed by the host program).
 * The resident region lists are replaced by the rescued ones (fetch with b200_batch_download); the host then calls
 * mem_sam_pe with MEM_F_NO_RESCUE set.  stats (may be NULL): [0]=ksw_align2 calls [1]=their DP cells; ms (may be
 * NULL): device time.  b200_batch_set_regs() loads region lists produced elsewhere into the resident batch. */
int b200_batch_rescue(b200_ctx_t *ctx, const mem_opt_t *opt, const mem_pestat_t *pes, int64_t *stats, float *ms);
int b200_batch_set_regs(b200_ctx_t *ctx, const mem_alnreg_t *regs, const int32_t *cnt);
/* work counters of the last b200_batch_run (for the roofline arithmetic):
 * [0]=bwt_extend calls [1]=of which touched two Occ blocks [2]=SA lookups [3]=LF steps walked
 * [4]=ksw_extend2 calls [5]=DP cells (sum over rows of end-beg) [6]=seeds [7]=chains kept [8]=regions
 * [9]=kernel launches */
int b200_batch_counters(b200_ctx_t *ctx, uint64_t *out /* 16 */);

/* Roofline denominators measured on this device: kind 0 = dependent random 32-byte sector
 * gathers over a table of table_bytes (`mlp` independent chains per lane), kind 1 = DPX instruction issue rate.
 * out[0] = milliseconds, out[1] = sectors gathered / DPX instructions executed. */
int b200_microbench(b200_ctx_t *ctx, int kind, int64_t table_bytes, int blocks_per_sm, int iters, int mlp, double *out);

/* =====================================================================================
 * Drop-in scalar entry points: the reference's own names and signatures.  Each call runs the
 * device code on a batch of one (parity tests call these one problem at a time).  The device
 * context/index is created lazily from the bwt_t identity on first use.
 * ===================================================================================== */
/* mem_process_seqs (bwamem.h:69, bwamem.c:1215) with everything on the device(s): worker1, the batch's mem_pestat (a host
 * reduction over device-made per-pair summaries), mate rescue, worker2 (mem_mark_primary_se, mem_pair, the MAPQ arithmetic and
 * flags of mem_sam_pe, mem_reg2aln) and the SAM text, through b200_reads_to_sam; no function of the host program is needed
 * (RG tag: b200_process_seqs_set_rg_id, else the host program's global bwa_rg_id if it is a dynamic symbol).
 * b200_process_seqs_host_worker2(1) selects the older route instead: worker1 and mate rescue on the device, then the
 * REFERENCE's own host code for mem_pestat / worker2 (mem_mark_primary_se, mem_reg2sam, mem_sam_pe), resolved from the program
 * this library is linked into (a cross-check).  seqs[i].sam is heap memory owned by the caller, as in the reference.
 * b200_process_seqs_set_ctx() selects an existing context (index already uploaded); otherwise one is created on
 * device 0 from (bwt, bns, pac) at the first call. */
void b200_process_seqs_set_ctx(b200_ctx_t *ctx);
/* Several devices in one host process: the batch is cut into contiguous blocks of reads (whole pairs), one per
 * context, each driven by its own host thread; results are put back in input order.  Either hand over contexts that
 * already hold the index, or ask for contexts on devices 0..n-1 to be created at the first call. */
void b200_process_seqs_set_ctxs(int n, b200_ctx_t **ctxs);
void b200_process_seqs_use_devices(int n_devices);
void b200_process_seqs_set_rg_id(const char *rg_id);   /* the host program's bwa_rg_id (bwa.c:44, set by -R) for the RG:Z tag; the pointer is kept */
void b200_process_seqs_host_worker2(int on);   /* 0 / -1 (default): worker2 and SAM on the device; 1: the host program's worker2 */
void b200_mem_process_seqs(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac,
                           int64_t n_processed, int n, bseq1_t *seqs, const mem_pestat_t *pes0);

#ifndef BWAMEM_B200_NO_DROPIN
void     mem_process_seqs(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac,
                          int64_t n_processed, int n, bseq1_t *seqs, const mem_pestat_t *pes0);             /* bwamem.h:69 */
void     bwt_occ4(const bwt_t *bwt, bwtint_t k, bwtint_t cnt[4]);                                   /* bwt.h:103 */
void     bwt_2occ4(const bwt_t *bwt, bwtint_t k, bwtint_t l, bwtint_t cntk[4], bwtint_t cntl[4]);  /* bwt.h:109 */
void     bwt_extend(const bwt_t *bwt, const bwtintv_t *ik, bwtintv_t ok[4], int is_back);           /* bwt.h:117 */
bwtint_t bwt_sa(const bwt_t *bwt, bwtint_t k);                                                      /* bwt.h:104 */
int      bwt_smem1a(const bwt_t *bwt, int len, const uint8_t *q, int x, int min_intv, uint64_t max_intv,
                    bwtintv_v *mem, bwtintv_v *tmpvec[2]);                                          /* bwt.h:124 */
int      ksw_extend2(int qlen, const uint8_t *query, int tlen, const uint8_t *target, int m, const int8_t *mat,
                     int o_del, int e_del, int o_ins, int e_ins, int w, int end_bonus, int zdrop, int h0,
                     int *qle, int *tle, int *gtle, int *gscore, int *max_off);                     /* ksw.h:108 */
kswr_t   ksw_align2(int qlen, uint8_t *query, int tlen, uint8_t *target, int m, const int8_t *mat, int o_del, int e_del,
                    int o_ins, int e_ins, int xtra, kswq_t **qry);                                   /* ksw.h:64 */
mem_alnreg_v mem_align1_core(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac,
                             int l_seq, char *seq, void *buf);                                      /* bwamem.c:1061 */
#endif

#ifdef __cplusplus
}
#endif
#endif
