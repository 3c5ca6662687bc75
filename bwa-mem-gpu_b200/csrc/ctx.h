// Host-side context: device index + grow-only device buffers for the batch pipeline.
#pragma once
#include "b200_internal.h"
#include <vector>
#include <mutex>

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
    int ensure(size_t need)            // grow-only; contents are NOT preserved
    {
        if (need <= bytes) return 0;
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
        size_t want = need + need / 8 + 256;
        if (cudaMalloc(&p, want) != cudaSuccess) { p = nullptr; b200_set_error("cudaMalloc(%zu) failed", want); return -1; }
        bytes = want;
        return 0;
    }
    int ensure_exact(size_t need)      // the same without the growth slack (index arrays: tens of GB, sized once)
    {
        if (need <= bytes) return 0;
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
        if (cudaMalloc(&p, need + 256) != cudaSuccess) { p = nullptr; b200_set_error("cudaMalloc(%zu) failed", need + 256); return -1; }
        bytes = need + 256;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
    template <class T> T *as() const { return (T *)p; }
};

struct b200_ctx {
    std::mutex call_mu;                        // serialises the same-name scalar shims (dropin.cu): they share this context's scratch and stream
    int device = 0;
    int n_sms = 148;
    cudaStream_t st = nullptr;
    cudaStream_t st_h2d = nullptr, st_d2h = nullptr;    // copy streams of b200_align_batch (created on first use)
    bool has_index = false;
    DevIndex ix;
    DevBuf d_occ32, d_kmer, d_retry, d_pk, d_spill, d_jobs0, d_jobs1, d_jobs2, d_jobs3, d_jsort, d_rstate, d_njobs, d_bwt, d_sa, d_sa_full, d_pac, d_ann_off, d_ann_len, d_ann_alt;
    const void *bwt_identity = nullptr;       // host bwt_t* this index was uploaded from (drop-in registry)
    bool stub_pac = false;             // index uploaded by a bwt-only primitive (no bns / pac yet)

    // resident input batch
    DevBuf d_seq, d_off;
    int n_reads = 0; i64 n_bytes = 0; int max_len = 0;
    std::vector<i64> h_off;
    const i64 *hoff = nullptr;                 // host view of the batch's read offsets (h_off, or the caller's array during b200_align_batch)

    // tunables
    int chunk_reads = 1 << 20; bool chunk_user = false;   // reads per pipeline chunk (tunable; unset: scaled down for long reads, pipeline.cu)
    int cap_intv = 0;                          // 0 = auto from read length
    int cap_retry = 1;                         // a chunk that overflows the interval slots is re-run with the reported need (pipeline.cu)
    int need_cap_intv = 0;                     // set by a chunk that overflowed
    int ext_path = 0;                          // 0 = auto, 1 = force the general warp-per-read kernel, 2 = force the round-based fast path
    int smem_path = 0;                         // 0 = auto (smem3 for short reads), 2 = smem2 kernel, 1 = general kernel only
    int l2_persist = 0;                        // 1: pin the seeding index in persisting L2 (set before the index upload)
    size_t l2_setaside = 0, l2_window = 0;
    int ext_bin0 = 0, ext_bin1 = 0;            // > 0: flank-length limits of extension bins 0 / 1 (default 32 / 64)
    int ext_bps = 0;                           // > 0: cap on resident blocks per SM of the extension DP kernel
    int ssw_path = 0;                          // mem_seed_sw: 0 = one alignment per lane (seedsw2.cu), 1 = lane-group restatement of the striped code (seedsw.cu)
    int ext_phased = 1;                        // byte-score extension kernel: warp-wide phases (left sides, retries, right sides, next jobs)
    int ext_sort = 1;                          // byte-score extension kernel: jobs handed out ordered by flank shape (ext3.cu)
    int seed_stats = 0;                        // seeding kernels count bwt_extend calls / two-block calls (counters extend, two_block); off = leaner kernel
    int ext_tail = 16384;                      // byte-score extension: a round with at most this many jobs is finished, with all later rounds, by ONE launch (0 = off)
    int ext_rowsync = 1;                       // byte-score extension kernel: a lane finishes its row before the warp's control slot (ext3.cu)
    int split_p3 = 1;                          // third seeding pass in its own kernel (smem_p3_kernel) ahead of the main one
    int force_wide = 0;                        // tests: run the 64-bit-position templates even when seq_len < 2^32 (1: packed 33-bit records, 2: two-quad records); set before the index upload
    int kmer_k = -1;                           // longest string in the k-mer interval table (0 = none, -1 = auto: 9, or 13 when seq_len >= 2^32:
                                               // the upper levels of a deep table only pay when the index itself is DRAM-priced); set before the index upload
    int smem_bps = 0;                          // > 0: cap on resident blocks per SM of the seeding kernel (experiments)
    int smem_lcap = 0;                         // 0 = auto: shared-memory list entries per read in the fast SMEM kernel
    i64 idx_group_cap = 0, idx_tie_cap = 0;    // device index builder (index_dev.cu): suffixes per sort group / colliding suffixes per group (0 = default)
    u64 idx_bwt_size = 0;                      // words of the reference-layout BWT held in d_bwt
    int idx_groups = 0; i64 idx_ties = 0;      // what the last device index build did

    // per-chunk intermediates
    DevBuf d_intv, d_intv_n, d_lists, d_seed_n, d_seed_off, d_l_rep, d_seeds, d_cseeds, d_chains, d_chn_n,
           d_cseed_n, d_scratch, d_stage, d_eh, d_srt, d_arena, d_reg_off, d_reg_n, d_dst_off, d_scan_tmp,
           d_ctl;                              // counters / flags / stats block
    // whole-batch result
    DevBuf d_final, d_final_n;
    i64 total_regs = 0;
    unsigned long long stats[CNT_N] = {0};
    DevBuf d_sam[28];                          // device SAM formatter / reads -> SAM driver (reads2sam.cu)
    cudaEvent_t sam_ev[2] = { nullptr, nullptr };  // D2H copies of the two text buffers
    DevBuf d_r2a[18];                          // b200_reg2aln_batch / b200_reg2sam_se_batch (reg2aln.cu)
    int r2a_warp = 1;                          // reg2aln: regions with a query part > 400 bases go to the warp-per-region kernel (tunable)
    int r2a_small = 1;                         // reg2aln pass 1 with ring-sized H / E rows (tunable)
    int log_cap = 0;                           // entries of the log(i) table in d_r2a[16]
    void *pin_p[16] = {}; size_t pin_cap[16] = {};   // page-locked host buffers of b200_align_sam_se_batch (b200_ctx_pinned)
    // scratch for the primitive batches
    DevBuf d_pa, d_pb, d_pc, d_pd, d_pe, d_pf, d_pg, d_ph;
};

// Position width of the seeding kernels (smem3.cu): 0 = u32 (seq_len < 2^32), 2 = u64 with packed 16-byte records
// (seq_len < 2^33: the 3.1 Gbp configs), 1 = u64 with two-quad records (anything larger)
static inline int b200_wide_mode(const b200_ctx *ctx)
{
    const u64 n = ctx->ix.seq_len;
    if (n < 0xfffffff0ull && !ctx->force_wide) return 0;
    if (ctx->force_wide == 2 || n >= 0x1fffffff0ull) return 1;
    return 2;
}

// layout of d_ctl (all 8-byte slots)
enum { CTL_WORK = 0, CTL_OVF = 1, CTL_OVF2 = 2, CTL_ARENA_USED = 3, CTL_TOTAL = 4, CTL_RETRY = 5, CTL_WORK2 = 6, CTL_STATS = 8, CTL_SLOTS = 8 + CNT_N };

// worker2 on the resident batch with the results LEFT on the device: `fn` is called once per chunk of whole reads while
// the chunk's records are valid (they are overwritten by the next chunk); reg2aln.cu
struct R2AChunk {
    int r_lo, r_hi; i64 g0; int n;              // reads r_lo .. r_hi of the batch = regions g0 .. g0 + n
    const mem_alnreg_t *regs; const b200_aln_t *aln; const b200_sel_t *sel; const u32 *cig; const char *md; i64 cig_base, md_base;
    const i64 *d_roff;                          // n_reads + 1 region offsets of the batch, on the device
    const i32 *mate_reg, *read_flag;            // paired-end (else null): per read of the batch, see samdev.cuh
};
typedef int (*r2a_chunk_fn)(void *user, const R2AChunk &c);
int64_t b200_batch_reg2sam_dev(b200_ctx *ctx, const mem_opt_t *opt, i64 n_processed, const i32 *cnt, r2a_chunk_fn fn, void *user,
                               const mem_pestat_t *pes = nullptr);

// per-pair summary for mem_pestat (finalize_pe.cuh): out[p] = 0, or (insert size << 2 | orientation) + 1
int launch_pestat_pairs(const mem_opt_t *opt, i64 l_pac, const mem_alnreg_t *regs, const i64 *roff, int n_pairs, u64 *out, cudaStream_t st);

int b200_run_chunk(b200_ctx *ctx, const mem_opt_t *opt, int r0, int n, int stop_after, float *stage_ms);
int b200_index_finish(b200_ctx *ctx, int flags);
int b200_pack_full_sa(b200_ctx *ctx);              // ctx.cu: 8-byte (or, with force_wide, 4-byte) full SA -> 5 bytes per row
b200_ctx *b200_ctx_for_bwt(const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac);

extern "C" void *b200_ctx_pinned(b200_ctx_t *ctx, int slot, size_t bytes);   // ctx.cu: grow-only page-locked buffer of the context (slots 0-15)
