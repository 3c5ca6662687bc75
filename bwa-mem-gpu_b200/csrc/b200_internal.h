// Internal declarations shared by the kernels and the host-side pipeline.
#pragma once
#include "simt.h"
#include "fmindex.cuh"
#include "../../include/bwamem_b200.h"

#define B200_CHECK(call)                                                                         \
    do {                                                                                         \
        cudaError_t e_ = (call);                                                                 \
        if (e_ != cudaSuccess) {                                                                 \
            b200_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return -1;                                                                           \
        }                                                                                        \
    } while (0)

void b200_set_error(const char *fmt, ...);

// ---- device-side records -------------------------------------------------------------------
struct Seed {            // one SA hit of one interval (mem_seed_t, bwa.h:121, without the fork's padding)
    i64 rbeg;            // K2a stores the SA row here, K2b replaces it by the text position
    i32 qbeg, len;
    i32 rid;             // bns_intv2rid result; < 0 = discarded (bwamem.c:294)
    i32 score;
};

struct ChainRec {        // one kept chain after mem_chain_flt
    i64 pos;
    i64 seed_off;        // into the chain-seed arena
    i32 n, rid, w, kept, first, is_alt;
    float frac_rep;
    i32 read;
};

enum {                   // work / traffic counters (u64 each)
    CNT_EXTEND = 0, CNT_TWO_BLOCK, CNT_SA, CNT_LF, CNT_KSW, CNT_CELLS, CNT_SEEDS, CNT_CHAINS, CNT_REGS,
    CNT_LAUNCH, CNT_SEEDSW, CNT_SEEDSW_CELLS, CNT_BIN0, CNT_BIN1, CNT_BIN2, CNT_DP8, CNT_N = 16
};

// ---- stage parameter blocks (passed by value to the kernels) ---------------------------------
struct SmemParams {
    DevIndex ix;
    const u8 *seq; const i64 *off; int n_reads;
    int min_seed_len, split_len, split_width; u64 max_mem_intv;
    Intv *out; int cap_out; i32 *out_n;          // cap_out slots per read
    Intv *lists; int list_cap;                   // 2 * list_cap per resident thread
    unsigned *work_counter;
    unsigned long long *stats; int *overflow;
    // single-call mode (b200_smem1a_batch): per problem x / min_intv, no length filter
    const i32 *px, *pmin_intv; u64 max_intv; i32 *out_ret;
    // fast kernel (smem2.cu): reads whose change-point list exceeds the shared-memory capacity are queued
    // here and re-done by the general kernel, which then takes its read indices from the queue
    unsigned *retry_n; i32 *retry_ids; int use_retry;
    // smem3.cu: 4-bit packed reads (pk_stride words per read) and the per-thread spill area (spill_cap entries)
    const u32 *pk; int pk_stride; int spill_cap;
    // third seeding pass (bwamem.c:167-182) in its own kernel: its intervals go to the END of the read's output slots
    // (out3_n of them, growing downwards); the main kernel then skips the pass and merges them before its final sort
    i32 *out3_n; int split_p3;
    int count_stats;                             // smem3.cu: count extensions / two-block extensions into stats[] (tunable seed_stats)
};

struct SeedParams {
    DevIndex ix;
    const i64 *off; int n_reads;
    const Intv *intv; int cap_intv; const i32 *intv_n;
    int max_occ;
    i32 *seed_n; const i64 *seed_off;            // per read
    Seed *seeds; i64 n_seeds;
    i32 *l_rep;                                  // per read (bwamem.c:272-279)
    unsigned long long *stats;
};

struct ChainParams {
    DevIndex ix;
    mem_opt_t opt;
    const i64 *off; int n_reads;
    const Seed *seeds; const i64 *seed_off; const i32 *seed_n; const i32 *l_rep;
    int *scratch;                                // 16 ints per seed + 128 per read
    int stage;                                   // 0: mem_chain only, 1: + mem_chain_flt
    // pass A: count; pass B: emit
    i32 *chn_n, *cseed_n;                        // per read: kept chains, their seeds
    const i64 *chn_off, *cseed_off;
    ChainRec *chains; Seed *cseeds;
    int *overflow;
};

struct SeedSwParams {                            // mem_flt_chained_seeds over the long reads of a chunk
    DevIndex ix;
    mem_opt_t opt;
    const u8 *seq; const i64 *off; int n_reads;
    ChainRec *chains; const i64 *chn_off; const i32 *chn_n; Seed *cseeds;
    const i32 *min_hsp;                          // per read: min_HSP_score (bwamem.c:608)
    const i32 *long_reads; int n_long;           // reads that pass the length test (bwamem.c:609)
    int slen_cap;
    unsigned *work_counter; u64 *stats; int *overflow;
};

struct ExtParams {
    DevIndex ix;
    mem_opt_t opt;
    const u8 *seq; const i64 *off; int n_reads;
    const ChainRec *chains; const i64 *chn_off; const i32 *chn_n;
    const Seed *cseeds;
    mem_alnreg_t *regs; int cap_regs;            // staging: cap_regs per resident warp
    i32 *eh; int eh_cap;                         // per resident warp: 2 * eh_cap ints (H,E rows)
    u64 *srt; int srt_cap;                       // per resident warp: seed sort keys
    mem_alnreg_t *out; i64 out_cap; unsigned long long *out_used;
    i64 *reg_off; i32 *reg_n;                    // per read
    unsigned *work_counter;
    unsigned long long *stats; int *overflow;
};

// ---- round-based extension (ext2.cu)
struct ExtJob {          // one seed extension: both sides + band retries (bwamem.c:722-778); 96 bytes
    i64 rbeg, rmax0, rmax1;          // seed start on the reference, chain window
    i32 read, qbeg, slen, lq;        // read, seed on the query, query length
    i32 wl0, wl1, wr0, wr1;          // bands of the two tries per side, already clamped like ksw.c:399-407
    i64 rb, re;                      // results
    i32 qb, qe, score, truesc, w, key;   // key: (left flank / 8) << 5 | (right flank / 8), the order jobs are handed to the DP lanes in
};
struct ExtRead { i32 ci, k, nav, pad; i64 rmax0, rmax1; };   // per-read cursor between rounds
struct Ext2Params {
    DevIndex ix;
    mem_opt_t opt;
    const u8 *seq; const i64 *off; int n_reads;
    const ChainRec *chains; const i64 *chn_off; const i32 *chn_n;
    const Seed *cseeds;
    mem_alnreg_t *regs;              // per read at chn_off[r], capacity = the read's seed count
    u64 *srt;                        // per chain at its seed_off
    ExtRead *rstate;
    // Jobs are binned by the longer of their two query flanks so that the DP kernel of a bin sizes its shared-memory
    // rows for that bin only (short flanks = many resident warps).  Bins 0 and 2 share one array (bin 0 grows up from
    // slot 0, bin 2 down from the last slot), bin 1 has its own; both double-buffered by round parity.
    ExtJob *jobs[2], *jobs_mid[2];
    int jobs_cap;                    // slots per array (>= reads in the chunk: at most one job per read and round)
    int bin_max[2];                  // longest flank of bin 0 / bin 1
    unsigned *njobs;                 // per round: 8 counters: [0..2] jobs emitted per bin, [4..6] jobs taken by the DP kernel
    u32 *perm; u32 *sort_cnt;        // optional: per bin (b * jobs_cap) the job slots ordered by key, longest flanks first; 3 x 1024 counters
    i32 *eh; int eh_cap;             // finish kernel scratch (mem_patch_reg's global alignment)
    i64 *reg_off; i32 *reg_n;
    unsigned *work_counter;
    unsigned long long *stats;
    ExtJob *tail_jobs;               // tail mode (ext3.cu): one job slot per resident DP lane
};

struct KswBatchParams {
    i64 n;
    const i32 *qlen; const i64 *qoff; const u8 *q;
    const i32 *tlen; const i64 *toff; const u8 *t;
    int8_t mat[25];
    int o_del, e_del, o_ins, e_ins; const i32 *w; int end_bonus, zdrop; const i32 *h0;
    i32 *out; i32 *eh; int eh_cap;
    unsigned *work_counter;
};

// ---- stage launchers (each returns 0 / -1) ----------------------------------------------------
int launch_pack_reads(u8 *seq, i64 n_bytes, cudaStream_t st);
int launch_smem_collect(const SmemParams &p, int grid, int block, cudaStream_t st);
int launch_smem_single(const SmemParams &p, int grid, int block, cudaStream_t st);
int launch_smem_fast(const SmemParams &p, int grid, int wide, cudaStream_t st);
int launch_pack4(u8 *seq, const i64 *off, int n_reads, u32 *pk, int stride, cudaStream_t st);
size_t smem_fast3_bytes(int wide, int lcap, int pk_stride);
int launch_smem_fast3(const SmemParams &p, int grid, int wide, int lcap, int qglob, cudaStream_t st);
int launch_smem_p3(const SmemParams &p, int n_sms, int wide, int qglob, cudaStream_t st);
int launch_build_occ32(const DevIndex &ix, uint4 *out, u64 n_blocks, cudaStream_t st);
int launch_build_kmer_tab(DevIndex &ix, uint4 *tab, int kmax, int wide, cudaStream_t st);
size_t kmer_tab_entries(int kmax);
int launch_seed_count(const SeedParams &p, cudaStream_t st);
int launch_seed_fill(const SeedParams &p, cudaStream_t st);
int launch_sa_locate(const SeedParams &p, cudaStream_t st);
int launch_chain(const ChainParams &p, int emit, cudaStream_t st);
int launch_extend(const ExtParams &p, int grid, int block, cudaStream_t st);
int launch_ext2_plan(const Ext2Params &p, int round, int n_threads, cudaStream_t st);
int launch_ext2_dp(const Ext2Params &p, int round, int bin, int he_cols, int grid, int rowsync, int phased, cudaStream_t st);
size_t ext2_dp8_smem(int n_groups);
int launch_ext2_dp8(const Ext2Params &p, int round, int bin, int n_groups, int grid, int rowsync, int phased, cudaStream_t st);
int launch_ext2_dp8_tail(const Ext2Params &p, int round, int n_groups, int grid, cudaStream_t st);
int launch_ext2_sort(const Ext2Params &p, int round, int n_jobs, int n_sms, cudaStream_t st);
int launch_ext2_finish(const Ext2Params &p, int grid, cudaStream_t st);
int launch_ksw_batch(const KswBatchParams &p, int grid, int block, cudaStream_t st);
int launch_gather_regs(const mem_alnreg_t *src, const i64 *src_off, const i32 *n, const i64 *dst_off,
                       int n_reads, mem_alnreg_t *dst, cudaStream_t st);
int launch_seed_sw(const SeedSwParams &P, int n_sms, cudaStream_t st);
struct SswJob;
bool seed_sw2_supported(const mem_opt_t *opt);
int launch_seed_sw2(const SeedSwParams &P, SswJob *jobs, i64 n_seeds_total, i32 *score_out, unsigned *n_jobs, int n_sms, cudaStream_t st);
int launch_exclusive_scan(const i32 *in, i64 *out /* n+1 */, int n, i64 *tmp, cudaStream_t st);
int launch_occ_batch(const DevIndex &ix, i64 n, const u64 *k, const u64 *l, u64 *ck, u64 *cl, cudaStream_t st);
int launch_extend_batch(const DevIndex &ix, i64 n, const u64 *ik, int is_back, u64 *ok, cudaStream_t st);
int launch_sa_batch(const DevIndex &ix, i64 n, const u64 *k, u64 *out, int use_full, cudaStream_t st);
int launch_build_full_sa(const DevIndex &ix, u32 *out32, u64 *out64, cudaStream_t st);
size_t sa40_bytes(u64 n_rows);
int launch_pack_sa40(const void *in, int in_is64, u32 *out, u64 n_rows, cudaStream_t st);
