// Batch pipeline (the new "batched read scheduler"): reads resident on the device are pushed through
// pack -> SMEM -> seeds/SA -> chain/filter -> extend/dedup in chunks that reuse grow-only intermediate
// buffers; results are compacted in read order.  Replaces mem_align_GPU / miniBatchMain
// (cuda/bwamem_GPU.cu:3457, cuda/minibatch_process.cpp:188) and worker1 of mem_process_seqs
// (bwamem.c:1183-1195, 1232).
#include <math.h>
#include "ctx.h"
#include <string.h>
#include <stdlib.h>
#include <stdio.h>
#include <algorithm>

static inline unsigned *ctl_u32(b200_ctx *c, int slot) { return (unsigned *)(c->d_ctl.as<u64>() + slot); }
static inline unsigned long long *ctl_u64(b200_ctx *c, int slot) { return (unsigned long long *)(c->d_ctl.as<u64>() + slot); }

static int ctl_zero(b200_ctx *c, int slot, int n = 1)
{
    B200_CHECK(cudaMemsetAsync(c->d_ctl.as<u64>() + slot, 0, 8 * (size_t)n, c->st));
    return 0;
}
static int ctl_read(b200_ctx *c, int slot, int n, u64 *out)
{
    B200_CHECK(cudaMemcpyAsync(out, c->d_ctl.as<u64>() + slot, 8 * (size_t)n, cudaMemcpyDeviceToHost, c->st));
    B200_CHECK(cudaStreamSynchronize(c->st));
    return 0;
}

extern "C" int b200_batch_upload(b200_ctx_t *ctx, int n_reads, const uint8_t *seqs, const int64_t *off)
{
    if (!ctx) { b200_set_error("null context"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    const i64 base = n_reads ? off[0] : 0, nbytes = n_reads ? off[n_reads] - base : 0;
    ctx->h_off.resize((size_t)n_reads + 1);
    int mx = 0;
    if (n_reads && base == 0) memcpy(ctx->h_off.data(), off, ((size_t)n_reads + 1) * 8);
    else for (int i = 0; i <= n_reads; ++i) ctx->h_off[i] = n_reads ? off[i] - base : 0;
    {
        i64 m = 0;
        const i64 *h = ctx->h_off.data();
        for (int i = 0; i < n_reads; ++i) { const i64 d = h[i + 1] - h[i]; m = d > m ? d : m; }
        mx = (int)m;
    }
    if (ctx->d_seq.ensure((size_t)nbytes + 64) || ctx->d_off.ensure(((size_t)n_reads + 1) * 8)) return -1;
    if (nbytes) B200_CHECK(cudaMemcpyAsync(ctx->d_seq.p, seqs + base, (size_t)nbytes, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_off.p, ctx->h_off.data(), ((size_t)n_reads + 1) * 8, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    ctx->n_reads = n_reads; ctx->n_bytes = nbytes; ctx->max_len = mx;
    ctx->total_regs = 0;
    ctx->hoff = ctx->h_off.data();
    return 0;
}

static int auto_cap_intv(const b200_ctx *ctx, const mem_opt_t *opt)
{
    if (ctx->cap_intv > 0) return ctx->cap_intv;
    // SMEMs of length >= min_seed_len: pass 1 can emit at most len - min_seed_len + 1, in practice a handful;
    // size for len/4 + 16 (>= 53 for 150 bp) and report an overflow instead of truncating silently.
    int c = ctx->max_len / 4 + 16;
    return c < 32 ? 32 : c;
}

struct StageTimer {
    cudaEvent_t ev[8]; int n = 0; bool on = false;
    void init(bool enable) { on = enable; if (on) for (auto &e : ev) cudaEventCreate(&e); }
    void mark(cudaStream_t st) { if (on && n < 8) cudaEventRecord(ev[n++], st); }
    void fini() { if (on) for (auto &e : ev) cudaEventDestroy(e); }
};

// stop_after: 1 = SMEM intervals, 2 = seeds, 3 = chains (opt stage in chain_stage), 5 = everything
static int g_chain_stage = 1;
#define SSW_MAX_Q 200                                    // mem_seed_sw only aligns windows shorter than MEM_SHORT_LEN

static int run_chunk_once(b200_ctx *ctx, const mem_opt_t *opt, int r0, int n, int stop_after, float *stage_ms);

// A chunk whose reads need more interval slots than the heuristic (or the tunable) gave them is run again with the
// count the device reported: the buffers are grow-only and nothing of a chunk is committed before it has succeeded
// (the reference has no such limit: bwtintv_v grows by realloc, bwt.c:306).  Tunable cap_retry=0 turns that off.
int b200_run_chunk(b200_ctx *ctx, const mem_opt_t *opt, int r0, int n, int stop_after, float *stage_ms)
{
    for (int attempt = 0; ; ++attempt) {
        ctx->need_cap_intv = 0;
        float ms[6] = { 0, 0, 0, 0, 0, 0 };
        const int rc = run_chunk_once(ctx, opt, r0, n, stop_after, stage_ms ? ms : nullptr);
        if (rc == -2 && ctx->need_cap_intv > 0 && ctx->cap_retry && attempt < 4) {
            ctx->cap_intv = ctx->need_cap_intv + 8;
            continue;
        }
        if (stage_ms) for (int i = 0; i < 6; ++i) stage_ms[i] += ms[i];
        return rc;
    }
}

static int run_chunk_once(b200_ctx *ctx, const mem_opt_t *opt, int r0, int n, int stop_after, float *stage_ms)
{
    cudaStream_t st = ctx->st;
    const i64 *d_off = ctx->d_off.as<i64>() + r0;
    if (!ctx->hoff) { b200_set_error("no resident batch (b200_batch_upload first)"); return -1; }
    const i64 byte0 = ctx->hoff[r0], nbytes = ctx->hoff[r0 + n] - byte0;
    (void)nbytes;
    StageTimer tm; tm.init(stage_ms != nullptr);
    if (ctx->d_ctl.ensure(CTL_SLOTS * 8)) return -1;
    if (ctl_zero(ctx, 0, CTL_SLOTS)) return -1;
    u64 launches = 0;

    tm.mark(st);
    // ---- K1 setup: which seeding kernel
    const int cap_intv = auto_cap_intv(ctx, opt);
    const int wmode = b200_wide_mode(ctx);                                    // 0 u32, 1 u64 two-quad records, 2 u64 packed records
    const bool wide = wmode != 0;
    const int pk_stride = ((ctx->max_len + 7) / 8 + 3) & ~3;                  // words per packed read, 16-byte multiple
    int lcap3 = ctx->smem_lcap > 0 ? ctx->smem_lcap : 8;      // 8: less shared memory = more L1 for the hot table levels (profiles/r4b_seed_ab.txt)
    lcap3 = lcap3 <= 8 ? 8 : 16;
    // reads too long to stage in shared memory: smem_fast_kernel (smem2.cu) by default -- measured faster on 1-10 kbp reads
    // (575 vs 758 ms per 200 k reads) -- or, with smem_path 3, the fast3 kernels fetching the read's 4-bit codes from global memory
    const bool fits3 = smem_fast3_bytes(wmode, lcap3, pk_stride) <= 100 * 1024;
    const bool qglob = ctx->smem_path == 3;
    if (qglob) lcap3 = 16;
    const bool use3 = (ctx->smem_path == 0 && fits3) || qglob;
    // ---- K0: ASCII -> nt4 (idempotent on already-packed input) [+ 4-bit packed copy for the seeding kernel]
    if (use3) {
        if (ctx->d_pk.ensure((size_t)n * pk_stride * 4 + 64)) return -1;
        if (launch_pack4(ctx->d_seq.as<u8>(), d_off, n, ctx->d_pk.as<u32>(), pk_stride, st)) return -1;
    } else if (launch_pack_reads(ctx->d_seq.as<u8>() + byte0, ctx->hoff[r0 + n] - byte0, st)) return -1;
    ++launches;
    tm.mark(st);

    // ---- K1: SMEM collection: fast kernel (shared-memory lists, occ32 index), then the general kernel on the
    //      (normally empty) queue of reads the fast kernel could not hold
    int blocks = std::min((n + 127) / 128, ctx->n_sms);
    const int list_cap = ctx->max_len + 1;
    {   // keep the general kernel's interval-list scratch under ~2 GB for long reads
        size_t per_block = (size_t)128 * 2 * list_cap * sizeof(Intv);
        size_t max_blocks = std::max<size_t>(1, ((size_t)2 << 30) / per_block);
        blocks = (int)std::min<size_t>(blocks, max_blocks);
    }
    if (ctx->smem_path == 1) blocks = std::min((n + 127) / 128, ctx->n_sms * 8);
    if (ctx->d_intv.ensure((size_t)n * cap_intv * sizeof(Intv)) || ctx->d_intv_n.ensure((size_t)n * 4) ||
        ctx->d_lists.ensure((size_t)blocks * 128 * 2 * list_cap * sizeof(Intv)) || ctx->d_retry.ensure((size_t)n * 4 + 4)) return -1;
    SmemParams SP;
    memset(&SP, 0, sizeof(SP));
    SP.ix = ctx->ix; SP.seq = ctx->d_seq.as<u8>(); SP.off = d_off; SP.n_reads = n;
    SP.min_seed_len = opt->min_seed_len;
    SP.split_len = (int)(opt->min_seed_len * opt->split_factor + .499);      // bwamem.c:141
    SP.split_width = opt->split_width; SP.max_mem_intv = opt->max_mem_intv;
    SP.out = ctx->d_intv.as<Intv>(); SP.cap_out = cap_intv; SP.out_n = ctx->d_intv_n.as<i32>();
    SP.work_counter = ctl_u32(ctx, CTL_WORK); SP.stats = ctl_u64(ctx, CTL_STATS); SP.overflow = (int *)ctl_u32(ctx, CTL_OVF);
    SP.retry_n = ctl_u32(ctx, CTL_RETRY); SP.retry_ids = ctx->d_retry.as<i32>();
    if (use3) {
        const size_t smem = smem_fast3_bytes(wmode, lcap3, qglob ? 0 : pk_stride);
        int per_sm = std::max(1, std::min(8, (int)((224 * 1024) / (smem + 1024))));
        if (ctx->smem_bps > 0) per_sm = std::min(per_sm, ctx->smem_bps);
        int fast_blocks = std::max(1, std::min((n + 127) / 128, ctx->n_sms * per_sm));
        const int spill_cap = std::max(1, ctx->max_len + 1 - lcap3);
        {   // keep the per-lane spill areas of long reads under ~8 GB
            const size_t per_block = (size_t)128 * spill_cap * (wmode == 1 ? 32 : 16);
            fast_blocks = (int)std::max<size_t>(1, std::min<size_t>(fast_blocks, ((size_t)8 << 30) / per_block));
        }
        if (ctx->d_spill.ensure((size_t)fast_blocks * 128 * spill_cap * (wmode == 1 ? 32 : 16))) return -1;
        SP.pk = ctx->d_pk.as<u32>(); SP.pk_stride = pk_stride; SP.spill_cap = spill_cap; SP.count_stats = ctx->seed_stats;
        SP.lists = (Intv *)ctx->d_spill.p;
        if (ctx->split_p3 && opt->max_mem_intv > 0) {                        // pass 3 first: uniform forward-only lanes
            SP.split_p3 = 1; SP.out3_n = ctx->d_retry.as<i32>();
            SP.work_counter = ctl_u32(ctx, CTL_WORK2);
            if (launch_smem_p3(SP, ctx->n_sms, wmode, qglob ? 1 : 0, st)) return -1;
            SP.work_counter = ctl_u32(ctx, CTL_WORK);
            ++launches;
        }
        if (launch_smem_fast3(SP, fast_blocks, wmode, lcap3, qglob ? 1 : 0, st)) return -1;
        ++launches;
    } else if (ctx->smem_path != 1) {
        const int lcap = ctx->smem_lcap > 0 ? ctx->smem_lcap : (wide ? 24 : 20);      // 64-bit indexes: ~17 interval-size changes per forward sweep at 3.1 Gbp; 16 sent every read to the general kernel (profiles/r4i_long_ab.jsonl)
        const int per_sm = std::max(1, (int)((200 * 1024) / ((size_t)lcap * 128 * (wide ? 32 : 16))));
        const int fast_blocks = std::max(1, std::min((n + 127) / 128, ctx->n_sms * std::min(per_sm, 6)));
        SP.list_cap = lcap;
        if (launch_smem_fast(SP, fast_blocks, wide ? 1 : 0, st)) return -1;
        ++launches;
    }
    if (!use3) {                                                             // general kernel: retry queue, or everything
        SP.lists = ctx->d_lists.as<Intv>(); SP.list_cap = list_cap; SP.use_retry = ctx->smem_path != 1;
        SP.work_counter = ctl_u32(ctx, CTL_WORK2);
        if (launch_smem_collect(SP, blocks, 128, st)) return -1;
        ++launches;
    }
    tm.mark(st);
    if (stop_after == 1) {
        u64 ovf;
        if (ctl_read(ctx, CTL_OVF, 1, &ovf)) return -1;
        if ((int)(ovf & 0xffffffff)) { ctx->need_cap_intv = (int)(ovf & 0xffffffff); tm.fini(); b200_set_error("interval capacity %d too small (need %d)", cap_intv, (int)(ovf & 0xffffffff)); return -2; }
        if (ovf >> 32) { b200_set_error("interval list scratch overflow"); return -2; }
        tm.fini();
        return 0;
    }

    // ---- K2: seeds + SA location
    if (ctx->d_seed_n.ensure((size_t)n * 4) || ctx->d_seed_off.ensure(((size_t)n + 1) * 8) || ctx->d_l_rep.ensure((size_t)n * 4) ||
        ctx->d_scan_tmp.ensure(((size_t)n / 1024 + 2) * 8)) return -1;
    SeedParams DP;
    memset(&DP, 0, sizeof(DP));
    DP.ix = ctx->ix; DP.off = d_off; DP.n_reads = n;
    DP.intv = ctx->d_intv.as<Intv>(); DP.cap_intv = cap_intv; DP.intv_n = ctx->d_intv_n.as<i32>();
    DP.max_occ = opt->max_occ;
    DP.seed_n = ctx->d_seed_n.as<i32>(); DP.seed_off = ctx->d_seed_off.as<i64>(); DP.l_rep = ctx->d_l_rep.as<i32>();
    DP.stats = ctl_u64(ctx, CTL_STATS);
    if (launch_seed_count(DP, st)) return -1;
    if (launch_exclusive_scan(DP.seed_n, ctx->d_seed_off.as<i64>(), n, ctx->d_scan_tmp.as<i64>(), st)) return -1;
    launches += 4;
    i64 tot_seeds = 0;
    u64 ovf = 0;
    B200_CHECK(cudaMemcpyAsync(&tot_seeds, ctx->d_seed_off.as<i64>() + n, 8, cudaMemcpyDeviceToHost, st));
    if (ctl_read(ctx, CTL_OVF, 1, &ovf)) return -1;                         // also synchronises
    if ((int)(ovf & 0xffffffff)) { ctx->need_cap_intv = (int)(ovf & 0xffffffff); tm.fini(); b200_set_error("interval capacity %d too small (need %d)", cap_intv, (int)(ovf & 0xffffffff)); return -2; }
    if (ovf >> 32) { b200_set_error("interval list scratch overflow"); return -2; }
    if (ctx->d_seeds.ensure((size_t)(tot_seeds + 1) * sizeof(Seed))) return -1;
    DP.seeds = ctx->d_seeds.as<Seed>(); DP.n_seeds = tot_seeds;
    if (launch_seed_fill(DP, st)) return -1;
    if (launch_sa_locate(DP, st)) return -1;
    launches += 2;
    ctx->stats[CNT_SEEDS] += tot_seeds;
    ctx->stats[CNT_SA] += tot_seeds;
    tm.mark(st);
    if (stop_after == 2) { B200_CHECK(cudaStreamSynchronize(st)); tm.fini(); return 0; }

    // ---- K3: chaining + filtering
    if (ctx->d_cseeds.ensure((size_t)(tot_seeds + 1) * sizeof(Seed)) || ctx->d_chains.ensure((size_t)(tot_seeds + 1) * sizeof(ChainRec)) ||
        ctx->d_chn_n.ensure((size_t)n * 4) || ctx->d_cseed_n.ensure((size_t)n * 4) ||
        ctx->d_scratch.ensure(((size_t)16 * tot_seeds + (size_t)128 * n + 64) * 4)) return -1;
    ChainParams CP;
    memset(&CP, 0, sizeof(CP));
    CP.ix = ctx->ix; CP.opt = *opt; CP.off = d_off; CP.n_reads = n;
    CP.seeds = ctx->d_seeds.as<Seed>(); CP.seed_off = ctx->d_seed_off.as<i64>(); CP.seed_n = ctx->d_seed_n.as<i32>();
    CP.l_rep = ctx->d_l_rep.as<i32>(); CP.scratch = ctx->d_scratch.as<int>(); CP.stage = stop_after == 3 ? g_chain_stage : 1;
    CP.chn_n = ctx->d_chn_n.as<i32>(); CP.cseed_n = ctx->d_cseed_n.as<i32>();
    CP.chn_off = ctx->d_seed_off.as<i64>(); CP.cseed_off = ctx->d_seed_off.as<i64>();
    CP.chains = ctx->d_chains.as<ChainRec>(); CP.cseeds = ctx->d_cseeds.as<Seed>();
    CP.overflow = (int *)ctl_u32(ctx, CTL_OVF2);
    if (launch_chain(CP, 1, st)) return -1;
    ++launches;
    // ---- K3b: mem_flt_chained_seeds (bwamem.c:605-622) for the reads long enough to need it
    if (stop_after != 3 || g_chain_stage >= 2) {
        // min_l and min_HSP_score depend on the read length only; log() is evaluated on the host like the reference does
        std::vector<i32> by_len((size_t)ctx->max_len + 1, -2), mh, lr;
        auto hsp_of = [&](int l_query) -> int {                          // -1: too short for the filter (bwamem.c:609)
            const double min_l = opt->min_chain_weight ? 1.1f * opt->min_chain_weight : 5.5f * log((double)l_query);
            return (l_query == 0 || min_l > 0.05f * l_query) ? -1 : (int)(opt->a * min_l + .499);
        };
        // min_l - 0.05 l is concave in l (a constant or 5.5 ln l, minus a line): when it is positive at both ends of
        // [min_seed_len, max_len] it is positive for every read that can carry a seed (bwamem.c:267), i.e. no read of the
        // chunk reaches the filter, and the per-read pass below (10^6 host iterations per chunk) is skipped
        const int l_lo = std::max(1, std::min(opt->min_seed_len, ctx->max_len));
        const bool none_long = ctx->max_len < 1 || (hsp_of(ctx->max_len) < 0 && hsp_of(l_lo) < 0);
        if (!none_long) mh.resize((size_t)n);
        for (int r = 0; r < n && !none_long; ++r) {
            const int l_query = (int)(ctx->hoff[r0 + r + 1] - ctx->hoff[r0 + r]);
            if (by_len[l_query] == -2) by_len[l_query] = hsp_of(l_query);
            mh[r] = by_len[l_query];
            if (by_len[l_query] >= 0) lr.push_back(r);
        }
        if (!lr.empty()) {
            if (ctx->d_pa.ensure((size_t)n * 4) || ctx->d_pb.ensure(lr.size() * 4)) return -1;
            B200_CHECK(cudaMemcpyAsync(ctx->d_pa.p, mh.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
            B200_CHECK(cudaMemcpyAsync(ctx->d_pb.p, lr.data(), lr.size() * 4, cudaMemcpyHostToDevice, st));
            if (ctl_zero(ctx, CTL_WORK)) return -1;
            SeedSwParams SP;
            memset(&SP, 0, sizeof(SP));
            SP.ix = ctx->ix; SP.opt = *opt; SP.seq = ctx->d_seq.as<u8>(); SP.off = d_off; SP.n_reads = n;
            SP.chains = ctx->d_chains.as<ChainRec>(); SP.chn_off = ctx->d_seed_off.as<i64>(); SP.chn_n = ctx->d_chn_n.as<i32>();
            SP.cseeds = ctx->d_cseeds.as<Seed>(); SP.min_hsp = ctx->d_pa.as<i32>();
            SP.long_reads = ctx->d_pb.as<i32>(); SP.n_long = (int)lr.size();
            SP.slen_cap = (SSW_MAX_Q + 7) / 8 + 1;
            SP.work_counter = ctl_u32(ctx, CTL_WORK); SP.stats = (u64 *)ctl_u64(ctx, CTL_STATS) + CNT_SEEDSW; SP.overflow = (int *)ctl_u32(ctx, CTL_OVF2);
            if (ctx->ssw_path != 1 && seed_sw2_supported(opt)) {            // one Smith-Waterman per lane (seedsw2.cu)
                if (ctx->d_pc.ensure((size_t)(tot_seeds + 1) * 32) || ctx->d_pd.ensure((size_t)(tot_seeds + 1) * 4)) return -1;
                if (launch_seed_sw2(SP, (SswJob *)ctx->d_pc.p, tot_seeds, ctx->d_pd.as<i32>(), ctl_u32(ctx, 7), ctx->n_sms, st)) return -1;
                launches += 3;
            } else {                                                         // lane-group restatement of the striped code (seedsw.cu)
                if (launch_seed_sw(SP, ctx->n_sms, st)) return -1;
                ++launches;
            }
            B200_CHECK(cudaStreamSynchronize(st));                       // mh / lr are stack-owned host buffers
        }
    }
    tm.mark(st);
    if (stop_after == 3) {
        if (ctl_read(ctx, CTL_OVF2, 1, &ovf)) return -1;
        if (ovf) { b200_set_error("chain B-tree pool overflow"); return -2; }
        tm.fini();
        return 0;
    }

    // ---- K4: extension + dedup
    const i64 arena_cap = tot_seeds + 1;
    if (ctx->d_arena.ensure((size_t)arena_cap * sizeof(mem_alnreg_t)) || ctx->d_reg_off.ensure((size_t)n * 8) ||
        ctx->d_reg_n.ensure((size_t)n * 4) || ctx->d_dst_off.ensure(((size_t)n + 1) * 8)) return -1;
    // round-based fast path (ext2.cu) when scores fit 14 bits and one H/E row per lane fits shared memory
    const int he_cols = ctx->max_len + 2 + 8;                                 // widest bin; + 8: the DP hot slot prefetches 8 columns
    bool fast_ext = (i64)ctx->max_len * opt->a < 16384 && (size_t)he_cols * 64 * 4 <= 96 * 1024;
    for (int k = 0; k < 25; ++k) if (opt->mat[k] > opt->a) fast_ext = false;     // bound above assumes max(mat) == a
    if (ctx->ext_path == 1) fast_ext = false;
    // byte-score DP kernel: every H / E / F of a read is at most l_query * a (ksw.c:399: h0 + qlen * max(mat))
    const bool dp8 = fast_ext && ctx->ext_path != 2 && (i64)ctx->max_len * opt->a <= 255 && opt->a <= 63;
    i32 *reg_n_ptr = ctx->d_reg_n.as<i32>();
    i64 *reg_off_ptr = ctx->d_reg_off.as<i64>();
    if (fast_ext) {
        const int MAX_ROUNDS = 4096;
        const int fin_blocks = std::max(1, std::min((n + 127) / 128, ctx->n_sms * 2));
        const int eh_cap = ctx->max_len + 2;
        if (ctx->d_jobs0.ensure((size_t)n * sizeof(ExtJob)) || ctx->d_jobs1.ensure((size_t)n * sizeof(ExtJob)) ||
            ctx->d_jobs2.ensure((size_t)n * sizeof(ExtJob)) || ctx->d_jobs3.ensure((size_t)n * sizeof(ExtJob)) ||
            ctx->d_rstate.ensure((size_t)n * sizeof(ExtRead)) || ctx->d_njobs.ensure((size_t)MAX_ROUNDS * 32) ||
            ctx->d_jsort.ensure(((size_t)3 * n + 3 * 1024) * 4) ||
            ctx->d_srt.ensure((size_t)(tot_seeds + 1) * 8) || ctx->d_eh.ensure((size_t)fin_blocks * 128 * 2 * eh_cap * 4)) return -1;
        B200_CHECK(cudaMemsetAsync(ctx->d_njobs.p, 0, (size_t)MAX_ROUNDS * 32, st));
        Ext2Params XP;
        memset(&XP, 0, sizeof(XP));
        XP.ix = ctx->ix; XP.opt = *opt; XP.seq = ctx->d_seq.as<u8>(); XP.off = d_off; XP.n_reads = n;
        XP.chains = ctx->d_chains.as<ChainRec>(); XP.chn_off = ctx->d_seed_off.as<i64>(); XP.chn_n = ctx->d_chn_n.as<i32>();
        XP.cseeds = ctx->d_cseeds.as<Seed>();
        XP.regs = ctx->d_arena.as<mem_alnreg_t>(); XP.srt = ctx->d_srt.as<u64>();
        XP.rstate = ctx->d_rstate.as<ExtRead>();
        XP.jobs[0] = ctx->d_jobs0.as<ExtJob>(); XP.jobs[1] = ctx->d_jobs1.as<ExtJob>();
        XP.jobs_mid[0] = ctx->d_jobs2.as<ExtJob>(); XP.jobs_mid[1] = ctx->d_jobs3.as<ExtJob>();
        XP.jobs_cap = n;
        // flank-length bins: the DP kernel of a bin holds flank + 1 columns (+ 8 for the hot slot's prefetch) per lane
        const int bin_flank[3] = { ctx->ext_bin0 > 0 ? ctx->ext_bin0 : 32, ctx->ext_bin1 > 0 ? ctx->ext_bin1 : 64, ctx->max_len };
        XP.bin_max[0] = std::min(bin_flank[0], ctx->max_len); XP.bin_max[1] = std::min(std::max(bin_flank[1], XP.bin_max[0]), ctx->max_len);
        XP.njobs = ctx->d_njobs.as<unsigned>();
        if (ctx->ext_sort) { XP.perm = ctx->d_jsort.as<u32>(); XP.sort_cnt = XP.perm + (size_t)3 * n; }
        XP.eh = ctx->d_eh.as<i32>(); XP.eh_cap = eh_cap;
        XP.reg_off = reg_off_ptr; XP.reg_n = reg_n_ptr;
        XP.work_counter = ctl_u32(ctx, CTL_WORK); XP.stats = ctl_u64(ctx, CTL_STATS);
        int threads = n;
        for (int round = 0; ; ++round) {
            if (round >= MAX_ROUNDS) { b200_set_error("extension: more than %d rounds", MAX_ROUNDS); return -2; }
            if (launch_ext2_plan(XP, round, threads, st)) return -1;
            ++launches;
            unsigned nj[3] = { 0, 0, 0 };
            B200_CHECK(cudaMemcpyAsync(nj, XP.njobs + 8 * round, 12, cudaMemcpyDeviceToHost, st));
            B200_CHECK(cudaStreamSynchronize(st));
            if (nj[0] + nj[1] + nj[2] == 0) break;
            if (dp8 && ctx->ext_tail > 0 && nj[0] + nj[1] + nj[2] <= (unsigned)ctx->ext_tail) {      // few jobs left: one launch finishes the chunk (ext3.cu, TAIL)
                const int n_groups = (ctx->max_len >> 3) + 1;
                const int dp_per_sm = std::max(1, std::min(8, (int)((200 * 1024) / (ext2_dp8_smem(n_groups) + 1024))));
                const int grid = std::max(1, std::min((int)((nj[0] + nj[1] + nj[2] + 63) / 64), ctx->n_sms * dp_per_sm));
                if (ctx->d_pe.ensure((size_t)grid * 64 * sizeof(ExtJob))) return -1;
                XP.tail_jobs = (ExtJob *)ctx->d_pe.p;
                if (launch_ext2_dp8_tail(XP, round, n_groups, grid, st)) return -1;
                ++launches;
                ctx->stats[CNT_DP8] += nj[0] + nj[1] + nj[2];
                break;
            }
            if (XP.perm) {                                                   // hand jobs out by shape (ext3.cu)
                if (launch_ext2_sort(XP, round, (int)(nj[0] + nj[1] + nj[2]), ctx->n_sms, st)) return -1;
                launches += 3;
            }
            for (int b = 0; b < 3; ++b) {
                if (nj[b] == 0) continue;
                const int flank = b == 2 ? ctx->max_len : XP.bin_max[b];
                if (dp8) {                                                   // byte scores, two columns per instruction (ext3.cu)
                    const int n_groups = (flank >> 3) + 1;                   // columns 0 .. flank
                    const int dp_per_sm = std::max(1, std::min(ctx->ext_bps > 0 ? ctx->ext_bps : 16, (int)((200 * 1024) / (ext2_dp8_smem(n_groups) + 1024))));
                    const int grid = std::max(1, std::min((int)((nj[b] + 63) / 64), ctx->n_sms * dp_per_sm));
                    if (launch_ext2_dp8(XP, round, b, n_groups, grid, ctx->ext_rowsync, ctx->ext_phased, st)) return -1;
                    ++launches;
                    ctx->stats[CNT_BIN0 + b] += nj[b];
                    ctx->stats[CNT_DP8] += nj[b];
                    continue;
                }
                const int cols = flank + 2 + 8;
                const int dp_per_sm = std::max(1, std::min(ctx->ext_bps > 0 ? ctx->ext_bps : 16, (int)((200 * 1024) / ((size_t)cols * 64 * 4))));
                const int grid = std::max(1, std::min((int)((nj[b] + 63) / 64), ctx->n_sms * dp_per_sm));
                if (launch_ext2_dp(XP, round, b, cols, grid, ctx->ext_rowsync, ctx->ext_phased, st)) return -1;
                ++launches;
                ctx->stats[CNT_BIN0 + b] += nj[b];
            }
            threads = (int)(nj[0] + nj[1] + nj[2]);
        }
        if (ctl_zero(ctx, CTL_WORK)) return -1;
        if (launch_ext2_finish(XP, fin_blocks, st)) return -1;
        ++launches;
    } else {
    const int ext_blocks = std::max(1, std::min((n + 3) / 4, ctx->n_sms * 8));
    const int cap_stage = 512, srt_cap = 8192, eh_cap = ctx->max_len + 2;
    if (ctx->d_stage.ensure((size_t)ext_blocks * 4 * cap_stage * sizeof(mem_alnreg_t)) ||
        ctx->d_eh.ensure((size_t)ext_blocks * 4 * 2 * eh_cap * 4) || ctx->d_srt.ensure((size_t)ext_blocks * 4 * srt_cap * 8)) return -1;
    if (ctl_zero(ctx, CTL_WORK)) return -1;
    ExtParams EP;
    memset(&EP, 0, sizeof(EP));
    EP.ix = ctx->ix; EP.opt = *opt; EP.seq = ctx->d_seq.as<u8>(); EP.off = d_off; EP.n_reads = n;
    EP.chains = ctx->d_chains.as<ChainRec>(); EP.chn_off = ctx->d_seed_off.as<i64>(); EP.chn_n = ctx->d_chn_n.as<i32>();
    EP.cseeds = ctx->d_cseeds.as<Seed>();
    EP.regs = ctx->d_stage.as<mem_alnreg_t>(); EP.cap_regs = cap_stage;
    EP.eh = ctx->d_eh.as<i32>(); EP.eh_cap = eh_cap; EP.srt = ctx->d_srt.as<u64>(); EP.srt_cap = srt_cap;
    EP.out = ctx->d_arena.as<mem_alnreg_t>(); EP.out_cap = arena_cap; EP.out_used = ctl_u64(ctx, CTL_ARENA_USED);
    EP.reg_off = reg_off_ptr; EP.reg_n = reg_n_ptr;
    EP.work_counter = ctl_u32(ctx, CTL_WORK); EP.stats = ctl_u64(ctx, CTL_STATS); EP.overflow = (int *)ctl_u32(ctx, CTL_OVF2);
    if (launch_extend(EP, ext_blocks, 128, st)) return -1;
    ++launches;
    }
    // read-ordered compaction into the batch result
    if (launch_exclusive_scan(reg_n_ptr, ctx->d_dst_off.as<i64>(), n, ctx->d_scan_tmp.as<i64>(), st)) return -1;
    launches += 3;
    i64 tot_regs = 0;
    u64 ctl[CTL_SLOTS];
    B200_CHECK(cudaMemcpyAsync(&tot_regs, ctx->d_dst_off.as<i64>() + n, 8, cudaMemcpyDeviceToHost, st));
    if (ctl_read(ctx, 0, CTL_SLOTS, ctl)) return -1;
    if (ctl[CTL_OVF2]) { b200_set_error("extension stage overflow (code %d)", (int)ctl[CTL_OVF2]); return -2; }
    // grow the final buffer preserving earlier chunks
    {
        size_t need = (size_t)(ctx->total_regs + tot_regs + 1) * sizeof(mem_alnreg_t);
        if (need > ctx->d_final.bytes) {
            DevBuf nb;
            if (nb.ensure(need * 2)) return -1;
            if (ctx->total_regs) B200_CHECK(cudaMemcpyAsync(nb.p, ctx->d_final.p, (size_t)ctx->total_regs * sizeof(mem_alnreg_t), cudaMemcpyDeviceToDevice, st));
            B200_CHECK(cudaStreamSynchronize(st));
            if (ctx->st_d2h) B200_CHECK(cudaStreamSynchronize(ctx->st_d2h));   // b200_align_batch: earlier chunks may still be on their way out
            ctx->d_final.release();
            ctx->d_final = nb;
        }
    }
    if (launch_gather_regs(ctx->d_arena.as<mem_alnreg_t>(), reg_off_ptr, reg_n_ptr, ctx->d_dst_off.as<i64>(), n,
                           ctx->d_final.as<mem_alnreg_t>() + ctx->total_regs, st)) return -1;
    ++launches;
    B200_CHECK(cudaMemcpyAsync(ctx->d_final_n.as<i32>() + r0, reg_n_ptr, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    tm.mark(st);
    B200_CHECK(cudaStreamSynchronize(st));
    ctx->total_regs += tot_regs;
    for (int i = 0; i < CNT_N; ++i) ctx->stats[i] += ctl[CTL_STATS + i];
    ctx->stats[CNT_REGS] += tot_regs;
    ctx->stats[CNT_LAUNCH] += launches;
    if (stage_ms) {
        for (int i = 0; i + 1 < tm.n && i < 5; ++i) { float ms = 0; cudaEventElapsedTime(&ms, tm.ev[i], tm.ev[i + 1]); stage_ms[i] += ms; }
        float ms = 0; cudaEventElapsedTime(&ms, tm.ev[0], tm.ev[tm.n - 1]); stage_ms[5] += ms;
    }
    tm.fini();
    return 0;
}

// Reads per pipeline chunk: the tunable when it was set, else 2^20 reads but at most ~2^30 bases (long reads: the
// per-read interval / list buffers grow with the read length, while the lane-per-read kernels want >= 10^5 reads in flight)
static int eff_chunk(const b200_ctx *ctx)
{
    if (ctx->chunk_user) return ctx->chunk_reads;
    const i64 by_bases = ((i64)1 << 30) / std::max(1, ctx->max_len);
    return (int)std::max<i64>(4096, std::min<i64>(ctx->chunk_reads, by_bases));
}

extern "C" int b200_batch_run(b200_ctx_t *ctx, const mem_opt_t *opt, float *stage_ms)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    if (stage_ms) for (int i = 0; i < 6; ++i) stage_ms[i] = 0;
    memset(ctx->stats, 0, sizeof(ctx->stats));
    ctx->total_regs = 0;
    if (ctx->d_final_n.ensure((size_t)ctx->n_reads * 4 + 4)) return -1;
    cudaEvent_t ea = nullptr, eb = nullptr;               // device-side bracket of the whole batch on the launching stream
    if (stage_ms) { cudaEventCreate(&ea); cudaEventCreate(&eb); cudaEventRecord(ea, ctx->st); }
    const int chunk = eff_chunk(ctx);
    for (int r0 = 0; r0 < ctx->n_reads; r0 += chunk) {
        int n = std::min(chunk, ctx->n_reads - r0);
        int rc = b200_run_chunk(ctx, opt, r0, n, 5, stage_ms);
        if (rc) return rc;
    }
    if (stage_ms) {
        cudaEventRecord(eb, ctx->st);
        cudaEventSynchronize(eb);
        cudaEventElapsedTime(&stage_ms[5], ea, eb);
        cudaEventDestroy(ea); cudaEventDestroy(eb);
    }
    return 0;
}

extern "C" int64_t b200_batch_download(b200_ctx_t *ctx, int64_t cap, mem_alnreg_t *out_regs, int32_t *out_n)
{
    if (!ctx) { b200_set_error("null context"); return B200_E_FAIL; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) { b200_set_error("cudaSetDevice failed"); return B200_E_FAIL; }
#define DL_CHECK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { b200_set_error("%s -> %s", #call, cudaGetErrorString(e_)); return B200_E_FAIL; } } while (0)
    if (out_n && ctx->n_reads) DL_CHECK(cudaMemcpyAsync(out_n, ctx->d_final_n.p, (size_t)ctx->n_reads * 4, cudaMemcpyDeviceToHost, ctx->st));
    if (ctx->total_regs > cap) { cudaStreamSynchronize(ctx->st); return -ctx->total_regs; }
    if (ctx->total_regs) DL_CHECK(cudaMemcpyAsync(out_regs, ctx->d_final.p, (size_t)ctx->total_regs * sizeof(mem_alnreg_t), cudaMemcpyDeviceToHost, ctx->st));
    DL_CHECK(cudaStreamSynchronize(ctx->st));
#undef DL_CHECK
    return ctx->total_regs;
}

// Replace the resident region lists (e.g. with lists produced elsewhere) for the stages that run after worker1.
extern "C" int b200_batch_set_regs(b200_ctx_t *ctx, const mem_alnreg_t *regs, const int32_t *cnt)
{
    if (!ctx) { b200_set_error("null context"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    i64 tot = 0;
    for (int i = 0; i < ctx->n_reads; ++i) tot += cnt[i];
    if (ctx->d_final.ensure((size_t)(tot + 1) * sizeof(mem_alnreg_t)) || ctx->d_final_n.ensure((size_t)ctx->n_reads * 4 + 4)) return -1;
    if (tot) B200_CHECK(cudaMemcpyAsync(ctx->d_final.p, regs, (size_t)tot * sizeof(mem_alnreg_t), cudaMemcpyHostToDevice, ctx->st));
    if (ctx->n_reads) B200_CHECK(cudaMemcpyAsync(ctx->d_final_n.p, cnt, (size_t)ctx->n_reads * 4, cudaMemcpyHostToDevice, ctx->st));
    if (ctx->n_bytes && launch_pack_reads(ctx->d_seq.as<u8>(), ctx->n_bytes, ctx->st)) return -1;      // reads as nt4 (idempotent)
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    ctx->total_regs = tot;
    return 0;
}

extern "C" int b200_batch_counters(b200_ctx_t *ctx, uint64_t *out)
{
    if (!ctx) return -1;
    for (int i = 0; i < CNT_N; ++i) out[i] = ctx->stats[i];
    return 0;
}

// Host buffers in, host buffers out, with the copies overlapped with the kernels: the reads of chunk k+1.. travel on
// their own stream while chunk k computes (the compute stream waits on a per-chunk event), and each chunk's records go
// back on a third stream as soon as its read-ordered gather is done.  With pinned host buffers only the first chunk's
// upload and the last chunk's download are exposed; pageable buffers make the same calls behave synchronously.
static int align_batch_impl(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, const uint8_t *seqs,
                            const int64_t *off, int64_t cap, mem_alnreg_t *out_regs, int32_t *out_n)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    if (!ctx->st_h2d) {
        B200_CHECK(cudaStreamCreateWithFlags(&ctx->st_h2d, cudaStreamNonBlocking));
        B200_CHECK(cudaStreamCreateWithFlags(&ctx->st_d2h, cudaStreamNonBlocking));
    }
    const i64 base = n_reads ? off[0] : 0, nbytes = n_reads ? off[n_reads] - base : 0;
    if (n_reads && base == 0) { ctx->h_off.clear(); ctx->hoff = off; }                  // offsets read in place for the duration of the call
    else {
        ctx->h_off.resize((size_t)n_reads + 1);
        for (int i = 0; i <= n_reads; ++i) ctx->h_off[i] = n_reads ? off[i] - base : 0;
        ctx->hoff = ctx->h_off.data();
    }
    int mx = 0;
    {
        i64 m = 0;
        const i64 *h = ctx->hoff;
        for (int i = 0; i < n_reads; ++i) { const i64 d = h[i + 1] - h[i]; m = d > m ? d : m; }
        mx = (int)m;
    }
    if (ctx->d_seq.ensure((size_t)nbytes + 64) || ctx->d_off.ensure(((size_t)n_reads + 1) * 8) ||
        ctx->d_final_n.ensure((size_t)n_reads * 4 + 4)) return -1;
    ctx->n_reads = n_reads; ctx->n_bytes = nbytes; ctx->max_len = mx;
    ctx->total_regs = 0;
    memset(ctx->stats, 0, sizeof(ctx->stats));
    const int chunk = eff_chunk(ctx);
    const int n_chunks = n_reads ? (n_reads + chunk - 1) / chunk : 0;
    std::vector<cudaEvent_t> up((size_t)n_chunks, nullptr);
    int rc = 0;
    for (int k = 0; k < n_chunks && !rc; ++k) {                                        // all uploads queued up front, one event per chunk
        const int r0 = k * chunk, n = std::min(chunk, n_reads - r0);
        const i64 b0 = ctx->hoff[r0], b1 = ctx->hoff[r0 + n];
        if (cudaEventCreateWithFlags(&up[k], cudaEventDisableTiming) != cudaSuccess) { b200_set_error("cudaEventCreate failed"); rc = -1; break; }
        cudaMemcpyAsync(ctx->d_off.as<i64>() + r0, ctx->hoff + r0, ((size_t)n + 1) * 8, cudaMemcpyHostToDevice, ctx->st_h2d);
        if (b1 > b0 && cudaMemcpyAsync(ctx->d_seq.as<u8>() + b0, seqs + base + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, ctx->st_h2d) != cudaSuccess) {
            b200_set_error("read upload failed"); rc = -1; break;
        }
        cudaEventRecord(up[k], ctx->st_h2d);
    }
    i64 copied = 0;                                                                     // records already on their way to the host
    for (int k = 0; k < n_chunks && !rc; ++k) {
        const int r0 = k * chunk, n = std::min(chunk, n_reads - r0);
        cudaStreamWaitEvent(ctx->st, up[k], 0);
        const i64 before = ctx->total_regs;
        rc = b200_run_chunk(ctx, opt, r0, n, 5, nullptr);                               // returns with the chunk's gather finished
        if (rc) break;
        if (out_n) cudaMemcpyAsync(out_n + r0, ctx->d_final_n.as<i32>() + r0, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->st_d2h);
        if (ctx->total_regs <= cap && ctx->total_regs > before) {
            cudaMemcpyAsync(out_regs + before, ctx->d_final.as<mem_alnreg_t>() + before, (size_t)(ctx->total_regs - before) * sizeof(mem_alnreg_t),
                            cudaMemcpyDeviceToHost, ctx->st_d2h);
            copied = ctx->total_regs;
        }
    }
    cudaStreamSynchronize(ctx->st_h2d);
    cudaError_t e = cudaStreamSynchronize(ctx->st_d2h);
    cudaStreamSynchronize(ctx->st);
    for (cudaEvent_t ev : up) if (ev) cudaEventDestroy(ev);
    if (ctx->hoff == off) ctx->hoff = nullptr;                                          // the caller's array is not ours to keep
    if (rc) return rc;
    if (e != cudaSuccess) { b200_set_error("result download failed: %s", cudaGetErrorString(e)); return -1; }
    (void)copied;
    return 0;
}

extern "C" int64_t b200_align_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, const uint8_t *seqs,
                                    const int64_t *off, int64_t cap, mem_alnreg_t *out_regs, int32_t *out_n)
{
    const int rc = align_batch_impl(ctx, opt, n_reads, seqs, off, cap, out_regs, out_n);
    if (rc) return rc == -2 ? B200_E_CAPACITY : B200_E_FAIL;                            // never confused with -(regions needed)
    return ctx->total_regs > cap ? -ctx->total_regs : ctx->total_regs;
}

// ---- stage seams (parity tests) -----------------------------------------------------------------
extern "C" int64_t b200_collect_intv_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, const uint8_t *seqs,
                                           const int64_t *off, int64_t cap_total, bwtintv_t *out_intv, int32_t *out_n)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    if (b200_batch_upload(ctx, n_reads, seqs, off)) return -1;
    i64 tot = 0;
    std::vector<Intv> tmp;
    const int chunk = eff_chunk(ctx);
    for (int r0 = 0; r0 < n_reads; r0 += chunk) {
        int n = std::min(chunk, n_reads - r0);
        int rc = b200_run_chunk(ctx, opt, r0, n, 1, nullptr);
        if (rc) return rc;
        const int cap_intv = auto_cap_intv(ctx, opt);
        tmp.resize((size_t)n * cap_intv);
        B200_CHECK(cudaMemcpyAsync(out_n + r0, ctx->d_intv_n.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->st));
        B200_CHECK(cudaMemcpyAsync(tmp.data(), ctx->d_intv.p, tmp.size() * sizeof(Intv), cudaMemcpyDeviceToHost, ctx->st));
        B200_CHECK(cudaStreamSynchronize(ctx->st));
        for (int r = 0; r < n; ++r) {
            int k = out_n[r0 + r];
            if (tot + k > cap_total) { b200_set_error("interval output capacity too small"); return -2; }
            memcpy(out_intv + tot, &tmp[(size_t)r * cap_intv], (size_t)k * sizeof(Intv));
            tot += k;
        }
    }
    return tot;
}

extern "C" int b200_chain_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int stage, int n_reads, const uint8_t *seqs,
                                const int64_t *off, int64_t cap_chn, int64_t *out_chn, int32_t *out_nchn,
                                int64_t cap_seed, int64_t *out_seed, int64_t *totals)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    if (b200_batch_upload(ctx, n_reads, seqs, off)) return -1;
    i64 nc = 0, ns = 0;
    g_chain_stage = stage;
    const int chunk = eff_chunk(ctx);
    for (int r0 = 0; r0 < n_reads; r0 += chunk) {
        int n = std::min(chunk, n_reads - r0);
        int rc = b200_run_chunk(ctx, opt, r0, n, 3, nullptr);
        if (rc) { g_chain_stage = 1; return rc; }
        std::vector<i64> soff((size_t)n + 1);
        std::vector<i32> cn(n), sn(n);
        B200_CHECK(cudaMemcpyAsync(soff.data(), ctx->d_seed_off.p, ((size_t)n + 1) * 8, cudaMemcpyDeviceToHost, ctx->st));
        B200_CHECK(cudaMemcpyAsync(cn.data(), ctx->d_chn_n.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->st));
        B200_CHECK(cudaMemcpyAsync(sn.data(), ctx->d_cseed_n.p, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->st));
        B200_CHECK(cudaStreamSynchronize(ctx->st));
        std::vector<ChainRec> ch((size_t)soff[n] + 1);
        std::vector<Seed> sd((size_t)soff[n] + 1);
        if (soff[n]) {
            B200_CHECK(cudaMemcpyAsync(ch.data(), ctx->d_chains.p, (size_t)soff[n] * sizeof(ChainRec), cudaMemcpyDeviceToHost, ctx->st));
            B200_CHECK(cudaMemcpyAsync(sd.data(), ctx->d_cseeds.p, (size_t)soff[n] * sizeof(Seed), cudaMemcpyDeviceToHost, ctx->st));
            B200_CHECK(cudaStreamSynchronize(ctx->st));
        }
        for (int r = 0; r < n; ++r) {
            out_nchn[r0 + r] = cn[r];
            for (int i = 0; i < cn[r]; ++i) {
                const ChainRec &c = ch[(size_t)soff[r] + i];
                if (nc >= cap_chn || ns + c.n > cap_seed) { g_chain_stage = 1; b200_set_error("chain output capacity too small"); return -2; }
                i64 *o = out_chn + 8 * nc;
                u32 fb; memcpy(&fb, &c.frac_rep, 4);
                o[0] = c.pos; o[1] = c.n; o[2] = c.rid; o[3] = c.w; o[4] = c.kept; o[5] = c.first; o[6] = c.is_alt; o[7] = fb;
                for (int j = 0; j < c.n; ++j) {
                    const Seed &s = sd[(size_t)c.seed_off + j];
                    i64 *q = out_seed + 4 * (ns + j);
                    q[0] = s.rbeg; q[1] = s.qbeg; q[2] = s.len; q[3] = s.score;
                }
                ++nc; ns += c.n;
            }
        }
    }
    g_chain_stage = 1;
    totals[0] = nc; totals[1] = ns;
    return 0;
}

extern "C" int b200_smem1a_batch(b200_ctx_t *ctx, int64_t n, const uint8_t *seqs, const int64_t *off, const int32_t *x,
                                 const int32_t *min_intv, uint64_t max_intv, int cap, bwtintv_t *out_intv,
                                 int32_t *out_n, int32_t *out_ret)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    if (n <= 0) return 0;
    if (b200_batch_upload(ctx, (int)n, seqs, off)) return -1;
    cudaStream_t st = ctx->st;
    if (launch_pack_reads(ctx->d_seq.as<u8>(), ctx->n_bytes, st)) return -1;
    int blocks = (int)std::min<i64>((n + 127) / 128, ctx->n_sms * 4);
    const int list_cap = ctx->max_len + 1;
    if (ctx->d_intv.ensure((size_t)n * cap * sizeof(Intv)) || ctx->d_intv_n.ensure((size_t)n * 4) || ctx->d_pa.ensure((size_t)n * 4) ||
        ctx->d_pb.ensure((size_t)n * 4) || ctx->d_pc.ensure((size_t)n * 4) ||
        ctx->d_lists.ensure((size_t)blocks * 128 * 2 * list_cap * sizeof(Intv)) || ctx->d_ctl.ensure(CTL_SLOTS * 8)) return -1;
    if (ctl_zero(ctx, 0, CTL_SLOTS)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_pa.p, x, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pb.p, min_intv, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    SmemParams SP;
    memset(&SP, 0, sizeof(SP));
    SP.ix = ctx->ix; SP.seq = ctx->d_seq.as<u8>(); SP.off = ctx->d_off.as<i64>(); SP.n_reads = (int)n;
    SP.out = ctx->d_intv.as<Intv>(); SP.cap_out = cap; SP.out_n = ctx->d_intv_n.as<i32>();
    SP.lists = ctx->d_lists.as<Intv>(); SP.list_cap = list_cap;
    SP.work_counter = ctl_u32(ctx, CTL_WORK); SP.stats = ctl_u64(ctx, CTL_STATS); SP.overflow = (int *)ctl_u32(ctx, CTL_OVF);
    SP.px = ctx->d_pa.as<i32>(); SP.pmin_intv = ctx->d_pb.as<i32>(); SP.max_intv = max_intv; SP.out_ret = ctx->d_pc.as<i32>();
    if (launch_smem_single(SP, blocks, 128, st)) return -1;
    u64 ovf = 0;
    if (ctl_read(ctx, CTL_OVF, 1, &ovf)) return -1;
    if ((int)(ovf & 0xffffffff)) { b200_set_error("interval capacity too small (need %d)", (int)(ovf & 0xffffffff)); return -(int)(ovf & 0xffffffff); }
    B200_CHECK(cudaMemcpyAsync(out_intv, ctx->d_intv.p, (size_t)n * cap * sizeof(Intv), cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaMemcpyAsync(out_n, ctx->d_intv_n.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaMemcpyAsync(out_ret, ctx->d_pc.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    return 0;
}
