// Context, device index loader and the batched FM-index primitive entry points of the C ABI.
#include "ctx.h"
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include <algorithm>

static thread_local char g_err[512] = "";

void b200_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *b200_last_error(void) { return g_err; }

extern "C" b200_ctx_t *b200_ctx_create(int device)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        b200_set_error("no CUDA device available (this library has no CPU path)");
        return nullptr;
    }
    if (device < 0 || device >= ndev) { b200_set_error("device %d out of range (%d devices)", device, ndev); return nullptr; }
    if (cudaSetDevice(device) != cudaSuccess) { b200_set_error("cudaSetDevice(%d) failed", device); return nullptr; }
    b200_ctx *ctx = new b200_ctx();
    ctx->device = device;
#ifdef B200_EMU
    ctx->n_sms = 2;
#else
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->n_sms = prop.multiProcessorCount;
#endif
    if (cudaStreamCreateWithFlags(&ctx->st, cudaStreamNonBlocking) != cudaSuccess) {
        b200_set_error("cudaStreamCreate failed");
        delete ctx;
        return nullptr;
    }
    memset(&ctx->ix, 0, sizeof(ctx->ix));
    return ctx;
}

extern "C" int b200_ctx_tune(b200_ctx_t *ctx, const char *key, int64_t value)
{
    if (!ctx || !key) { b200_set_error("null argument"); return -1; }
    if (!strcmp(key, "chunk_reads")) { if (value < 1) value = 1; ctx->chunk_reads = (int)value; ctx->chunk_user = true; }
    else if (!strcmp(key, "smem_lcap")) ctx->smem_lcap = (int)value;
    else if (!strcmp(key, "cap_intv")) ctx->cap_intv = (int)value;
    else if (!strcmp(key, "ext_path")) ctx->ext_path = (int)value;
    else if (!strcmp(key, "smem_path")) ctx->smem_path = (int)value;
    else if (!strcmp(key, "smem_bps")) ctx->smem_bps = (int)value;
    else if (!strcmp(key, "split_p3")) ctx->split_p3 = (int)value;
    else if (!strcmp(key, "force_wide")) ctx->force_wide = (int)value;       // before the index upload
    else if (!strcmp(key, "cap_retry")) ctx->cap_retry = (int)value;
    else if (!strcmp(key, "ext_tail")) ctx->ext_tail = (int)value;
    else if (!strcmp(key, "seed_stats")) ctx->seed_stats = (int)value;
    else if (!strcmp(key, "r2a_warp")) ctx->r2a_warp = (int)value;
#ifndef B200_EMU
    else if (!strcmp(key, "l2_fetch")) {                                      // L2 fetch granularity hint in bytes (32 / 64 / 128), device-wide
        if (cudaSetDevice(ctx->device) != cudaSuccess || cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)value) != cudaSuccess) {
            cudaGetLastError(); b200_set_error("cudaLimitMaxL2FetchGranularity %lld refused", (long long)value); return -1;
        }
    }
#endif
    else if (!strcmp(key, "kmer_k")) ctx->kmer_k = (int)value;               // before the index upload
    else if (!strcmp(key, "ext_bin0")) ctx->ext_bin0 = (int)value;
    else if (!strcmp(key, "ext_bin1")) ctx->ext_bin1 = (int)value;
    else if (!strcmp(key, "ext_bps")) ctx->ext_bps = (int)value;
    else if (!strcmp(key, "ext_rowsync")) ctx->ext_rowsync = (int)value;
    else if (!strcmp(key, "ext_sort")) ctx->ext_sort = (int)value;
    else if (!strcmp(key, "ext_phased")) ctx->ext_phased = (int)value;
    else if (!strcmp(key, "ssw_path")) ctx->ssw_path = (int)value;
    else if (!strcmp(key, "r2a_small")) ctx->r2a_small = (int)value;
    else if (!strcmp(key, "l2_persist")) ctx->l2_persist = (int)value;
    else if (!strcmp(key, "idx_group_cap")) ctx->idx_group_cap = value > ((i64)1 << 30) ? ((i64)1 << 30) : value;   // before b200_index_build_device
    else if (!strcmp(key, "idx_tie_cap")) ctx->idx_tie_cap = value;
    else { b200_set_error("unknown tunable '%s'", key); return -1; }
    return 0;
}

// page-locked host memory for callers that want the copies of the batched calls to run at full speed and overlap
extern "C" void *b200_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); b200_set_error("cudaHostAlloc(%zu) failed", bytes); return nullptr; }
    return p;
}
extern "C" void b200_host_free(void *p) { if (p) cudaFreeHost(p); }
// grow-only page-locked buffer owned by the context (released by b200_ctx_destroy); contents are not preserved on growth
extern "C" void *b200_ctx_pinned(b200_ctx_t *ctx, int slot, size_t bytes)
{
    if (!ctx || slot < 0 || slot >= 16) { b200_set_error("b200_ctx_pinned: bad slot"); return nullptr; }
    if (bytes <= ctx->pin_cap[slot] && ctx->pin_p[slot]) return ctx->pin_p[slot];
    cudaSetDevice(ctx->device);
    b200_host_free(ctx->pin_p[slot]);
    ctx->pin_cap[slot] = bytes + bytes / 4 + 4096;
    ctx->pin_p[slot] = b200_host_alloc(ctx->pin_cap[slot]);
    if (!ctx->pin_p[slot]) ctx->pin_cap[slot] = 0;
    return ctx->pin_p[slot];
}

extern "C" void b200_ctx_destroy(b200_ctx_t *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->st);
    for (void *&q : ctx->pin_p) { b200_host_free(q); q = nullptr; }
    DevBuf *bufs[] = { &ctx->d_occ32, &ctx->d_kmer, &ctx->d_retry, &ctx->d_pk, &ctx->d_spill, &ctx->d_jobs0, &ctx->d_jobs1, &ctx->d_jobs2, &ctx->d_jobs3, &ctx->d_jsort, &ctx->d_rstate, &ctx->d_njobs, &ctx->d_bwt, &ctx->d_sa, &ctx->d_sa_full, &ctx->d_pac, &ctx->d_ann_off, &ctx->d_ann_len, &ctx->d_ann_alt,
                       &ctx->d_seq, &ctx->d_off, &ctx->d_intv, &ctx->d_intv_n, &ctx->d_lists, &ctx->d_seed_n, &ctx->d_seed_off,
                       &ctx->d_l_rep, &ctx->d_seeds, &ctx->d_cseeds, &ctx->d_chains, &ctx->d_chn_n, &ctx->d_cseed_n,
                       &ctx->d_scratch, &ctx->d_stage, &ctx->d_eh, &ctx->d_srt, &ctx->d_arena, &ctx->d_reg_off, &ctx->d_reg_n,
                       &ctx->d_dst_off, &ctx->d_scan_tmp, &ctx->d_ctl, &ctx->d_final, &ctx->d_final_n,
                       &ctx->d_pa, &ctx->d_pb, &ctx->d_pc, &ctx->d_pd, &ctx->d_pe, &ctx->d_pf, &ctx->d_pg, &ctx->d_ph };
    for (DevBuf *b : bufs) b->release();
    for (DevBuf &b : ctx->d_r2a) b.release();
    for (DevBuf &b : ctx->d_sam) b.release();
    for (cudaEvent_t &e : ctx->sam_ev) if (e) { cudaEventDestroy(e); e = nullptr; }
    if (ctx->st_h2d) cudaStreamDestroy(ctx->st_h2d);
    if (ctx->st_d2h) cudaStreamDestroy(ctx->st_d2h);
    cudaStreamDestroy(ctx->st);
    delete ctx;
}

// Gives the device memory of the batch pipeline back (resident batch, stage buffers, worker2 / SAM scratch); the index and
// the context stay usable, buffers grow again on the next call.
extern "C" int b200_ctx_trim(b200_ctx_t *ctx)
{
    if (!ctx) { b200_set_error("null context"); return -1; }
    std::lock_guard<std::mutex> lk(ctx->call_mu);
    B200_CHECK(cudaSetDevice(ctx->device));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    DevBuf *bufs[] = { &ctx->d_retry, &ctx->d_pk, &ctx->d_spill, &ctx->d_jobs0, &ctx->d_jobs1, &ctx->d_jobs2, &ctx->d_jobs3, &ctx->d_jsort, &ctx->d_rstate, &ctx->d_njobs,
                       &ctx->d_seq, &ctx->d_off, &ctx->d_intv, &ctx->d_intv_n, &ctx->d_lists, &ctx->d_seed_n, &ctx->d_seed_off,
                       &ctx->d_l_rep, &ctx->d_seeds, &ctx->d_cseeds, &ctx->d_chains, &ctx->d_chn_n, &ctx->d_cseed_n,
                       &ctx->d_scratch, &ctx->d_stage, &ctx->d_eh, &ctx->d_srt, &ctx->d_arena, &ctx->d_reg_off, &ctx->d_reg_n,
                       &ctx->d_dst_off, &ctx->d_scan_tmp, &ctx->d_final, &ctx->d_final_n,
                       &ctx->d_pa, &ctx->d_pb, &ctx->d_pc, &ctx->d_pd, &ctx->d_pe, &ctx->d_pf, &ctx->d_pg, &ctx->d_ph };
    for (DevBuf *b : bufs) b->release();
    for (DevBuf &b : ctx->d_r2a) b.release();
    for (DevBuf &b : ctx->d_sam) b.release();
    ctx->n_reads = 0; ctx->n_bytes = 0;
    return 0;
}

// The full suffix array of a text of 2^32 or more rows as 5 bytes per row instead of 8 (50 -> 31 GB at 3.1 Gbp, on the
// device, in the index file and in the device-to-device fan-out); a lookup is two aligned 32-bit loads (fmindex.cuh).  With
// force_wide (tests) a 4-byte array is packed too, so that small genomes exercise the path.
int b200_pack_full_sa(b200_ctx *ctx)
{
    DevIndex &ix = ctx->ix;
    const bool from64 = ix.sa_full64 != nullptr;
    if (!from64 && !(ix.sa_full32 && ctx->force_wide)) return 0;
    const u64 n = ix.seq_len + 1;
    DevBuf packed;
    if (packed.ensure_exact(sa40_bytes(n))) return -1;
    if (launch_pack_sa40(from64 ? (const void *)ix.sa_full64 : (const void *)ix.sa_full32, from64, packed.as<u32>(), n, ctx->st)) { packed.release(); return -1; }
    if (cudaStreamSynchronize(ctx->st) != cudaSuccess) { packed.release(); b200_set_error("packing the suffix array failed"); return -1; }
    ctx->d_sa_full.release();
    ctx->d_sa_full = packed;
    ix.sa_full40 = ctx->d_sa_full.as<u32>(); ix.sa_full32 = nullptr; ix.sa_full64 = nullptr;
    return 0;
}

// Derived device structures of an index whose reference-layout arrays are resident (uploaded or built on the device):
// occ32 re-layout, k-mer interval table, and the full suffix array unless it is there already.
int b200_index_finish(b200_ctx *ctx, int flags)
{
    DevIndex &ix = ctx->ix;
    const u64 seq_len = ix.seq_len;
    for (int s = 0; s < 8; ++s)
        for (int c = 0; c < 4; ++c) { u64 g = 0; for (int d = c + 1; d < 4; ++d) g += ix.sb_cnt[s][d]; ix.sb_gt[s][c] = g; }
    {   // occ32 re-layout for the seeding kernel (occ32.cuh); superblock bases already in ix.sb_cnt
        const u64 n_blk = (seq_len >> 6) + 2;
        if (ctx->d_bwt.bytes < ((n_blk + 1) / 2) * 64) { b200_set_error("internal: bwt padding too small"); return -1; }
        if (ctx->d_occ32.ensure_exact((size_t)n_blk * 32)) return -1;
        if (launch_build_occ32(ix, ctx->d_occ32.as<uint4>(), n_blk, ctx->st)) return -1;
        ix.occ32 = ctx->d_occ32.as<uint4>();
        ix.kmer_tab = nullptr; ix.kmer_k = 0;
        // default depth: the length at which a random string becomes unique in this text, floor(log4 seq_len), between 9
        // and 15: 15 at 3.1 Gbp (22.9 GB of table next to the 57 GB index), 13 at 100 Mbp (1.4 GB), 11 at 5 Mbp.  Deeper
        // is faster up to there (profiles/r6q_seed_table_depth.txt): more list entries are bits, fewer extensions pay for
        // two index sectors and the popcounts.
        int auto_k = 0;
        for (u64 t = seq_len; t >= 4; t >>= 2) ++auto_k;
        auto_k = std::max(9, std::min(15, auto_k));
#ifdef B200_EMU
        auto_k = 9;                                                  // the emulated builder is slow; tests ask for deeper tables by name
#endif
        const int want_k = ctx->kmer_k >= 0 ? ctx->kmer_k : auto_k;
        if (want_k > 0) {                                            // k-mer interval table (kmer_tab.cu)
            const int wide = b200_wide_mode(ctx);                               // same rule as the seeding kernel template (pipeline.cu)
            const int kk = want_k > 15 ? 15 : want_k;                  // kmer_off[16]
            if (ctx->d_kmer.ensure_exact(kmer_tab_entries(kk) * (wide == 1 ? 32 : 16))) return -1;
            if (launch_build_kmer_tab(ix, ctx->d_kmer.as<uint4>(), kk, wide, ctx->st)) return -1;
            ix.kmer_tab = ctx->d_kmer.as<uint4>(); ix.kmer_k = kk;
        }
#ifndef B200_EMU
        {   // optional (tunable l2_persist=1, before the upload): persisting-L2 window over the seeding index.
            // Measured on B200: seeding -2 %, but the other stages lose L2 and the batch gets slower, so it is off.
            cudaDeviceProp prop;
            if (cudaGetDeviceProperties(&prop, ctx->device) == cudaSuccess && prop.persistingL2CacheMaxSize > 0 && ctx->l2_persist) {
                const size_t setaside = (size_t)prop.persistingL2CacheMaxSize;
                cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, setaside);
                cudaStreamAttrValue av;
                memset(&av, 0, sizeof(av));
                size_t win = std::min<size_t>((size_t)n_blk * 32, (size_t)prop.accessPolicyMaxWindowSize);
                av.accessPolicyWindow.base_ptr = ctx->d_occ32.p;
                av.accessPolicyWindow.num_bytes = win;
                av.accessPolicyWindow.hitRatio = win <= setaside ? 1.0f : (float)((double)setaside / (double)win);
                av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                if (cudaStreamSetAttribute(ctx->st, cudaStreamAttributeAccessPolicyWindow, &av) != cudaSuccess) cudaGetLastError();
                ctx->l2_setaside = setaside; ctx->l2_window = win;
            }
        }
#endif
    }
    if ((flags & B200_IDX_FULL_SA) && !ix.sa_full32 && !ix.sa_full64 && !ix.sa_full40) {
        const bool narrow = seq_len < 0xffffffffull;
        if (ctx->d_sa_full.ensure_exact((size_t)(seq_len + 1) * (narrow ? 4 : 8))) return -1;
        if (launch_build_full_sa(ix, narrow ? ctx->d_sa_full.as<u32>() : nullptr, narrow ? nullptr : ctx->d_sa_full.as<u64>(), ctx->st)) return -1;
        if (narrow) ix.sa_full32 = ctx->d_sa_full.as<u32>(); else ix.sa_full64 = ctx->d_sa_full.as<u64>();
    }
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    return b200_pack_full_sa(ctx);
}

extern "C" int b200_index_upload_raw(b200_ctx_t *ctx, const uint32_t *bwt, uint64_t bwt_size, uint64_t primary,
                                     const uint64_t L2[5], uint64_t seq_len, const uint64_t *sa, uint64_t n_sa, int sa_intv,
                                     const uint8_t *pac, int64_t l_pac, int n_seqs, const int64_t *ann_offset,
                                     const int32_t *ann_len, const int32_t *ann_is_alt, int flags)
{
    if (!ctx) { b200_set_error("null context"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    if (sa_intv <= 0 || (sa_intv & (sa_intv - 1))) { b200_set_error("sa_intv must be a power of two"); return -1; }
    // Occ blocks: verbatim, padded to a whole number of 64-byte blocks (+1 so that row seq_len-1 is always inside)
    size_t bwt_bytes = ((size_t)bwt_size * 4 + 63) / 64 * 64 + 128;
    if (ctx->d_bwt.ensure(bwt_bytes)) return -1;
    B200_CHECK(cudaMemsetAsync(ctx->d_bwt.p, 0, bwt_bytes, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_bwt.p, bwt, (size_t)bwt_size * 4, cudaMemcpyHostToDevice, ctx->st));
    if (ctx->d_sa.ensure((size_t)n_sa * 8)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_sa.p, sa, (size_t)n_sa * 8, cudaMemcpyHostToDevice, ctx->st));
    size_t pac_bytes = (size_t)(l_pac / 4 + 1);
    if (ctx->d_pac.ensure(pac_bytes)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_pac.p, pac, pac_bytes, cudaMemcpyHostToDevice, ctx->st));
    if (ctx->d_ann_off.ensure((size_t)n_seqs * 8) || ctx->d_ann_len.ensure((size_t)n_seqs * 4) || ctx->d_ann_alt.ensure((size_t)n_seqs * 4)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_ann_off.p, ann_offset, (size_t)n_seqs * 8, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_ann_len.p, ann_len, (size_t)n_seqs * 4, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_ann_alt.p, ann_is_alt, (size_t)n_seqs * 4, cudaMemcpyHostToDevice, ctx->st));

    DevIndex &ix = ctx->ix;
    memset(&ix, 0, sizeof(ix));
    ix.bwt = ctx->d_bwt.as<uint4>();
    ix.primary = primary; ix.seq_len = seq_len;
    for (int i = 0; i < 5; ++i) ix.L2[i] = L2[i];
    ix.sa = ctx->d_sa.as<u64>(); ix.n_sa = n_sa; ix.sa_intv = sa_intv;
    ix.pac = ctx->d_pac.as<u8>(); ix.l_pac = l_pac; ix.n_seqs = n_seqs;
    ix.ann_off = ctx->d_ann_off.as<i64>(); ix.ann_len = ctx->d_ann_len.as<i32>(); ix.ann_alt = ctx->d_ann_alt.as<i32>();
    for (int s = 0; s < 8; ++s)                                      // superblock bases from the host checkpoints
        for (int c = 0; c < 4; ++c) ix.sb_cnt[s][c] = 0;
    if (seq_len >> 35) { b200_set_error("seq_len >= 2^35 not supported by the occ32 layout"); return -1; }
    for (u64 s = 1; (s << 32) < seq_len; ++s) {
        const uint32_t *blk = bwt + (((s << 32) >> 7) << 4);
        for (int c = 0; c < 4; ++c) ix.sb_cnt[s][c] = (u64)blk[2 * c] | (u64)blk[2 * c + 1] << 32;
    }
    ctx->idx_bwt_size = bwt_size;
    if (b200_index_finish(ctx, flags)) return -1;
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    ctx->has_index = true;
    return 0;
}

extern "C" int b200_index_upload(b200_ctx_t *ctx, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac, int flags)
{
    if (!bwt || !bns || !pac) { b200_set_error("null index component"); return -1; }
    std::vector<int64_t> off(bns->n_seqs);
    std::vector<int32_t> len(bns->n_seqs), alt(bns->n_seqs);
    for (int i = 0; i < bns->n_seqs; ++i) { off[i] = bns->anns[i].offset; len[i] = bns->anns[i].len; alt[i] = bns->anns[i].is_alt; }
    int rc = b200_index_upload_raw(ctx, bwt->bwt, bwt->bwt_size, bwt->primary, bwt->L2, bwt->seq_len, bwt->sa, bwt->n_sa,
                                   bwt->sa_intv, pac, bns->l_pac, bns->n_seqs, off.data(), len.data(), alt.data(), flags);
    if (rc == 0) ctx->bwt_identity = bwt;
    return rc;
}

// ---- drop-in registry: one context per host bwt_t (created lazily)
static std::mutex g_reg_mu;
static std::vector<b200_ctx *> g_reg;

b200_ctx *b200_ctx_for_bwt(const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac)
{
    std::lock_guard<std::mutex> lk(g_reg_mu);
    b200_ctx *c = nullptr;
    for (b200_ctx *r : g_reg) if (r->bwt_identity == bwt) c = r;
    if (c && (!pac || !c->stub_pac)) return c;
    const bool fresh = c == nullptr;            // else: created by an FM-index primitive, now needs the real reference
    if (fresh) c = b200_ctx_create(0);
    if (!c) { fprintf(stderr, "[b200] %s\n", b200_last_error()); exit(1); }
    // primitives that only need the FM-index may arrive without bns/pac: upload a 1-contig stub
    bntseq_t stub; bntann1_t ann; uint8_t zero = 0;
    if (!bns) {
        memset(&stub, 0, sizeof(stub)); memset(&ann, 0, sizeof(ann));
        stub.l_pac = (int64_t)(bwt->seq_len / 2); stub.n_seqs = 1; stub.anns = &ann; ann.len = (int32_t)stub.l_pac;
        bns = &stub;
    }
    int rc;
    if (pac) rc = b200_index_upload(c, bwt, bns, pac, B200_IDX_FULL_SA);
    else {
        std::vector<uint8_t> z((size_t)(bns->l_pac / 4 + 1), zero);
        rc = b200_index_upload(c, bwt, bns, z.data(), 0);
    }
    if (rc) { fprintf(stderr, "[b200] %s\n", b200_last_error()); exit(1); }
    c->bwt_identity = bwt;
    c->stub_pac = pac == nullptr;
    if (fresh) g_reg.push_back(c);
    return c;
}

// ---- one index, several contexts ---------------------------------------------------------------
// The FM-index is read-only, so contexts can have it in common.  Same device: `dst` borrows src's arrays (no copy, no
// memory; src must outlive dst) -- two contexts per device let two host threads keep one GPU busy across each other's host
// gaps and synchronisations.  Another device: the arrays are copied device to device (cudaMemcpyPeerAsync: NVLink where the
// box has it) instead of being uploaded or built again -- the "one H2D + 7 peer copies" fan-out of SURVEY.md 8(e).
extern "C" int b200_index_replicate(b200_ctx_t *dst, b200_ctx_t *src)
{
    if (!dst || !src || !src->has_index) { b200_set_error("b200_index_replicate: source context has no index"); return -1; }
    if (dst == src) return 0;
    dst->force_wide = src->force_wide; dst->kmer_k = src->kmer_k;
    dst->idx_bwt_size = src->idx_bwt_size; dst->idx_groups = src->idx_groups; dst->idx_ties = src->idx_ties;
    dst->bwt_identity = nullptr; dst->stub_pac = src->stub_pac;
    if (dst->device == src->device) {
        dst->ix = src->ix;
        dst->has_index = true;
        return 0;
    }
#ifdef B200_EMU
    b200_set_error("b200_index_replicate: one device only in the emulated build"); return -1;
#else
    B200_CHECK(cudaSetDevice(dst->device));
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, dst->device, src->device) == cudaSuccess && can) {
        cudaError_t e = cudaDeviceEnablePeerAccess(src->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else cudaGetLastError();
    }
    DevBuf *sb[] = { &src->d_bwt, &src->d_sa, &src->d_sa_full, &src->d_pac, &src->d_ann_off, &src->d_ann_len, &src->d_ann_alt, &src->d_occ32, &src->d_kmer };
    DevBuf *db[] = { &dst->d_bwt, &dst->d_sa, &dst->d_sa_full, &dst->d_pac, &dst->d_ann_off, &dst->d_ann_len, &dst->d_ann_alt, &dst->d_occ32, &dst->d_kmer };
    const int nb = (int)(sizeof(sb) / sizeof(sb[0]));
    for (int i = 0; i < nb; ++i) {
        if (!sb[i]->p) continue;
        if (db[i]->ensure_exact(sb[i]->bytes)) return -1;
        B200_CHECK(cudaMemcpyPeerAsync(db[i]->p, dst->device, sb[i]->p, src->device, sb[i]->bytes, dst->st));
    }
    B200_CHECK(cudaStreamSynchronize(dst->st));
    auto move = [&](const void *p) -> const void * {       // the same offset inside the copy of the array that holds p
        if (!p) return nullptr;
        for (int i = 0; i < nb; ++i) {
            const char *b = (const char *)sb[i]->p;
            if (b && (const char *)p >= b && (const char *)p < b + sb[i]->bytes) return (const char *)db[i]->p + ((const char *)p - b);
        }
        return nullptr;
    };
    DevIndex ix = src->ix;
    ix.bwt = (const uint4 *)move(src->ix.bwt); ix.sa = (const u64 *)move(src->ix.sa);
    ix.sa_full32 = (const u32 *)move(src->ix.sa_full32); ix.sa_full64 = (const u64 *)move(src->ix.sa_full64);
    ix.sa_full40 = (const u32 *)move(src->ix.sa_full40);
    ix.pac = (const u8 *)move(src->ix.pac); ix.ann_off = (const i64 *)move(src->ix.ann_off);
    ix.ann_len = (const i32 *)move(src->ix.ann_len); ix.ann_alt = (const i32 *)move(src->ix.ann_alt);
    ix.occ32 = (const uint4 *)move(src->ix.occ32); ix.kmer_tab = (const uint4 *)move(src->ix.kmer_tab);
    if (!ix.bwt || !ix.pac || !ix.occ32 || (src->ix.kmer_tab && !ix.kmer_tab) || (src->ix.sa_full64 && !ix.sa_full64) || (src->ix.sa_full32 && !ix.sa_full32) || (src->ix.sa_full40 && !ix.sa_full40)) {
        b200_set_error("b200_index_replicate: an index array lies outside the context's buffers"); return -1;
    }
    dst->ix = ix;
    dst->has_index = true;
    return 0;
#endif
}

// ---- primitive batches -----------------------------------------------------------------------
static int need_index(b200_ctx *ctx)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) { b200_set_error("cudaSetDevice failed"); return -1; }
    return 0;
}

extern "C" int b200_occ4_batch(b200_ctx_t *ctx, int64_t n, const uint64_t *k, uint64_t *cnt)
{
    if (need_index(ctx)) return -1;
    if (n <= 0) return 0;
    if (ctx->d_pa.ensure(n * 8) || ctx->d_pb.ensure(n * 32)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_pa.p, k, n * 8, cudaMemcpyHostToDevice, ctx->st));
    if (launch_occ_batch(ctx->ix, n, ctx->d_pa.as<u64>(), nullptr, ctx->d_pb.as<u64>(), nullptr, ctx->st)) return -1;
    B200_CHECK(cudaMemcpyAsync(cnt, ctx->d_pb.p, n * 32, cudaMemcpyDeviceToHost, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    return 0;
}

extern "C" int b200_2occ4_batch(b200_ctx_t *ctx, int64_t n, const uint64_t *k, const uint64_t *l, uint64_t *cntk, uint64_t *cntl)
{
    if (need_index(ctx)) return -1;
    if (n <= 0) return 0;
    if (ctx->d_pa.ensure(n * 8) || ctx->d_pb.ensure(n * 32) || ctx->d_pc.ensure(n * 8) || ctx->d_pd.ensure(n * 32)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_pa.p, k, n * 8, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pc.p, l, n * 8, cudaMemcpyHostToDevice, ctx->st));
    if (launch_occ_batch(ctx->ix, n, ctx->d_pa.as<u64>(), ctx->d_pc.as<u64>(), ctx->d_pb.as<u64>(), ctx->d_pd.as<u64>(), ctx->st)) return -1;
    B200_CHECK(cudaMemcpyAsync(cntk, ctx->d_pb.p, n * 32, cudaMemcpyDeviceToHost, ctx->st));
    B200_CHECK(cudaMemcpyAsync(cntl, ctx->d_pd.p, n * 32, cudaMemcpyDeviceToHost, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    return 0;
}

extern "C" int b200_extend_batch(b200_ctx_t *ctx, int64_t n, const uint64_t *ik, int is_back, uint64_t *ok)
{
    if (need_index(ctx)) return -1;
    if (n <= 0) return 0;
    if (ctx->d_pa.ensure(n * 24) || ctx->d_pb.ensure(n * 96)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_pa.p, ik, n * 24, cudaMemcpyHostToDevice, ctx->st));
    if (launch_extend_batch(ctx->ix, n, ctx->d_pa.as<u64>(), is_back, ctx->d_pb.as<u64>(), ctx->st)) return -1;
    B200_CHECK(cudaMemcpyAsync(ok, ctx->d_pb.p, n * 96, cudaMemcpyDeviceToHost, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    return 0;
}

extern "C" int b200_sa_batch(b200_ctx_t *ctx, int64_t n, const uint64_t *k, uint64_t *sa, int use_full_sa)
{
    if (need_index(ctx)) return -1;
    if (n <= 0) return 0;
    if (ctx->d_pa.ensure(n * 8) || ctx->d_pb.ensure(n * 8)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_pa.p, k, n * 8, cudaMemcpyHostToDevice, ctx->st));
    if (launch_sa_batch(ctx->ix, n, ctx->d_pa.as<u64>(), ctx->d_pb.as<u64>(), use_full_sa, ctx->st)) return -1;
    B200_CHECK(cudaMemcpyAsync(sa, ctx->d_pb.p, n * 8, cudaMemcpyDeviceToHost, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    return 0;
}

extern "C" int b200_ksw_extend2_batch(b200_ctx_t *ctx, int64_t n, const int32_t *qlen, const int64_t *qoff, const uint8_t *q,
                                      const int32_t *tlen, const int64_t *toff, const uint8_t *t, const int8_t *mat,
                                      int o_del, int e_del, int o_ins, int e_ins, const int32_t *w, int end_bonus,
                                      int zdrop, const int32_t *h0, int32_t *out)
{
    if (!ctx) { b200_set_error("null context"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    if (n <= 0) return 0;
    i64 qbytes = 0, tbytes = 0; int maxq = 0;
    for (i64 i = 0; i < n; ++i) {
        if (qoff[i] + qlen[i] > qbytes) qbytes = qoff[i] + qlen[i];
        if (toff[i] + tlen[i] > tbytes) tbytes = toff[i] + tlen[i];
        if (qlen[i] > maxq) maxq = qlen[i];
        if (h0[i] <= 0) { b200_set_error("ksw_extend2 requires h0 > 0 (ksw.c:385)"); return -1; }
    }
    const int grid = ctx->n_sms * 4;
    const int eh_cap = maxq + 2;
    if (ctx->d_pa.ensure(n * 4) || ctx->d_pb.ensure(n * 8) || ctx->d_pc.ensure(qbytes + 1) || ctx->d_pd.ensure(n * 4) ||
        ctx->d_pe.ensure(n * 8) || ctx->d_pf.ensure(tbytes + 1) || ctx->d_pg.ensure(n * 8) || ctx->d_ph.ensure(n * 24) ||
        ctx->d_eh.ensure((size_t)grid * 4 * 2 * eh_cap * 4) || ctx->d_ctl.ensure(CTL_SLOTS * 8)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_pa.p, qlen, n * 4, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pb.p, qoff, n * 8, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pc.p, q, qbytes, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pd.p, tlen, n * 4, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pe.p, toff, n * 8, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pf.p, t, tbytes, cudaMemcpyHostToDevice, ctx->st));
    i32 *wh = ctx->d_pg.as<i32>();
    B200_CHECK(cudaMemcpyAsync(wh, w, n * 4, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(wh + n, h0, n * 4, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemsetAsync(ctx->d_ctl.p, 0, CTL_SLOTS * 8, ctx->st));
    KswBatchParams P;
    P.n = n; P.qlen = ctx->d_pa.as<i32>(); P.qoff = ctx->d_pb.as<i64>(); P.q = ctx->d_pc.as<u8>();
    P.tlen = ctx->d_pd.as<i32>(); P.toff = ctx->d_pe.as<i64>(); P.t = ctx->d_pf.as<u8>();
    memcpy(P.mat, mat, 25);
    P.o_del = o_del; P.e_del = e_del; P.o_ins = o_ins; P.e_ins = e_ins; P.w = wh; P.end_bonus = end_bonus; P.zdrop = zdrop;
    P.h0 = wh + n; P.out = ctx->d_ph.as<i32>(); P.eh = ctx->d_eh.as<i32>(); P.eh_cap = eh_cap;
    P.work_counter = (unsigned *)(ctx->d_ctl.as<u64>() + CTL_WORK);
    if (launch_ksw_batch(P, grid, 128, ctx->st)) return -1;
    B200_CHECK(cudaMemcpyAsync(out, ctx->d_ph.p, n * 24, cudaMemcpyDeviceToHost, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    return 0;
}
