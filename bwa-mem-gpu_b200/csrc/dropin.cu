// Drop-in scalar entry points with the reference's own names and signatures.
// Every call runs the DEVICE code on a batch of one; there is no host implementation behind them.
// Error behaviour mirrors the reference: no status codes, failures print and exit (err_fatal style).
#include "ctx.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>

static void die_on(int rc, const char *what)
{
    if (rc < 0) { fprintf(stderr, "[b200::%s] %s\n", what, b200_last_error()); exit(1); }
}

extern "C" void bwt_occ4(const bwt_t *bwt, bwtint_t k, bwtint_t cnt[4])
{
    b200_ctx *ctx = b200_ctx_for_bwt(bwt, nullptr, nullptr);
    std::lock_guard<std::mutex> lk(ctx->call_mu);
    die_on(b200_occ4_batch(ctx, 1, &k, cnt), "bwt_occ4");
}

extern "C" void bwt_2occ4(const bwt_t *bwt, bwtint_t k, bwtint_t l, bwtint_t cntk[4], bwtint_t cntl[4])
{
    b200_ctx *ctx = b200_ctx_for_bwt(bwt, nullptr, nullptr);
    std::lock_guard<std::mutex> lk(ctx->call_mu);
    die_on(b200_2occ4_batch(ctx, 1, &k, &l, cntk, cntl), "bwt_2occ4");
}

extern "C" void bwt_extend(const bwt_t *bwt, const bwtintv_t *ik, bwtintv_t ok[4], int is_back)
{
    uint64_t in[3] = { ik->x[0], ik->x[1], ik->x[2] }, out[12];
    b200_ctx *ctx = b200_ctx_for_bwt(bwt, nullptr, nullptr);
    {
        std::lock_guard<std::mutex> lk(ctx->call_mu);
        die_on(b200_extend_batch(ctx, 1, in, is_back, out), "bwt_extend");
    }
    for (int c = 0; c < 4; ++c) { ok[c].x[0] = out[3 * c]; ok[c].x[1] = out[3 * c + 1]; ok[c].x[2] = out[3 * c + 2]; }
}

extern "C" bwtint_t bwt_sa(const bwt_t *bwt, bwtint_t k)
{
    uint64_t out = 0;
    b200_ctx *ctx = b200_ctx_for_bwt(bwt, nullptr, nullptr);
    std::lock_guard<std::mutex> lk(ctx->call_mu);
    die_on(b200_sa_batch(ctx, 1, &k, &out, 0), "bwt_sa");
    return out;
}

extern "C" int bwt_smem1a(const bwt_t *bwt, int len, const uint8_t *q, int x, int min_intv, uint64_t max_intv,
                          bwtintv_v *mem, bwtintv_v *tmpvec[2])
{
    (void)tmpvec;                       // scratch vectors of the CPU version; the device keeps its own
    b200_ctx *ctx = b200_ctx_for_bwt(bwt, nullptr, nullptr);
    int64_t off[2] = { 0, len };
    int32_t n = 0, ret = 0, xx = x, mi = min_intv;
    int cap = len + 1;
    bwtintv_t *tmp = (bwtintv_t *)malloc(sizeof(bwtintv_t) * (size_t)cap);
    {
        std::lock_guard<std::mutex> lk(ctx->call_mu);
        die_on(b200_smem1a_batch(ctx, 1, q, off, &xx, &mi, max_intv, cap, tmp, &n, &ret), "bwt_smem1a");
    }
    if (mem->m < n) { mem->m = n; mem->a = (bwtintv_t *)realloc(mem->a, sizeof(bwtintv_t) * (size_t)n); }
    mem->n = n;
    if (n) memcpy(mem->a, tmp, sizeof(bwtintv_t) * (size_t)n);
    free(tmp);
    return ret;
}

// Threading contract of the same-name shims: like the reference's functions they may be called from several host
// threads at once (kt_for workers call bwt_smem1a / ksw_extend2 / ksw_align2); calls that share a context are serialised
// by that context's mutex (they share its scratch buffers and stream), so they are safe, not concurrent.
static b200_ctx *g_ksw_ctx = nullptr;
static std::once_flag g_ksw_once;
static b200_ctx *ksw_ctx(const char *who)
{
    std::call_once(g_ksw_once, []() { g_ksw_ctx = b200_ctx_create(0); });
    if (!g_ksw_ctx) die_on(-1, who);
    return g_ksw_ctx;
}

extern "C" int ksw_extend2(int qlen, const uint8_t *query, int tlen, const uint8_t *target, int m, const int8_t *mat,
                           int o_del, int e_del, int o_ins, int e_ins, int w, int end_bonus, int zdrop, int h0,
                           int *qle, int *tle, int *gtle, int *gscore, int *max_off)
{
    if (m != 5) { fprintf(stderr, "[b200::ksw_extend2] only the 5-letter nt4 alphabet is supported (m=%d)\n", m); exit(1); }
    b200_ctx *kc = ksw_ctx("ksw_extend2");
    std::lock_guard<std::mutex> lk(kc->call_mu);
    int32_t ql = qlen, tl = tlen, ww = w, hh = h0, out[6];
    int64_t zero = 0;
    die_on(b200_ksw_extend2_batch(kc, 1, &ql, &zero, query, &tl, &zero, target, mat, o_del, e_del, o_ins, e_ins,
                                  &ww, end_bonus, zdrop, &hh, out), "ksw_extend2");
    if (qle) *qle = out[1];
    if (tle) *tle = out[2];
    if (gtle) *gtle = out[3];
    if (gscore) *gscore = out[4];
    if (max_off) *max_off = out[5];
    return out[0];
}

// The query profile of the reference (kswq_t, ksw.c:50-56) is private to ksw.c; callers only pass the pointer back and
// free() it.  Ours records the score width chosen at the first call, which is all that later calls depend on (ksw.c:350-353).
struct _kswq_t { int qlen, size; };

extern "C" kswr_t ksw_align2(int qlen, uint8_t *query, int tlen, uint8_t *target, int m, const int8_t *mat, int o_del,
                             int e_del, int o_ins, int e_ins, int xtra, kswq_t **qry)
{
    if (m != 5) { fprintf(stderr, "[b200::ksw_align2] only the 5-letter nt4 alphabet is supported (m=%d)\n", m); exit(1); }
    b200_ctx *kc = ksw_ctx("ksw_align2");
    std::lock_guard<std::mutex> lk(kc->call_mu);
    int size = (xtra & 0x10000) ? 1 : 2;                        // KSW_XBYTE, ksw.c:350
    if (qry && *qry) size = (*qry)->size;
    else if (qry) { *qry = (kswq_t *)malloc(sizeof(kswq_t)); (*qry)->qlen = qlen; (*qry)->size = size; }
    int32_t ql = qlen, tl = tlen, xx = (xtra & ~0x10000) | (size == 1 ? 0x10000 : 0), out[7];
    int64_t zero = 0;
    die_on(b200_ksw_align2_batch(kc, 1, &ql, &zero, query, &tl, &zero, target, mat, o_del, e_del, o_ins, e_ins, &xx, out),
           "ksw_align2");
    kswr_t r;
    r.score = out[0]; r.te = out[1]; r.qe = out[2]; r.score2 = out[3]; r.te2 = out[4]; r.tb = out[5]; r.qb = out[6];
    return r;
}

extern "C" mem_alnreg_v mem_align1_core(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac,
                                        int l_seq, char *seq, void *buf)
{
    (void)buf;
    b200_ctx *ctx = b200_ctx_for_bwt(bwt, bns, pac);
    // the reference converts the caller's buffer to nt4 in place (bwamem.c:1067-1068); keep that side effect
    static const unsigned char tab[4] = { 'A', 'C', 'G', 'T' };
    for (int i = 0; i < l_seq; ++i) {
        unsigned char c = (unsigned char)seq[i];
        if (c >= 4) {
            unsigned char u = c & 0xdf; int v = 4;
            for (int k = 0; k < 4; ++k) if (u == tab[k]) v = k;
            if (c == '-') v = 5;
            seq[i] = (char)v;
        }
    }
    int64_t off[2] = { 0, l_seq };
    mem_alnreg_v regs; regs.n = regs.m = 0; regs.a = nullptr;
    int32_t n = 0;
    int64_t cap = 64;
    std::lock_guard<std::mutex> lk(ctx->call_mu);
    for (;;) {
        regs.a = (mem_alnreg_t *)realloc(regs.a, sizeof(mem_alnreg_t) * (size_t)cap);
        int64_t tot = b200_align_batch(ctx, opt, 1, (const uint8_t *)seq, off, cap, regs.a, &n);
        if (tot >= 0) { regs.n = (int)tot; regs.m = (int)cap; break; }
        if (B200_IS_ERROR(tot)) die_on(-1, "mem_align1_core");
        cap = -tot;
    }
    return regs;
}
