// Device functions shared by the extension kernels: cal_max_gap (bwamem.c:628-635), the score-only global
// alignment used by mem_patch_reg (ksw_global2 ksw.c:504-606, bwa_gen_cigar2 bwa.c:261-347, mem_patch_reg
// bwamem.c:413-442) and mem_sort_dedup_patch (bwamem.c:444-496).
#pragma once
#include "b200_internal.h"
#include "ksort.cuh"

// cal_max_gap (bwamem.c:628-635)
__device__ __forceinline__ int cal_max_gap(const mem_opt_t &opt, int qlen)
{
    int l_del = (int)((double)(qlen * opt.a - opt.o_del) / opt.e_del + 1.);
    int l_ins = (int)((double)(qlen * opt.a - opt.o_ins) / opt.e_ins + 1.);
    int l = l_del > l_ins ? l_del : l_ins;
    l = l > 1 ? l : 1;
    return l < opt.w << 1 ? l : opt.w << 1;
}

struct RegEndLess { __device__ __forceinline__ bool operator()(const mem_alnreg_t &a, const mem_alnreg_t &b) const { return a.re < b.re; } };   // alnreg_slt2
struct RegScoreLess {                                                                                                                       // alnreg_slt
    __device__ __forceinline__ bool operator()(const mem_alnreg_t &a, const mem_alnreg_t &b) const
    { return a.score > b.score || (a.score == b.score && (a.rb < b.rb || (a.rb == b.rb && a.qb < b.qb))); }
};
struct U64Less { __device__ __forceinline__ bool operator()(u64 a, u64 b) const { return a < b; } };

// ksw_global2 (ksw.c:504-606), score only, one lane.  Query/target through accessors.
template <class QF, class TF>
__device__ inline int ksw_global_score(int qlen, QF qf, int tlen, TF tf, const int8_t *mat, int o_del, int e_del,
                                       int o_ins, int e_ins, int w, i32 *H, i32 *E)
{
    const int MINUS_INF = -0x40000000;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins;
    int j;
    H[0] = 0; E[0] = MINUS_INF;
    for (j = 1; j <= qlen && j <= w; ++j) { H[j] = -(o_ins + e_ins * j); E[j] = MINUS_INF; }
    for (; j <= qlen; ++j) H[j] = E[j] = MINUS_INF;
    for (int i = 0; i < tlen; ++i) {
        int f = MINUS_INF, h1, t;
        const int8_t *mrow = mat + tf(i) * 5;
        int beg = i > w ? i - w : 0;
        int end = i + w + 1 < qlen ? i + w + 1 : qlen;
        h1 = beg == 0 ? -(o_del + e_del * (i + 1)) : MINUS_INF;
        for (j = beg; j < end; ++j) {
            int m = H[j], e = E[j], h;
            H[j] = h1;
            m += mrow[qf(j)];
            h = m >= e ? m : e;
            h = h >= f ? h : f;
            h1 = h;
            t = m - oe_del; e -= e_del; e = e > t ? e : t; E[j] = e;
            t = m - oe_ins; f -= e_ins; f = f > t ? f : t;
        }
        H[end] = h1; E[end] = MINUS_INF;
    }
    return H[qlen];
}

struct SeqFwd { const u8 *q; __device__ __forceinline__ int operator()(int j) const { return q[j]; } };
struct SeqRev { const u8 *q; int n; __device__ __forceinline__ int operator()(int j) const { return q[n - 1 - j]; } };
struct RefSeqFwd { const DevIndex *ix; i64 rb; __device__ __forceinline__ int operator()(int i) const { return dev_ref_base(*ix, rb + i); } };
struct RefSeqRev { const DevIndex *ix; i64 rb; i64 rlen; __device__ __forceinline__ int operator()(int i) const { return dev_ref_base(*ix, rb + rlen - 1 - i); } };

// bwa_gen_cigar2 (bwa.c:261-347) restricted to the score (n_cigar == NULL), as used by mem_patch_reg
__device__ inline int gen_cigar_score(const DevIndex &ix, const mem_opt_t &opt, int w_, int l_query, const u8 *query,
                                      i64 rb, i64 re, i32 *H, i32 *E, int *ok)
{
    const i64 l_pac = ix.l_pac;
    *ok = 0;
    if (l_query <= 0 || rb >= re || (rb < l_pac && re > l_pac)) return 0;
    i64 rlen = re - rb;
    *ok = 1;
    const int8_t *mat = opt.mat;
    if (l_query == re - rb && w_ == 0) {
        int score = 0;
        for (int i = 0; i < l_query; ++i) {
            int r = rb >= l_pac ? dev_ref_base(ix, rb + rlen - 1 - i) : dev_ref_base(ix, rb + i);
            int q = rb >= l_pac ? query[l_query - 1 - i] : query[i];
            score += mat[r * 5 + q];
        }
        return score;
    }
    int max_ins = (int)((double)(((l_query + 1) >> 1) * mat[0] - opt.o_ins) / opt.e_ins + 1.);
    int max_del = (int)((double)(((l_query + 1) >> 1) * mat[0] - opt.o_del) / opt.e_del + 1.);
    int max_gap = max_ins > max_del ? max_ins : max_del;
    max_gap = max_gap > 1 ? max_gap : 1;
    int dl = (int)rlen - l_query; dl = dl < 0 ? -dl : dl;
    int w = (max_gap + dl + 1) >> 1;
    w = w < w_ ? w : w_;
    int min_w = dl + 3;
    w = w > min_w ? w : min_w;
    if (rb >= l_pac)
        return ksw_global_score(l_query, SeqRev{query, l_query}, (int)rlen, RefSeqRev{&ix, rb, rlen}, mat,
                                opt.o_del, opt.e_del, opt.o_ins, opt.e_ins, w, H, E);
    return ksw_global_score(l_query, SeqFwd{query}, (int)rlen, RefSeqFwd{&ix, rb}, mat,
                            opt.o_del, opt.e_del, opt.o_ins, opt.e_ins, w, H, E);
}

// mem_patch_reg (bwamem.c:413-442)
__device__ inline int patch_reg(const DevIndex &ix, const mem_opt_t &opt, const u8 *query, const mem_alnreg_t *a,
                                const mem_alnreg_t *b, int *_w, i32 *H, i32 *E)
{
    int w, score, q_s, r_s, ok;
    double r;
    if (query == nullptr) return 0;                                       // bwamem.c:417 (callers without a query: mem_matesw)
    if (a->rb < ix.l_pac && b->rb >= ix.l_pac) return 0;
    if (a->qb >= b->qb || a->qe >= b->qe || a->re >= b->re) return 0;
    w = (int)((a->re - b->rb) - (a->qe - b->qb));
    w = w > 0 ? w : -w;
    r = (double)(a->re - b->rb) / (b->re - a->rb) - (double)(a->qe - b->qb) / (b->qe - a->qb);
    r = r > 0. ? r : -r;
    if (a->re < b->rb || a->qe < b->qb) {
        if (w > opt.w << 1 || r >= 0.05f) return 0;                       // PATCH_MAX_R_BW
    } else if (w > opt.w << 2 || r >= 0.05f * 2) return 0;
    w += a->w + b->w;
    w = w < opt.w << 2 ? w : opt.w << 2;
    score = gen_cigar_score(ix, opt, w, b->qe - a->qb, query + a->qb, a->rb, b->re, H, E, &ok);
    q_s = (int)((double)(b->qe - a->qb) / ((b->qe - b->qb) + (a->qe - a->qb)) * (b->score + a->score) + .499);
    r_s = (int)((double)(b->re - a->rb) / ((b->re - b->rb) + (a->re - a->rb)) * (b->score + a->score) + .499);
    if ((double)score / (q_s > r_s ? q_s : r_s) < 0.90f) return 0;         // PATCH_MIN_SC_RATIO
    *_w = w;
    return score;
}

// mem_sort_dedup_patch (bwamem.c:444-496), one lane
__device__ inline int sort_dedup_patch(const DevIndex &ix, const mem_opt_t &opt, const u8 *query, int n, mem_alnreg_t *a,
                                       i32 *H, i32 *E)
{
    int m, i, j;
    if (n <= 1) return n;
    ks_introsort(a, n, RegEndLess());
    for (i = 0; i < n; ++i) a[i].n_comp = 1;
    for (i = 1; i < n; ++i) {
        mem_alnreg_t *p = &a[i];
        if (p->rid != a[i - 1].rid || p->rb >= a[i - 1].re + opt.max_chain_gap) continue;
        for (j = i - 1; j >= 0 && p->rid == a[j].rid && p->rb < a[j].re + opt.max_chain_gap; --j) {
            mem_alnreg_t *q = &a[j];
            i64 orr, oq, mr, mq;
            int score, w;
            if (q->qe == q->qb) continue;
            orr = q->re - p->rb;
            oq = q->qb < p->qb ? q->qe - p->qb : p->qe - q->qb;
            mr = q->re - q->rb < p->re - p->rb ? q->re - q->rb : p->re - p->rb;
            mq = q->qe - q->qb < p->qe - p->qb ? q->qe - q->qb : p->qe - p->qb;
            if ((float)orr > opt.mask_level_redun * (float)mr && (float)oq > opt.mask_level_redun * (float)mq) {
                if (p->score < q->score) { p->qe = p->qb; break; }
                else q->qe = q->qb;
            } else if (q->rb < p->rb && (score = patch_reg(ix, opt, query, q, p, &w, H, E)) > 0) {
                p->n_comp += q->n_comp + 1;
                p->seedcov = p->seedcov > q->seedcov ? p->seedcov : q->seedcov;
                p->sub = p->sub > q->sub ? p->sub : q->sub;
                p->csub = p->csub > q->csub ? p->csub : q->csub;
                p->qb = q->qb; p->rb = q->rb;
                p->truesc = p->score = score;
                p->w = w;
                q->qb = q->qe;
            }
        }
    }
    for (i = 0, m = 0; i < n; ++i)
        if (a[i].qe > a[i].qb) { if (m != i) a[m++] = a[i]; else ++m; }
    n = m;
    ks_introsort(a, n, RegScoreLess());
    for (i = 1; i < n; ++i)
        if (a[i].score == a[i - 1].score && a[i].rb == a[i - 1].rb && a[i].qb == a[i - 1].qb) a[i].qe = a[i].qb;
    for (i = 1, m = 1; i < n; ++i)
        if (a[i].qe > a[i].qb) { if (m != i) a[m++] = a[i]; else ++m; }
    return m;
}

