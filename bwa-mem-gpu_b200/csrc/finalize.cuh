// SURVEY.md 8(f) rank 2, single-end half: what worker2 (bwamem.c:1195-1205) does to a read's regions before any SAM
// text exists -- mem_mark_primary_se (bwamem.c:500-565), mem_reorder_primary5 (:988-1010), the record selection, flags
// and MAPQ cap of mem_reg2sam (:1013-1058), mem_approx_mapq_se (:962-986) and the choice of XA entries of mem_gen_alt
// (bwamem_extra.c:116-170).  One read per lane; a read has one or two regions almost always, so everything is done in
// place on the read's slice of the region array with no scratch.
//
// Exactness: hash_64 is a bijection, so both sort orders (alnreg_hlt, alnreg_hlt2) are strict total orders over a
// read's regions and any sorting algorithm gives the reference's permutation.  MAPQ is double arithmetic around
// log(integer): the logarithms come from a table the host fills with its own libm (identical bits to the reference
// on the same box), every other operation is a single correctly-rounded IEEE operation written with the _rn
// intrinsics so that no multiply-add is contracted.
#pragma once
#include "ctx.h"

#ifndef MEM_F_ALL
#define MEM_F_ALL 0x8
#endif
#ifndef MEM_F_NO_MULTI
#define MEM_F_NO_MULTI 0x10
#endif
#ifndef MEM_F_PRIMARY5
#define MEM_F_PRIMARY5 0x800
#endif
#ifndef MEM_F_KEEP_SUPP_MAPQ
#define MEM_F_KEEP_SUPP_MAPQ 0x1000
#endif

struct FinParams {
    mem_opt_t opt;
    mem_alnreg_t *regs;                  // the chunk's regions
    const i64 *roff;                     // n_reads + 1 region offsets (global numbering); the chunk starts at g0
    i64 g0; int r0, n_reads;             // reads r0 .. r0 + n_reads of the call
    i64 id0;                             // n_processed
    b200_sel_t *sel;                     // per region of the chunk
    const double *logtab; int log_cap;   // log(i), i < log_cap
    int *overflow;
};

__host__ __device__ __forceinline__ u64 fin_hash_64(u64 key)       // utils.h:97-108
{
    key += ~(key << 32); key ^= (key >> 22);
    key += ~(key << 13); key ^= (key >> 8);
    key += (key << 3);   key ^= (key >> 15);
    key += ~(key << 27); key ^= (key >> 31);
    return key;
}

#ifdef B200_EMU
static inline double fin_dmul(double a, double b) { volatile double r = a * b; return r; }
static inline double fin_dadd(double a, double b) { volatile double r = a + b; return r; }
static inline double fin_ddiv(double a, double b) { volatile double r = a / b; return r; }
static inline float fin_fmul(float a, float b) { volatile float r = a * b; return r; }
#else
__device__ __forceinline__ double fin_dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double fin_dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double fin_ddiv(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ float fin_fmul(float a, float b) { return __fmul_rn(a, b); }
#endif

__device__ __forceinline__ bool fin_hlt(const mem_alnreg_t &a, const mem_alnreg_t &b)    // alnreg_hlt, bwamem.c:404
{
    return a.score > b.score || (a.score == b.score && (a.is_alt < b.is_alt || (a.is_alt == b.is_alt && a.hash < b.hash)));
}
__device__ __forceinline__ bool fin_hlt2(const mem_alnreg_t &a, const mem_alnreg_t &b)   // alnreg_hlt2, bwamem.c:407
{
    return a.is_alt < b.is_alt || (a.is_alt == b.is_alt && (a.score > b.score || (a.score == b.score && a.hash < b.hash)));
}

template <int WHICH>
__device__ __forceinline__ bool fin_less(const mem_alnreg_t &a, const mem_alnreg_t &b) { return WHICH ? fin_hlt2(a, b) : fin_hlt(a, b); }

// in-place sort of a strict total order: insertion for the usual handful of regions, heap sort beyond
template <int WHICH>
__device__ void fin_sort(mem_alnreg_t *a, int n)
{
    if (n < 2) return;
    if (n <= 8) {
        for (int i = 1; i < n; ++i) {
            if (!fin_less<WHICH>(a[i], a[i - 1])) continue;
            const mem_alnreg_t t = a[i];
            int j = i;
            for (; j > 0 && fin_less<WHICH>(t, a[j - 1]); --j) a[j] = a[j - 1];
            a[j] = t;
        }
        return;
    }
    auto sift = [&](int i, int m) {
        const mem_alnreg_t t = a[i];
        for (;;) {
            int k = 2 * i + 1;
            if (k >= m) break;
            if (k + 1 < m && fin_less<WHICH>(a[k], a[k + 1])) ++k;
            if (!fin_less<WHICH>(t, a[k])) break;
            a[i] = a[k]; i = k;
        }
        a[i] = t;
    };
    for (int i = n / 2 - 1; i >= 0; --i) sift(i, n);
    for (int m = n - 1; m > 0; --m) {
        const mem_alnreg_t t = a[0]; a[0] = a[m]; a[m] = t;
        sift(0, m);
    }
}

// mem_mark_primary_se_core (bwamem.c:500-527).  The reference's list z holds, in increasing order, exactly the earlier
// regions that were not made secondary; they are recognised here by secondary < 0.
__device__ void fin_mark_core(const mem_opt_t &opt, int n, mem_alnreg_t *a)
{
    int tmp = opt.a + opt.b;
    tmp = opt.o_del + opt.e_del > tmp ? opt.o_del + opt.e_del : tmp;
    tmp = opt.o_ins + opt.e_ins > tmp ? opt.o_ins + opt.e_ins : tmp;
    for (int i = 1; i < n; ++i) {
        const int qbi = a[i].qb, qei = a[i].qe, sci = a[i].score, alti = a[i].is_alt;
        for (int j = 0; j < i; ++j) {
            if (a[j].secondary >= 0) continue;
            const int qbj = a[j].qb, qej = a[j].qe;
            const int b_max = qbj > qbi ? qbj : qbi, e_min = qej < qei ? qej : qei;
            if (e_min > b_max) {
                const int min_l = qei - qbi < qej - qbj ? qei - qbi : qej - qbj;
                if ((float)(e_min - b_max) >= fin_fmul((float)min_l, opt.mask_level)) {
                    if (a[j].sub == 0) a[j].sub = sci;
                    if (a[j].score - sci <= tmp && (a[j].is_alt || !alti)) ++a[j].sub_n;
                    a[i].secondary = j;
                    break;
                }
            }
        }
    }
}

__device__ int fin_mark_primary_se(const mem_opt_t &opt, int n, mem_alnreg_t *a, i64 id)     // bwamem.c:529-565
{
    int n_pri = 0;
    for (int i = 0; i < n; ++i) {
        a[i].sub = 0; a[i].alt_sc = 0; a[i].secondary = -1; a[i].secondary_all = -1; a[i].hash = fin_hash_64((u64)(id + i));
        if (!a[i].is_alt) ++n_pri;
    }
    fin_sort<0>(a, n);
    fin_mark_core(opt, n, a);
    for (int i = 0; i < n; ++i) {
        a[i].secondary_all = i;
        if (!a[i].is_alt && a[i].secondary >= 0 && a[a[i].secondary].is_alt) a[i].alt_sc = a[a[i].secondary].score;
    }
    if (n_pri < n) {
        if (n_pri > 0) fin_sort<1>(a, n);
        for (int i = 0; i < n; ++i) {                  // parent's rank in the first order -> its index now
            const int s = a[i].secondary;
            if (s < 0) continue;
            int k = 0;
            while (a[k].secondary_all != s) ++k;
            a[i].secondary = k;
        }
        for (int i = 0; i < n; ++i) {
            if (a[i].secondary >= 0) {
                a[i].secondary_all = a[i].secondary;
                if (a[i].is_alt) a[i].secondary = 0x7fffffff;
            } else a[i].secondary_all = -1;
        }
        if (n_pri > 0) {
            for (int i = 0; i < n_pri; ++i) { a[i].sub = 0; a[i].secondary = -1; }
            fin_mark_core(opt, n_pri, a);
        }
    } else {
        for (int i = 0; i < n; ++i) a[i].secondary_all = a[i].secondary;
    }
    return n_pri;
}

__device__ void fin_reorder_primary5(int T, int n, mem_alnreg_t *a)       // bwamem.c:988-1010
{
    int n_pri = 0, left_st = 0x7fffffff, left_k = -1;
    for (int k = 0; k < n; ++k) if (a[k].secondary < 0 && !a[k].is_alt && a[k].score >= T) ++n_pri;
    if (n_pri <= 1) return;
    for (int k = 0; k < n; ++k) {
        if (a[k].secondary >= 0 || a[k].is_alt || a[k].score < T) continue;
        if (a[k].qb < left_st) { left_st = a[k].qb; left_k = k; }
    }
    if (left_k == 0) return;
    const mem_alnreg_t t = a[0]; a[0] = a[left_k]; a[left_k] = t;
    for (int k = 1; k < n; ++k) {
        if (a[k].secondary == 0) a[k].secondary = left_k;
        else if (a[k].secondary == left_k) a[k].secondary = 0;
        if (a[k].secondary_all == 0) a[k].secondary_all = left_k;
        else if (a[k].secondary_all == left_k) a[k].secondary_all = 0;
    }
}

__device__ __forceinline__ double fin_log(const FinParams &P, int v, int *ovf)
{
    if (v < 0 || v >= P.log_cap) { *ovf = 3; return 0.; }
    return P.logtab[v];
}

__device__ int fin_approx_mapq_se(const FinParams &P, const mem_alnreg_t &a, int *ovf)      // bwamem.c:962-986
{
    const mem_opt_t &opt = P.opt;
    int mapq, sub = a.sub ? a.sub : opt.min_seed_len * opt.a;
    sub = a.csub > sub ? a.csub : sub;
    if (sub >= a.score) return 0;
    const int l = a.qe - a.qb > (int)(a.re - a.rb) ? a.qe - a.qb : (int)(a.re - a.rb);
    const double identity = fin_dadd(1., -fin_ddiv(fin_ddiv((double)(l * opt.a - a.score), (double)(opt.a + opt.b)), (double)l));
    if (a.score == 0) mapq = 0;
    else if (opt.mapQ_coef_len > 0) {
        double tmp = (float)l < opt.mapQ_coef_len ? 1. : fin_ddiv((double)opt.mapQ_coef_fac, fin_log(P, l, ovf));
        tmp = fin_dmul(tmp, fin_dmul(identity, identity));
        double v = fin_dmul(6.02, (double)(a.score - sub));      // 6.02 * (score - sub) / a * tmp * tmp + .499, left to right
        v = fin_ddiv(v, (double)opt.a);
        v = fin_dmul(fin_dmul(v, tmp), tmp);
        mapq = (int)fin_dadd(v, .499);
    } else {
        double v = fin_dmul(30., fin_dadd(1., -fin_ddiv((double)sub, (double)a.score)));      // MEM_MAPQ_COEF 30.0, bwa.h:68
        v = fin_dmul(v, fin_log(P, a.seedcov, ovf));
        mapq = (int)fin_dadd(v, .499);
        if (identity < 0.95) mapq = (int)fin_dadd(fin_dmul(fin_dmul((double)mapq, identity), identity), .499);
    }
    if (a.sub_n > 0) mapq -= (int)fin_dadd(fin_dmul(4.343, fin_log(P, a.sub_n + 1, ovf)), .499);
    if (mapq > 60) mapq = 60;
    if (mapq < 0) mapq = 0;
    mapq = (int)fin_dadd(fin_dmul((double)mapq, fin_dadd(1., -(double)a.frac_rep)), .499);
    return mapq;
}

__device__ __forceinline__ int fin_pri_idx(double xa_ratio, const mem_alnreg_t *a, int i)     // get_pri_idx, bwamem_extra.c:116-121
{
    const int k = a[i].secondary_all;
    if (k >= 0 && (double)a[i].score >= fin_dmul((double)a[k].score, xa_ratio)) return k;
    return -1;
}

// mem_gen_alt (bwamem_extra.c:124-170): which regions become XA entries, and of which record
__device__ void fin_mark_xa(const mem_opt_t &opt, int n, const mem_alnreg_t *a, b200_sel_t *sel)
{
    const double xr = (double)opt.XA_drop_ratio;
    for (int i = 0; i < n; ++i) {
        const int pr = fin_pri_idx(xr, a, i);
        if (pr < 0) continue;
        int cnt = 0, has_alt = 0;
        for (int j = 0; j < n; ++j)
            if (fin_pri_idx(xr, a, j) == pr) { ++cnt; if (a[j].is_alt) has_alt = 1; }
        if (cnt > opt.max_XA_hits_alt || (!has_alt && cnt > opt.max_XA_hits)) continue;
        sel[i].kind |= 2; sel[i].xa_of = pr;           // kind 3: a record of its own as well (a primary-assembly hit under an ALT hit)
    }
}

// mem_reg2sam's walk over one read's regions (bwamem.c:1024-1042): which become records, their MAPQ and flags
__device__ void fin_reg2sam_walk(const FinParams &P, int n, const mem_alnreg_t *a, b200_sel_t *sel, int extra_flag, int *ovf)
{
    const mem_opt_t &opt = P.opt;
    int l = 0, mapq0 = 0;
    for (int k = 0; k < n; ++k) {
        const mem_alnreg_t &p = a[k];
        b200_sel_t s; s.kind = 0; s.mapq = 0; s.xa_of = -1; s.flag = 0;
        bool keep = p.score >= opt.T;
        if (keep && p.secondary >= 0 && (p.is_alt || !(opt.flag & MEM_F_ALL))) keep = false;
        if (keep && p.secondary >= 0 && p.secondary < 0x7fffffff &&
            (float)p.score < fin_fmul((float)a[p.secondary].score, opt.drop_ratio)) keep = false;
        if (keep) {
            s.kind = 1;
            s.flag = extra_flag;
            s.mapq = p.secondary < 0 ? fin_approx_mapq_se(P, p, ovf) : 0;                    // mem_reg2aln, bwamem.c:1150
            if (l && p.secondary < 0) s.flag |= (opt.flag & MEM_F_NO_MULTI) ? 0x10000 : 0x800;
            if (!(opt.flag & MEM_F_KEEP_SUPP_MAPQ) && l && !p.is_alt && s.mapq > mapq0) s.mapq = mapq0;
            if (l == 0) mapq0 = s.mapq;
            ++l;
        }
        sel[k] = s;
    }
    if (!(opt.flag & MEM_F_ALL)) fin_mark_xa(opt, n, a, sel);
}

__global__ void __launch_bounds__(128) finalize_se_kernel(FinParams P)
{
    const mem_opt_t &opt = P.opt;
    int ovf = 0;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < P.n_reads; r += gridDim.x * blockDim.x) {
        const i64 b = P.roff[P.r0 + r] - P.g0;
        const int n = (int)(P.roff[P.r0 + r + 1] - P.roff[P.r0 + r]);
        if (n == 0) continue;
        mem_alnreg_t *a = P.regs + b;
        b200_sel_t *sel = P.sel + b;
        fin_mark_primary_se(opt, n, a, P.id0 + P.r0 + r);
        if (opt.flag & MEM_F_PRIMARY5) fin_reorder_primary5(opt.T, n, a);
        fin_reg2sam_walk(P, n, a, sel, 0, &ovf);
    }
    if (ovf) atomicMax(P.overflow, ovf);
}
