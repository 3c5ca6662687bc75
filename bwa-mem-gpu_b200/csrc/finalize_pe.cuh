// SURVEY.md 8(f) rank 2, paired-end half: what mem_sam_pe (bwamem_pair.c:276-419) does to a pair's regions AFTER mate rescue
// and before any SAM text exists -- mem_mark_primary_se x 2 (:303-304), mem_reorder_primary5 (:305-308), mem_pair
// (:208-269), the MAPQ arithmetic of the paired branch (:311-345), the secondary_all switch (:346-355), the choice of the
// records h / g and their flags (:361-377), and the no_pairing branch (:395-416: the records mem_reg2sam would write with
// the pair flags, the proper-pair test on the top hits).  One pair per lane, in place on the pair's slice of the region
// array (the two reads' lists are adjacent).  Also the per-pair summary mem_pestat (:72-135) needs.
//
// Exactness.  hash_64 makes every sort key unique, so any sorting algorithm reproduces ks_introsort_128's result.  The
// one transcendental expression, .721 * log(2 * erfc(|ns| / sqrt 2)) * a with ns = (dist - avg) / std (:243-244), depends
// only on the orientation and the INTEGER insert size dist in [low, high]: the host fills a table per orientation with
// its own libm (the same bits the reference computes on this box) and the kernel does the two remaining additions as
// single correctly rounded operations.  log(n_sub + 1) comes from the integer log table of finalize.cuh.
#pragma once
#include "finalize.cuh"

#ifndef MEM_F_NOPAIRING
#define MEM_F_NOPAIRING 0x4
#endif

struct FinPeParams {
    FinParams F;                         // F.r0 / F.n_reads: reads of the chunk (both even); F.id0 = n_processed of the batch (even)
    mem_pestat_t pes[4];
    const double *ptab[4]; int ptab_n[4];    // ptab[d][dist - pes[d].low] = .721 * log(2 * erfc(|ns| * M_SQRT1_2)) * opt.a
    i32 *mate_reg;                       // per read of the batch: chunk-relative region of the MATE's record h, -1 = mate unmapped
    i32 *read_flag;                      // per read of the batch: extra flag of its unmapped record
    u64 *vscr;                           // 2 x u64 per region of the chunk: mem_pair's array v
    i64 l_pac; const i64 *ann_off;
};

__device__ __forceinline__ int fin_infer_dir(i64 l_pac, i64 b1, i64 b2, i64 *dist)       // mem_infer_dir, bwamem_pair.c:49-56
{
    const int r1 = b1 >= l_pac, r2 = b2 >= l_pac;
    const i64 p2 = r1 == r2 ? b2 : (l_pac << 1) - 1 - b2;
    *dist = p2 > b1 ? p2 - b1 : b1 - p2;
    return (r1 == r2 ? 0 : 1) ^ (p2 > b1 ? 0 : 3);
}

__device__ __forceinline__ bool fin_p64_lt(u64 ax, u64 ay, u64 bx, u64 by) { return ax < bx || (ax == bx && ay < by); }   // pair64_lt, utils.h

__device__ void fin_sort_p64(u64 *v, int n)              // ascending by (x, y); keys are unique
{
    if (n <= 12) {
        for (int i = 1; i < n; ++i) {
            const u64 tx = v[2 * i], ty = v[2 * i + 1];
            int j = i;
            for (; j > 0 && fin_p64_lt(tx, ty, v[2 * j - 2], v[2 * j - 1]); --j) { v[2 * j] = v[2 * j - 2]; v[2 * j + 1] = v[2 * j - 1]; }
            v[2 * j] = tx; v[2 * j + 1] = ty;
        }
        return;
    }
    auto sift = [&](int i, int m) {
        const u64 tx = v[2 * i], ty = v[2 * i + 1];
        for (;;) {
            int k = 2 * i + 1;
            if (k >= m) break;
            if (k + 1 < m && fin_p64_lt(v[2 * k], v[2 * k + 1], v[2 * k + 2], v[2 * k + 3])) ++k;
            if (!fin_p64_lt(tx, ty, v[2 * k], v[2 * k + 1])) break;
            v[2 * i] = v[2 * k]; v[2 * i + 1] = v[2 * k + 1]; i = k;
        }
        v[2 * i] = tx; v[2 * i + 1] = ty;
    };
    for (int i = n / 2 - 1; i >= 0; --i) sift(i, n);
    for (int m = n - 1; m > 0; --m) {
        const u64 tx = v[0], ty = v[1];
        v[0] = v[2 * m]; v[1] = v[2 * m + 1]; v[2 * m] = tx; v[2 * m + 1] = ty;
        sift(0, m);
    }
}

// The candidate pairs of mem_pair's double loop (bwamem_pair.c:226-255), generated in the reference's order and handed to
// `f(x, y)`; u is never stored: the caller keeps what it needs of it (two passes).
template <class F>
__device__ __forceinline__ void fin_pair_scan(const FinPeParams &P, const u64 *v, int nv, i64 id, F f)
{
    int y[4] = { -1, -1, -1, -1 };
    const int id32 = (int)id;                             // mem_pair takes `int id`
    const u64 idx = (u64)(i64)(int)((unsigned)id32 << 8); // id<<8 as int, sign-extended into the xor
    for (int i = 0; i < nv; ++i) {
        const u64 vy = v[2 * i + 1];
        for (int r = 0; r < 2; ++r) {
            const int dir = r << 1 | (int)(vy >> 1 & 1);
            if (P.pes[dir].failed) continue;
            const int which = r << 1 | (int)((vy & 1) ^ 1);
            if (y[which] < 0) continue;
            for (int k = y[which]; k >= 0; --k) {
                if ((int)(v[2 * k + 1] & 3) != which) continue;
                const i64 dist = (i64)v[2 * i] - (i64)v[2 * k];
                if (dist > P.pes[dir].high) break;
                if (dist < P.pes[dir].low) continue;
                const u64 ssum = (vy >> 32) + (v[2 * k + 1] >> 32);
                const i64 ti = dist - P.pes[dir].low;
                const double t = ti < P.ptab_n[dir] ? P.ptab[dir][ti] : 0.;
                int q = (int)fin_dadd(fin_dadd((double)ssum, t), .499);
                if (q < 0) q = 0;
                const u64 py = (u64)k << 32 | (u64)i;
                const u64 px = (u64)q << 32 | (fin_hash_64(py ^ idx) & 0xffffffffull);
                f(px, py);
            }
        }
        y[vy & 3] = i;
    }
}

__device__ __forceinline__ int fin_raw_mapq(int diff, int a)     // raw_mapq, bwamem_pair.c:274
{
    return (int)fin_dadd(fin_ddiv(fin_dmul(6.02, (double)diff), (double)a), .499);
}

__global__ void __launch_bounds__(128) finalize_pe_kernel(FinPeParams P)
{
    const FinParams &F = P.F;
    const mem_opt_t &opt = F.opt;
    int ovf = 0;
    const int n_pairs = F.n_reads >> 1;
    for (int pp = blockIdx.x * blockDim.x + threadIdx.x; pp < n_pairs; pp += gridDim.x * blockDim.x) {
        const int r0 = F.r0 + 2 * pp;                     // reads r0, r0 + 1 of the batch
        const i64 b0 = F.roff[r0] - F.g0, b1 = F.roff[r0 + 1] - F.g0;
        const int n0 = (int)(F.roff[r0 + 1] - F.roff[r0]), n1 = (int)(F.roff[r0 + 2] - F.roff[r0 + 1]);
        mem_alnreg_t *a[2] = { F.regs + b0, F.regs + b1 };
        b200_sel_t *sel[2] = { F.sel + b0, F.sel + b1 };
        const i64 bb[2] = { b0, b1 };
        const int nn[2] = { n0, n1 };
        const i64 id = (F.id0 >> 1) + (r0 >> 1);          // bwamem.c:1210: (n_processed >> 1) + i
        int n_pri[2];
        n_pri[0] = fin_mark_primary_se(opt, n0, a[0], id << 1 | 0);
        n_pri[1] = fin_mark_primary_se(opt, n1, a[1], id << 1 | 1);
        if (opt.flag & MEM_F_PRIMARY5) { fin_reorder_primary5(opt.T, n0, a[0]); fin_reorder_primary5(opt.T, n1, a[1]); }
        for (int i = 0; i < 2; ++i)
            for (int k = 0; k < nn[i]; ++k) { b200_sel_t s; s.kind = 0; s.mapq = 0; s.xa_of = -1; s.flag = 0; sel[i][k] = s; }
        int extra_flag = 1;
        bool paired = false;
        int z[2] = { 0, 0 };
        if (!(opt.flag & MEM_F_NOPAIRING) && n_pri[0] && n_pri[1]) {
            // ---- mem_pair (bwamem_pair.c:208-269)
            u64 *v = P.vscr + 2 * b0;
            int nv = 0;
            for (int r = 0; r < 2; ++r)
                for (int i = 0; i < n_pri[r]; ++i) {
                    const mem_alnreg_t &e = a[r][i];
                    u64 kx = (u64)(e.rb < P.l_pac ? e.rb : (P.l_pac << 1) - 1 - e.rb);
                    kx = (u64)e.rid << 32 | (kx - (u64)P.ann_off[e.rid]);
                    v[2 * nv] = kx;
                    v[2 * nv + 1] = (u64)e.score << 32 | (u64)(i << 2) | (u64)(e.rb >= P.l_pac) << 1 | (u64)r;
                    ++nv;
                }
            fin_sort_p64(v, nv);
            u64 bx = 0, by = 0, sx = 0, sy = 0; i64 nu = 0;       // best and second best of u in (x, y) order
            fin_pair_scan(P, v, nv, id, [&](u64 px, u64 py) {
                if (nu == 0 || fin_p64_lt(bx, by, px, py)) { sx = bx; sy = by; bx = px; by = py; }
                else if (nu == 1 || fin_p64_lt(sx, sy, px, py)) { sx = px; sy = py; }
                ++nu;
            });
            int o = 0, subo = 0, n_sub = 0;
            if (nu) {
                int tmp = opt.a + opt.b;
                tmp = tmp > opt.o_del + opt.e_del ? tmp : opt.o_del + opt.e_del;
                tmp = tmp > opt.o_ins + opt.e_ins ? tmp : opt.o_ins + opt.e_ins;
                const int i_ = (int)(by >> 32), k_ = (int)(by & 0xffffffffull);
                z[v[2 * i_ + 1] & 1] = (int)((v[2 * i_ + 1] << 32) >> 34);
                z[v[2 * k_ + 1] & 1] = (int)((v[2 * k_ + 1] << 32) >> 34);
                o = (int)(bx >> 32);
                subo = nu > 1 ? (int)(sx >> 32) : 0;
                i64 cnt = 0;
                fin_pair_scan(P, v, nv, id, [&](u64 px, u64) { if (subo - (int)(px >> 32) <= tmp) ++cnt; });
                n_sub = (int)(cnt - 1);                       // every entry but the best one (which always qualifies)
            }
            if (o > 0) {
                bool multi = false;
                for (int i = 0; i < 2 && !multi; ++i)
                    for (int j = 1; j < n_pri[i]; ++j)
                        if (a[i][j].secondary < 0 && a[i][j].score >= opt.T) { multi = true; break; }
                if (!multi) {
                    paired = true;
                    // ---- MAPQ of the pair (bwamem_pair.c:322-345)
                    const int score_un = a[0][0].score + a[1][0].score - opt.pen_unpaired;
                    subo = subo > score_un ? subo : score_un;
                    int q_pe = fin_raw_mapq(o - subo, opt.a);
                    if (n_sub > 0) q_pe -= (int)fin_dadd(fin_dmul(4.343, fin_log(F, n_sub + 1, &ovf)), .499);
                    if (q_pe < 0) q_pe = 0;
                    if (q_pe > 60) q_pe = 60;
                    {
#ifdef B200_EMU
                        volatile float fs = a[0][0].frac_rep + a[1][0].frac_rep; const float frs = fs;
#else
                        const float frs = __fadd_rn(a[0][0].frac_rep, a[1][0].frac_rep);
#endif
                        q_pe = (int)fin_dadd(fin_dmul((double)q_pe, fin_dadd(1., -fin_dmul(.5, (double)frs))), .499);
                    }
                    int q_se[2];
                    if (o > score_un) {
                        mem_alnreg_t *c[2] = { &a[0][z[0]], &a[1][z[1]] };
                        for (int i = 0; i < 2; ++i) {
                            if (c[i]->secondary >= 0) { c[i]->sub = a[i][c[i]->secondary].score; c[i]->secondary = -2; }
                            q_se[i] = fin_approx_mapq_se(F, *c[i], &ovf);
                        }
                        for (int i = 0; i < 2; ++i) {
                            q_se[i] = q_se[i] > q_pe ? q_se[i] : q_pe < q_se[i] + 40 ? q_pe : q_se[i] + 40;
                        }
                        extra_flag |= 2;
                        for (int i = 0; i < 2; ++i) {
                            const int cap = fin_raw_mapq(c[i]->score - c[i]->csub, opt.a);
                            q_se[i] = q_se[i] < cap ? q_se[i] : cap;
                        }
                    } else {
                        z[0] = z[1] = 0;
                        q_se[0] = fin_approx_mapq_se(F, a[0][0], &ovf);
                        q_se[1] = fin_approx_mapq_se(F, a[1][0], &ovf);
                    }
                    for (int i = 0; i < 2; ++i) {             // switch secondary and primary if both are non-ALT (:346-355)
                        const int k = a[i][z[i]].secondary_all;
                        if (k >= 0 && k < n_pri[i]) {
                            for (int j = 0; j < nn[i]; ++j)
                                if (a[i][j].secondary_all == k || j == k) a[i][j].secondary_all = z[i];
                            a[i][z[i]].secondary_all = -1;
                        }
                    }
                    if (!(opt.flag & MEM_F_ALL)) { fin_mark_xa(opt, n0, a[0], sel[0]); fin_mark_xa(opt, n1, a[1], sel[1]); }
                    for (int i = 0; i < 2; ++i) {             // the records h and g (:361-377)
                        sel[i][z[i]].kind |= 1; sel[i][z[i]].mapq = q_se[i]; sel[i][z[i]].flag = 0x40 << i | extra_flag;
                        if (n_pri[i] < nn[i]) {
                            const mem_alnreg_t &p = a[i][n_pri[i]];
                            if (!(p.score < opt.T || p.secondary >= 0 || !p.is_alt)) {
                                b200_sel_t &s = sel[i][n_pri[i]];
                                s.kind |= 1; s.mapq = fin_approx_mapq_se(F, p, &ovf); s.flag = 0x800 | 0x40 << i | extra_flag;
                            }
                        }
                        P.mate_reg[r0 + (i ^ 1)] = (i32)(bb[i] + z[i]);      // the other read's mate record is this read's h
                        P.read_flag[r0 + i] = 0;
                    }
                }
            }
        }
        if (!paired) {                                        // no_pairing (bwamem_pair.c:395-416)
            int which[2] = { -1, -1 };
            for (int i = 0; i < 2; ++i) {
                if (nn[i]) {
                    if (a[i][0].score >= opt.T) which[i] = 0;
                    else if (n_pri[i] < nn[i] && a[i][n_pri[i]].score >= opt.T) which[i] = n_pri[i];
                }
            }
            extra_flag = 1;
            // h[i].rid: mem_reg2aln of a region is on its contig (rid of the region); unmapped = -1
            const int rid0 = which[0] >= 0 ? a[0][which[0]].rid : -1, rid1 = which[1] >= 0 ? a[1][which[1]].rid : -1;
            if (!(opt.flag & MEM_F_NOPAIRING) && rid0 == rid1 && rid0 >= 0) {
                i64 dist;
                const int d = fin_infer_dir(P.l_pac, a[0][0].rb, a[1][0].rb, &dist);
                if (!P.pes[d].failed && dist >= P.pes[d].low && dist <= P.pes[d].high) extra_flag |= 2;
            }
            for (int i = 0; i < 2; ++i) {
                const int xf = (i ? 0x81 : 0x41) | extra_flag;
                fin_reg2sam_walk(F, nn[i], a[i], sel[i], xf, &ovf);
                if (which[i] >= 0) sel[i][which[i]].kind |= 4;            // its alignment is needed as the mate's record even if it is not printed
                P.mate_reg[r0 + (i ^ 1)] = which[i] >= 0 ? (i32)(bb[i] + which[i]) : -1;
                P.read_flag[r0 + i] = xf;
            }
        }
    }
    if (ovf) atomicMax(F.overflow, ovf);
}

// ---- per-pair summary for mem_pestat (bwamem_pair.c:72-90) on worker1's regions: 0 = the pair does not count, else
// (insert size << 2 | orientation) + 1
__device__ __forceinline__ int fin_cal_sub(const mem_opt_t &opt, int n, const mem_alnreg_t *a)     // cal_sub, bwamem_pair.c:58-70
{
    int j;
    for (j = 1; j < n; ++j) {
        const int b_max = a[j].qb > a[0].qb ? a[j].qb : a[0].qb;
        const int e_min = a[j].qe < a[0].qe ? a[j].qe : a[0].qe;
        if (e_min > b_max) {
            const int min_l = a[j].qe - a[j].qb < a[0].qe - a[0].qb ? a[j].qe - a[j].qb : a[0].qe - a[0].qb;
            if ((float)(e_min - b_max) >= fin_fmul((float)min_l, opt.mask_level)) break;
        }
    }
    return j < n ? a[j].score : opt.min_seed_len * opt.a;
}

__global__ void pestat_pairs_kernel(mem_opt_t opt, i64 l_pac, const mem_alnreg_t *regs, const i64 *roff, int n_pairs, u64 *out)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    const mem_alnreg_t *a0 = regs + roff[2 * p], *a1 = regs + roff[2 * p + 1];
    const int n0 = (int)(roff[2 * p + 1] - roff[2 * p]), n1 = (int)(roff[2 * p + 2] - roff[2 * p + 1]);
    u64 r = 0;
    if (n0 && n1 && !((double)fin_cal_sub(opt, n0, a0) > fin_dmul(0.8, (double)a0[0].score)) &&
        !((double)fin_cal_sub(opt, n1, a1) > fin_dmul(0.8, (double)a1[0].score)) && a0[0].rid == a1[0].rid) {
        i64 is;
        const int dir = fin_infer_dir(l_pac, a0[0].rb, a1[0].rb, &is);
        if (is && is <= opt.max_ins) r = ((u64)is << 2 | (u64)dir) + 1;
    }
    out[p] = r;
}
