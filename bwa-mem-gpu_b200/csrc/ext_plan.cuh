// Planning step of the round-based extension (ext2.cu / ext3.cu): see ext2.cu's header comment.  Shared by the per-round
// planning kernel and by the tail mode of the DP kernels, where a lane plans the next seed of its read itself.
#pragma once
#include "b200_internal.h"
#include "extend_common.cuh"

__device__ __forceinline__ int clamp_band(const mem_opt_t &opt, int qlen, int w, int end_bonus)
{   // ksw.c:399-407 with max = max(mat)
    int mx = 0;
    for (int k = 0; k < 25; ++k) mx = mx > opt.mat[k] ? mx : opt.mat[k];
    int max_ins = (int)((double)(qlen * mx + end_bonus - opt.o_ins) / opt.e_ins + 1.);
    max_ins = max_ins > 1 ? max_ins : 1;
    w = w < max_ins ? w : max_ins;
    int max_del = (int)((double)(qlen * mx + end_bonus - opt.o_del) / opt.e_del + 1.);
    max_del = max_del > 1 ? max_del : 1;
    return w < max_del ? w : max_del;
}

// One planning step of read r: consume the finished job `done` (null on the read's first step), walk on through the
// chains / seeds in the reference's order and hand back the next seed extension in *out.  Returns false when the read is
// finished.  first = the read's first step (no state yet).
static __device__ bool ext_plan_read(const Ext2Params &P, int r, const ExtJob *done, bool first, ExtJob *out)
{
    const mem_opt_t &opt = P.opt;
    const DevIndex &ix = P.ix;
    const i64 l_pac = ix.l_pac;
    const u8 *query = P.seq + P.off[r];
    const int l_query = (int)(P.off[r + 1] - P.off[r]);
    const int nchn = P.chn_n[r];
    const ChainRec *chains = P.chains + P.chn_off[r];
    mem_alnreg_t *av = P.regs + P.chn_off[r];
    ExtRead er;
    if (first) { er.ci = 0; er.k = -2; er.nav = 0; er.rmax0 = er.rmax1 = 0; }
    else er = P.rstate[r];
    (void)query;

    if (done) {                                                            // finish the region of the last job
        const ChainRec c = chains[er.ci];
        const Seed *seeds = P.cseeds + c.seed_off;
        mem_alnreg_t a;
        memset(&a, 0, sizeof(a));
        a.rb = done->rb; a.re = done->re; a.qb = done->qb; a.qe = done->qe;
        a.score = done->score; a.truesc = done->truesc; a.w = done->w; a.rid = c.rid;
        a.seedcov = 0;                                                     // bwamem.c:781-790
        for (int i = 0; i < c.n; ++i) {
            const Seed s = seeds[i];
            if (s.qbeg >= a.qb && s.qbeg + s.len <= a.qe && s.rbeg >= a.rb && s.rbeg + s.len <= a.re) a.seedcov += s.len;
        }
        a.seedlen0 = done->slen;
        a.frac_rep = c.frac_rep;
        av[er.nav++] = a;
        --er.k;
    }

    for (;;) {
        if (er.ci >= nchn) { er.k = -3; break; }                           // read finished
        const ChainRec c = chains[er.ci];
        const Seed *seeds = P.cseeds + c.seed_off;
        u64 *srt = P.srt + c.seed_off;
        if (er.k == -2) {                                                  // new chain
            if (c.n == 0) { ++er.ci; continue; }
            i64 rmax0 = l_pac << 1, rmax1 = 0;                             // bwamem.c:649-664
            for (int i = 0; i < c.n; ++i) {
                const Seed s = seeds[i];
                i64 b = s.rbeg - (s.qbeg + cal_max_gap(opt, s.qbeg));
                i64 e = s.rbeg + s.len + ((l_query - s.qbeg - s.len) + cal_max_gap(opt, l_query - s.qbeg - s.len));
                rmax0 = rmax0 < b ? rmax0 : b;
                rmax1 = rmax1 > e ? rmax1 : e;
            }
            rmax0 = rmax0 > 0 ? rmax0 : 0;
            rmax1 = rmax1 < l_pac << 1 ? rmax1 : l_pac << 1;
            if (rmax0 < l_pac && l_pac < rmax1) {
                if (seeds[0].rbeg < l_pac) rmax1 = l_pac; else rmax0 = l_pac;
            }
            {   // bns_fetch_seq clamp to the contig of seeds[0].rbeg (bntseq.c:432-443)
                int is_rev;
                int rid = dev_pos2rid(ix, dev_depos(ix, seeds[0].rbeg, &is_rev));
                i64 far_beg = ix.ann_off[rid], far_end = far_beg + ix.ann_len[rid];
                if (is_rev) { i64 tmp = far_beg; far_beg = (l_pac << 1) - far_end; far_end = (l_pac << 1) - tmp; }
                rmax0 = rmax0 > far_beg ? rmax0 : far_beg;
                rmax1 = rmax1 < far_end ? rmax1 : far_end;
            }
            er.rmax0 = rmax0; er.rmax1 = rmax1;
            for (int i = 0; i < c.n; ++i) srt[i] = (u64)(u32)seeds[i].score << 32 | (u32)i;   // bwamem.c:669-672 (unique keys)
            ks_introsort(srt, c.n, U64Less());
            er.k = c.n - 1;
        }
        bool emitted = false;
        while (er.k >= 0) {
            const int k = er.k;
            const Seed s = seeds[(u32)srt[k]];
            int i;
            for (i = 0; i < er.nav; ++i) {                                 // bwamem.c:678-694
                const mem_alnreg_t *p = &av[i];
                i64 rd; int qd, w, max_gap;
                if (s.rbeg < p->rb || s.rbeg + s.len > p->re || s.qbeg < p->qb || s.qbeg + s.len > p->qe) continue;
                if (s.len - p->seedlen0 > .1 * l_query) continue;
                qd = s.qbeg - p->qb; rd = s.rbeg - p->rb;
                max_gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
                w = max_gap < p->w ? max_gap : p->w;
                if (qd - rd < w && rd - qd < w) break;
                qd = p->qe - (s.qbeg + s.len); rd = p->re - (s.rbeg + s.len);
                max_gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
                w = max_gap < p->w ? max_gap : p->w;
                if (qd - rd < w && rd - qd < w) break;
            }
            if (i < er.nav) {                                              // bwamem.c:695-713
                for (i = k + 1; i < c.n; ++i) {
                    if (srt[i] == 0) continue;
                    const Seed u = seeds[(u32)srt[i]];
                    if (u.len < s.len * .95) continue;
                    if (s.qbeg <= u.qbeg && s.qbeg + s.len - u.qbeg >= s.len >> 2 && u.qbeg - s.qbeg != u.rbeg - s.rbeg) break;
                    if (u.qbeg <= s.qbeg && u.qbeg + u.len - s.qbeg >= s.len >> 2 && s.qbeg - u.qbeg != s.rbeg - u.rbeg) break;
                }
                if (i == c.n) { srt[k] = 0; --er.k; continue; }
            }
            // extend this seed: one job
            const int ql = s.qbeg, qr = l_query - (s.qbeg + s.len);
            ExtJob j;
            j.rbeg = s.rbeg; j.rmax0 = er.rmax0; j.rmax1 = er.rmax1;
            j.read = r; j.qbeg = s.qbeg; j.slen = s.len; j.lq = l_query;
            j.wl0 = clamp_band(opt, ql, opt.w, opt.pen_clip5); j.wl1 = clamp_band(opt, ql, opt.w << 1, opt.pen_clip5);
            j.wr0 = clamp_band(opt, qr, opt.w, opt.pen_clip3); j.wr1 = clamp_band(opt, qr, opt.w << 1, opt.pen_clip3);
            j.rb = j.re = 0; j.qb = j.qe = j.score = j.truesc = j.w = 0;
            j.key = ((ql >> 3) > 31 ? 31 : (ql >> 3)) << 5 | ((qr >> 3) > 31 ? 31 : (qr >> 3));
            *out = j;
            emitted = true;
            break;
        }
        if (emitted) break;
        ++er.ci; er.k = -2;
    }
    P.rstate[r] = er;
    return er.k != -3;
}

