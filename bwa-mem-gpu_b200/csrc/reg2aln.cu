// mem_reg2aln (bwamem.c:1099-1169) on the device -- for every region of a batch the final
// position, CIGAR, NM and MD: infer_bw (bwamem.c:799-806), the band-doubling loop around bwa_gen_cigar2 (bwa.c:261-347),
// ksw_global2 with its traceback matrix (ksw.c:504-606), the squeeze of leading / trailing deletions, soft clips,
// bns_depos / bns_pos2rid.  MAPQ (mem_approx_mapq_se, double arithmetic with log) stays with the caller.
//
// One region per lane (persistent lanes, atomic work counter).  The H/E row and the traceback bytes live in a
// lane-private global scratch (L1/L2-resident at the batch sizes used: <= 100 KB per lane for 150-bp reads); CIGAR
// words and MD characters go to fixed-stride slots of a per-chunk arena and are compacted by a scan + gather, so only
// the bytes that exist travel back to the host.  The traceback reproduces the reference's direction bits (ties: M over
// E over F, `>=` for H, `>` for continuing a gap), so CIGARs are identical, not just equivalent.
#include "ctx.h"
#include "extend_common.cuh"
#include "finalize_pe.cuh"
#include <cmath>

struct R2AParams {
    DevIndex ix; mem_opt_t opt;
    const u8 *seq; const i64 *off;
    const mem_alnreg_t *regs; const i32 *reg_read; int n_regs;
    b200_aln_t *aln;
    u32 *cig; int cig_stride; char *md; int md_stride;
    i32 *he; int he_cap;                 // per lane: 2 * he_cap ints
    u8 *z; i64 z_cap;                    // per lane: traceback bytes
    unsigned *work_counter; int *overflow;
    // two passes: pass 0 finishes the regions that need no DP (equal lengths, inferred band 0) and lists the others;
    // pass 1 runs the listed regions, so that the lanes of a warp are all inside ksw_global2 together
    // the list is kept as four lists by inferred band (one lane = one region, a warp costs what its widest band costs:
    // lanes that fetch consecutive entries get bands of the same class), widest class first
    i32 *dp_list; unsigned *dp_n; int pass; int dp_stride;
    // pass 1 runs with a small H / E row per lane (he_cap entries, rings of 2w + 2); regions whose band does not fit are
    // listed again and run in pass 2 with a full row
    i32 *dp_list2; unsigned *dp_n2;
    // DP passes: the region's query codes staged in shared memory, 8 per word, [word][lane] (0 = read them from global
    // memory): with the H / E rows taking most of the SM's shared memory little L1 is left, and the DP reads one query
    // code per cell
    int q_words;
    int he_smem;                         // H / E rows in shared memory (2 * he_cap * 64 ints per block)
    const b200_sel_t *sel;               // b200_reg2sam_se_batch: which regions reach the SAM (null: all)
    int warp_list;                       // 1: regions with a query part longer than R2A_WARP_LQ are listed for reg2aln_warp_kernel (list 4)
};

#define R2A_WARP_LQ 400

#define R2A_MINUS_INF (-0x40000000)

struct SeqSmem {                         // query codes in DP order, 4 bits each, word k of a lane at Q[k * 64]
    const u32 *Q;
    __device__ __forceinline__ int operator()(int j) const { return (int)(Q[(j >> 3) * 64] >> ((j & 7) << 2) & 0xfu); }
};

__device__ __forceinline__ int r2a_infer_bw(int l1, int l2, int score, int a, int q, int r)       // bwamem.c:799-806
{
    if (l1 == l2 && l1 * a - score < (q + r - a) << 1) return 0;
    int w = (int)((double)((l1 < l2 ? l1 : l2) * a - score - q) / r + 2.);
    const int d = l1 > l2 ? l1 - l2 : l2 - l1;
    if (w < d) w = d;
    return w;
}

struct R2AMd {                           // kstring appends of bwa_gen_cigar2 (kputw / kputc)
    char *s; int l, cap; int *ovf;
    __device__ __forceinline__ void putc_(char c) { if (l < cap) s[l] = c; else *ovf = 1; ++l; }
    __device__ inline void putw(int v)
    {
        char buf[12]; int n = 0;
        unsigned x = v < 0 ? (unsigned)-v : (unsigned)v;
        do { buf[n++] = (char)('0' + x % 10); x /= 10; } while (x);
        if (v < 0) putc_('-');
        while (n) putc_(buf[--n]);
    }
};

__device__ __forceinline__ void r2a_push(u32 *cigar, int &n_cigar, int cap, int op, int len, int *ovf)   // push_cigar, ksw.c:492-502
{
    if (n_cigar == 0 || op != (int)(cigar[(n_cigar - 1 < cap ? n_cigar - 1 : cap - 1)] & 0xf)) {
        if (n_cigar < cap) cigar[n_cigar] = (u32)len << 4 | (u32)op; else *ovf = 1;
        ++n_cigar;
    } else if (n_cigar - 1 < cap) cigar[n_cigar - 1] += (u32)len << 4;
}

// ksw_global2 (ksw.c:504-606) with traceback; returns the score, CIGAR (already reversed into reading order) in cigar[].
// H / E: entry j at H[j * hs] (hs = 1: lane-private global scratch; hs = block size: interleaved shared memory).
// The traceback matrix keeps the reference's three direction fields as 4 bits per cell (h: 2 bits, "E continues", "F
// continues"), eight cells per 32-bit word, rows padded to whole words: one store per eight cells instead of one per cell.
// Row i reads columns [i - w, i + w] and writes [i - w, i + w + 1], all of which row i - 1 wrote, so the H / E row is kept
// as a ring of M = 2w + 2 entries when that is less than qlen + 1 (M >= qlen + 1: no wrap-around): a lane needs
// 16 (w + 1) bytes instead of 8 (qlen + 1), which is what lets many more warps share an SM's shared memory.
template <class QF, class TF>
__device__ inline int r2a_global2(int qlen, QF qf, int tlen, TF tf, const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins,
                                  int w, int M, i32 *H, i32 *E, int hs, u8 *z, i64 z_cap, u32 *cigar, int cig_cap, int &n_cigar, int *ovf)
{
    const bool ring = M < qlen + 1;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins;
    const int n_col = qlen < 2 * w + 1 ? qlen : 2 * w + 1;
    const int row_words = (n_col + 7) >> 3;
    u32 *zw = (u32 *)z;
    n_cigar = 0;
    if ((i64)row_words * 4 * tlen > z_cap) { *ovf = 2; return R2A_MINUS_INF; }
    int j;
    H[0] = 0; E[0] = R2A_MINUS_INF;
    for (j = 1; j <= qlen && j <= w; ++j) { H[j * hs] = -(o_ins + e_ins * j); E[j * hs] = R2A_MINUS_INF; }
    if (!ring) for (; j <= qlen; ++j) H[j * hs] = E[j * hs] = R2A_MINUS_INF;      // ring: columns > w are written before they are read
    for (int i = 0; i < tlen; ++i) {
        int f = R2A_MINUS_INF, t;
        const int8_t *mrow = mat + tf(i) * 5;
        const int beg = i > w ? i - w : 0;
        const int end = i + w + 1 < qlen ? i + w + 1 : qlen;
        int h1 = beg == 0 ? -(o_del + e_del * (i + 1)) : R2A_MINUS_INF;
        u32 *zi = zw + (i64)i * row_words;
        u32 acc = 0;
        int jm = ring ? beg % M : beg;                     // ring slot of column j
        for (j = beg; j < end; ++j) {
            const int js = jm * hs;
            if (++jm == M) jm = 0;                         // M > qlen when there is no ring: never taken
            int m = H[js], e = E[js], h;
            u32 d;
            H[js] = h1;
            m += mrow[qf(j)];
            d = m >= e ? 0u : 1u;
            h = m >= e ? m : e;
            d = h >= f ? d : 2u;
            h = h >= f ? h : f;
            h1 = h;
            t = m - oe_del;
            e -= e_del;
            d |= e > t ? 4u : 0u;
            e = e > t ? e : t;
            E[js] = e;
            t = m - oe_ins;
            f -= e_ins;
            d |= f > t ? 8u : 0u;
            f = f > t ? f : t;
            const int c = j - beg;
            acc |= d << ((c & 7) << 2);
            if ((c & 7) == 7) { zi[c >> 3] = acc; acc = 0; }
        }
        if ((end - beg) & 7) zi[(end - beg) >> 3] = acc;
        H[jm * hs] = h1; E[jm * hs] = R2A_MINUS_INF;       // column `end`
    }
    const int last_end = tlen + w < qlen ? tlen + w : qlen;
    const int score = !ring ? H[qlen * hs] : (last_end == qlen ? H[(qlen % M) * hs] : R2A_MINUS_INF);
    {   // backtrack (ksw.c:587-602)
        int which = 0;
        int i = tlen - 1, k = (i + w + 1 < qlen ? i + w + 1 : qlen) - 1;
        while (i >= 0 && k >= 0) {
            const int c = k - (i > w ? i - w : 0);
            const u32 d = zw[(i64)i * row_words + (c >> 3)] >> ((c & 7) << 2) & 0xfu;
            which = which == 0 ? (int)(d & 3u) : which == 1 ? (int)(d >> 2 & 1u) : (int)(d >> 3 & 1u) << 1;
            if (which == 0) { r2a_push(cigar, n_cigar, cig_cap, 0, 1, ovf); --i; --k; }
            else if (which == 1) { r2a_push(cigar, n_cigar, cig_cap, 2, 1, ovf); --i; }
            else { r2a_push(cigar, n_cigar, cig_cap, 1, 1, ovf); --k; }
        }
        if (i >= 0) r2a_push(cigar, n_cigar, cig_cap, 2, i + 1, ovf);
        if (k >= 0) r2a_push(cigar, n_cigar, cig_cap, 1, k + 1, ovf);
        const int nc = n_cigar < cig_cap ? n_cigar : cig_cap;
        for (int a = 0; a < nc >> 1; ++a) { const u32 tmp = cigar[a]; cigar[a] = cigar[nc - 1 - a]; cigar[nc - 1 - a] = tmp; }
    }
    return score;
}

// bwa_gen_cigar2 (bwa.c:261-347): CIGAR + score + NM + MD of query[0, l_query) against text [rb, re)
template <bool DP, class QF, class TF>
__device__ inline void r2a_gen_cigar2(const R2AParams &P, const int8_t *mat, int w_, int l_query, QF qf, i64 rb, i64 rlen, TF tf, bool rev,
                                      i32 *H, i32 *E, int hs, u8 *z, u32 *cigar, int &n_cigar, int &score, int &NM, R2AMd &md, int *ovf,
                                      int he_avail, int *need_big)
{
    const mem_opt_t &opt = P.opt;
    if (l_query == rlen && w_ == 0) {                                          // no gap; no need to do DP
        if (P.cig_stride > 0) cigar[0] = (u32)l_query << 4;
        n_cigar = 1;
        score = 0;
        for (int i = 0; i < l_query; ++i) score += mat[tf(i) * 5 + qf(i)];
    } else if (DP) {
        int max_ins = (int)((double)(((l_query + 1) >> 1) * mat[0] - opt.o_ins) / opt.e_ins + 1.);
        int max_del = (int)((double)(((l_query + 1) >> 1) * mat[0] - opt.o_del) / opt.e_del + 1.);
        int max_gap = max_ins > max_del ? max_ins : max_del;
        max_gap = max_gap > 1 ? max_gap : 1;
        const int dl = (int)rlen > l_query ? (int)rlen - l_query : l_query - (int)rlen;
        int w = (max_gap + dl + 1) >> 1;
        w = w < w_ ? w : w_;
        const int min_w = dl + 3;
        w = w > min_w ? w : min_w;
        int M = 2 * w + 2;
        if (M >= l_query + 1) M = l_query + 2;             // the whole row: no ring
        if ((M < l_query + 1 ? M : l_query + 1) > he_avail) { *need_big = 1; n_cigar = 0; score = 0; return; }     // H / E of this pass too small
        score = r2a_global2(l_query, qf, (int)rlen, tf, mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins, w, M, H, E, hs, z, P.z_cap,
                            cigar, P.cig_stride, n_cigar, ovf);
    }
    {   // NM and MD (bwa.c:311-338)
        int x = 0, y = 0, u = 0, n_mm = 0, n_gap = 0;
        const int nc = n_cigar < P.cig_stride ? n_cigar : P.cig_stride;
        md.l = 0;
        for (int k = 0; k < nc; ++k) {
            const int op = (int)(cigar[k] & 0xf), len = (int)(cigar[k] >> 4);
            if (op == 0) {
                for (int i = 0; i < len; ++i) {
                    const int rc = tf(y + i);
                    if (qf(x + i) != rc) { md.putw(u); md.putc_((rev ? "TGCAN" : "ACGTN")[rc]); ++n_mm; u = 0; }
                    else ++u;
                }
                x += len; y += len;
            } else if (op == 2) {
                if (k > 0 && k < n_cigar - 1) {
                    md.putw(u); md.putc_('^');
                    for (int i = 0; i < len; ++i) md.putc_((rev ? "TGCAN" : "ACGTN")[tf(y + i)]);
                    u = 0; n_gap += len;
                }
                y += len;
            } else if (op == 1) { x += len; n_gap += len; }
        }
        md.putw(u);
        NM = n_mm + n_gap;
    }
}

// MODE 0: pass 0 (no DP code in the kernel: few registers, no shared-memory rows); MODE 1: DP passes with the query codes
// staged in shared memory (every read of the batch fits q_words); MODE 2: DP passes reading the query from global memory
template <int MODE>
__global__ void __launch_bounds__(64) reg2aln_kernel(R2AParams P)
{
    __shared__ int8_t s_mat[25];
    if (threadIdx.x < 25) s_mat[threadIdx.x] = P.opt.mat[threadIdx.x];
    __syncthreads();
    B200_DYN_SMEM(r2a_raw);
    const size_t lane_id = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    i32 *H = P.he + lane_id * 2 * (size_t)P.he_cap, *E = H + P.he_cap;
    int hs = 1;
    if (P.he_smem) { H = (i32 *)r2a_raw + threadIdx.x; E = H + (size_t)P.he_cap * 64; hs = 64; }    // [column][lane], conflict-free
    u32 *QW = (u32 *)r2a_raw + (P.he_smem ? (size_t)2 * P.he_cap * 64 : 0) + threadIdx.x;
    u8 *z = P.z + lane_id * (size_t)P.z_cap;
    const mem_opt_t &opt = P.opt;
    const DevIndex &ix = P.ix;
    const i64 l_pac = ix.l_pac;
    int ovf = 0;
    for (;;) {
        unsigned g = atomicAdd(P.work_counter, 1u);
        if (P.pass == 1) {
            const unsigned n3 = P.dp_n[3], n2 = P.dp_n[2], n1 = P.dp_n[1], n0 = P.dp_n[0];
            if (g >= n3 + n2 + n1 + n0) break;
            if (g < n3) g = (unsigned)P.dp_list[3 * (size_t)P.dp_stride + g];
            else if (g < n3 + n2) g = (unsigned)P.dp_list[2 * (size_t)P.dp_stride + (g - n3)];
            else if (g < n3 + n2 + n1) g = (unsigned)P.dp_list[(size_t)P.dp_stride + (g - n3 - n2)];
            else g = (unsigned)P.dp_list[g - n3 - n2 - n1];
        }
        else if (P.pass == 2) { if (g >= *P.dp_n2) break; g = (unsigned)P.dp_list2[g]; }
        else if (g >= (unsigned)P.n_regs) break;
        const mem_alnreg_t ar = P.regs[g];
        const int r = P.reg_read[g];
        const u8 *query = P.seq + P.off[r];
        const int l_seq = (int)(P.off[r + 1] - P.off[r]);
        u32 *cigar = P.cig + (size_t)g * P.cig_stride;
        R2AMd md; md.s = P.md + (size_t)g * P.md_stride; md.l = 0; md.cap = P.md_stride; md.ovf = &ovf;
        b200_aln_t a;
        memset(&a, 0, sizeof(a));
        if (P.sel && P.sel[g].kind == 0) { a.rid = -1; a.pos = -1; P.aln[g] = a; continue; }        // not in the SAM
        if (ar.rb < 0 || ar.re < 0) { a.rid = -1; a.pos = -1; a.flag = 0x4; P.aln[g] = a; continue; }   // unmapped record
        const int qb = ar.qb, qe = ar.qe;
        const i64 rb = ar.rb, re = ar.re;
        if (ar.secondary >= 0) a.flag |= 0x100;
        int tmp = r2a_infer_bw(qe - qb, (int)(re - rb), ar.truesc, opt.a, opt.o_del, opt.e_del);
        int w2 = r2a_infer_bw(qe - qb, (int)(re - rb), ar.truesc, opt.a, opt.o_ins, opt.e_ins);
        w2 = w2 > tmp ? w2 : tmp;
        if (w2 > opt.w) w2 = w2 < ar.w ? w2 : ar.w;
        int n_cigar = 0, NM = -1, score = 0, last_sc = -(1 << 30), it = 0;
        const int lq = qe - qb;
        const i64 rlen = re - rb;
        const bool ok = lq > 0 && rb < re && !(rb < l_pac && re > l_pac);      // bwa.c:272
        const bool rev = rb >= l_pac;
        if (P.pass == 0 && ok && !(lq == rlen && (w2 < opt.w << 2 ? w2 : opt.w << 2) == 0)) {   // needs ksw_global2: second pass
            // long regions: one per warp (reg2aln_warp_kernel); the others by inferred band, one per lane
            const int cls = lq > R2A_WARP_LQ && P.warp_list ? 4 : w2 <= 12 ? 0 : w2 <= 18 ? 1 : w2 <= 26 ? 2 : 3;
            P.dp_list[cls * (size_t)P.dp_stride + atomicAdd(P.dp_n + cls, 1u)] = (i32)g;
            continue;
        }
        int need_big = 0;
        const bool staged = MODE == 1 && ok;
        if (staged) {
            for (int k = 0; k * 8 < lq; ++k) {
                u32 wd = 0;
                for (int b = 0; b < 8 && k * 8 + b < lq; ++b) {
                    const int j = k * 8 + b;
                    wd |= (u32)(rev ? query[qb + lq - 1 - j] : query[qb + j]) << (b << 2);
                }
                QW[k * 64] = wd;
            }
        }
        do {
            w2 = w2 < opt.w << 2 ? w2 : opt.w << 2;
            n_cigar = 0; NM = -1;
            if (MODE == 1) {
                if (!ok) ;
                else if (rev) r2a_gen_cigar2<true>(P, s_mat, w2, lq, SeqSmem{QW}, rb, rlen, RefSeqRev{&ix, rb, rlen}, true, H, E, hs, z, cigar, n_cigar, score, NM, md, &ovf, P.he_cap, &need_big);
                else r2a_gen_cigar2<true>(P, s_mat, w2, lq, SeqSmem{QW}, rb, rlen, RefSeqFwd{&ix, rb}, false, H, E, hs, z, cigar, n_cigar, score, NM, md, &ovf, P.he_cap, &need_big);
            } else if (ok) {
                if (rev) r2a_gen_cigar2<MODE != 0>(P, s_mat, w2, lq, SeqRev{query + qb, lq}, rb, rlen, RefSeqRev{&ix, rb, rlen}, true, H, E, hs, z, cigar, n_cigar, score, NM, md, &ovf, P.he_cap, &need_big);
                else r2a_gen_cigar2<MODE != 0>(P, s_mat, w2, lq, SeqFwd{query + qb}, rb, rlen, RefSeqFwd{&ix, rb}, false, H, E, hs, z, cigar, n_cigar, score, NM, md, &ovf, P.he_cap, &need_big);
            }
            if (need_big) break;
            if (score == last_sc || w2 == opt.w << 2) break;
            last_sc = score;
            w2 <<= 1;
        } while (++it < 3 && score < ar.truesc - opt.a);
        if (need_big) {
            if (P.pass == 1) P.dp_list2[atomicAdd(P.dp_n2, 1u)] = (i32)g;      // again in pass 2, with a full row
            else ovf = 2;                                                       // cannot happen: pass 0 has no DP, pass 2 has the full row
            continue;
        }
        a.NM = NM;
        int is_rev;
        i64 pos = dev_depos(ix, rb < l_pac ? rb : re - 1, &is_rev);
        a.is_rev = is_rev;
        const int cap = P.cig_stride;
        if (n_cigar > 0 && n_cigar <= cap) {                                       // squeeze out leading or trailing deletions
            if ((cigar[0] & 0xf) == 2) {
                pos += cigar[0] >> 4;
                --n_cigar;
                for (int k = 0; k < n_cigar; ++k) cigar[k] = cigar[k + 1];
            } else if ((cigar[n_cigar - 1] & 0xf) == 2) --n_cigar;
        }
        if (qb != 0 || qe != l_seq) {                                              // add clipping to CIGAR
            const int clip5 = is_rev ? l_seq - qe : qb, clip3 = is_rev ? qb : l_seq - qe;
            if (n_cigar + 2 > cap) ovf = 1;
            else {
                if (clip5) { for (int k = n_cigar; k > 0; --k) cigar[k] = cigar[k - 1]; cigar[0] = (u32)clip5 << 4 | 3; ++n_cigar; }
                if (clip3) cigar[n_cigar++] = (u32)clip3 << 4 | 3;
            }
        }
        a.n_cigar = n_cigar; a.l_md = md.l;
        a.rid = dev_pos2rid(ix, pos);
        a.pos = pos - (a.rid >= 0 ? ix.ann_off[a.rid] : 0);
        a.score = ar.score; a.sub = ar.sub > ar.csub ? ar.sub : ar.csub;
        a.is_alt = ar.is_alt; a.alt_sc = ar.alt_sc;
        if (P.sel) {                                                               // mem_reg2sam, bwamem.c:1035-1038
            a.flag |= P.sel[g].flag;
            if (ar.secondary >= 0) a.sub = -1;
        }
        P.aln[g] = a;
    }
    if (ovf) atomicMax(P.overflow, ovf);
}

#include "reg2aln_warp.cuh"

// mem_reg2aln (bwamem.c:1099-1169) for the regions of list 4: one region per warp.  The scalar decisions are those of
// reg2aln_kernel (every lane computes them, they are warp-uniform); the DP is r2a_global2_warp; lane 0 does the rest.
__global__ void __launch_bounds__(128) reg2aln_warp_kernel(R2AParams P)
{
    __shared__ int8_t s_mat[25];
    if (threadIdx.x < 25) s_mat[threadIdx.x] = P.opt.mat[threadIdx.x];
    __syncthreads();
    const unsigned lane = threadIdx.x & 31;
    const size_t wi = (size_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    i32 *H = P.he + wi * 2 * (size_t)P.he_cap, *E = H + P.he_cap;
    u8 *z = P.z + wi * (size_t)P.z_cap;
    const mem_opt_t &opt = P.opt;
    const DevIndex &ix = P.ix;
    const i64 l_pac = ix.l_pac;
    const unsigned n_list = P.dp_n[4];
    int ovf = 0;
    for (;;) {
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(P.work_counter, 1u);
        t = __shfl_sync(B200_FULL_MASK, t, 0);
        if (t >= n_list) break;
        const unsigned g = (unsigned)P.dp_list[4 * (size_t)P.dp_stride + t];
        const mem_alnreg_t ar = P.regs[g];
        const int r = P.reg_read[g];
        const u8 *query = P.seq + P.off[r];
        const int l_seq = (int)(P.off[r + 1] - P.off[r]);
        u32 *cigar = P.cig + (size_t)g * P.cig_stride;
        R2AMd md; md.s = P.md + (size_t)g * P.md_stride; md.l = 0; md.cap = P.md_stride; md.ovf = &ovf;
        b200_aln_t a;
        memset(&a, 0, sizeof(a));
        const int qb = ar.qb, qe = ar.qe;
        const i64 rb = ar.rb, re = ar.re;
        if (ar.secondary >= 0) a.flag |= 0x100;
        int tmp = r2a_infer_bw(qe - qb, (int)(re - rb), ar.truesc, opt.a, opt.o_del, opt.e_del);
        int w2 = r2a_infer_bw(qe - qb, (int)(re - rb), ar.truesc, opt.a, opt.o_ins, opt.e_ins);
        w2 = w2 > tmp ? w2 : tmp;
        if (w2 > opt.w) w2 = w2 < ar.w ? w2 : ar.w;
        int n_cigar = 0, NM = -1, score = 0, last_sc = -(1 << 30), it = 0;
        const int lq = qe - qb;
        const i64 rlen = re - rb;
        const bool ok = lq > 0 && rb < re && !(rb < l_pac && re > l_pac);      // bwa.c:272
        const bool rev = rb >= l_pac;
        do {
            w2 = w2 < opt.w << 2 ? w2 : opt.w << 2;
            n_cigar = 0; NM = -1;
            if (ok) {
                // bwa_gen_cigar2 (bwa.c:261-347): band as in r2a_gen_cigar2, the DP across the warp, NM / MD by lane 0
                int max_ins = (int)((double)(((lq + 1) >> 1) * s_mat[0] - opt.o_ins) / opt.e_ins + 1.);
                int max_del = (int)((double)(((lq + 1) >> 1) * s_mat[0] - opt.o_del) / opt.e_del + 1.);
                int max_gap = max_ins > max_del ? max_ins : max_del;
                max_gap = max_gap > 1 ? max_gap : 1;
                const int dl = (int)rlen > lq ? (int)rlen - lq : lq - (int)rlen;
                int w = (max_gap + dl + 1) >> 1;
                w = w < w2 ? w : w2;
                const int min_w = dl + 3;
                w = w > min_w ? w : min_w;
                if (lq + 2 > P.he_cap) { ovf = 2; score = 0; }
                else if (lq == rlen && w2 == 0) {                               // no gap (cannot be listed here, kept for completeness)
                    if (lane == 0 && P.cig_stride > 0) cigar[0] = (u32)lq << 4;
                    n_cigar = 1; score = 0;
                    for (int i = 0; i < lq; ++i) score += s_mat[(rev ? RefSeqRev{&ix, rb, rlen}(i) : RefSeqFwd{&ix, rb}(i)) * 5 + (rev ? SeqRev{query + qb, lq}(i) : SeqFwd{query + qb}(i))];
                } else if (rev) score = r2a_global2_warp(lq, SeqRev{query + qb, lq}, (int)rlen, RefSeqRev{&ix, rb, rlen}, s_mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins,
                                                         w, H, E, z, P.z_cap, cigar, P.cig_stride, n_cigar, &ovf);
                else score = r2a_global2_warp(lq, SeqFwd{query + qb}, (int)rlen, RefSeqFwd{&ix, rb}, s_mat, opt.o_del, opt.e_del, opt.o_ins, opt.e_ins,
                                              w, H, E, z, P.z_cap, cigar, P.cig_stride, n_cigar, &ovf);
            }
            if (score == last_sc || w2 == opt.w << 2) break;
            last_sc = score;
            w2 <<= 1;
        } while (++it < 3 && score < ar.truesc - opt.a);
        if (lane == 0) {
            if (ok) {                                                          // NM and MD (bwa.c:311-338)
                int x = 0, y = 0, u = 0, n_mm = 0, n_gap = 0;
                const int nc = n_cigar < P.cig_stride ? n_cigar : P.cig_stride;
                for (int k = 0; k < nc; ++k) {
                    const int op = (int)(cigar[k] & 0xf), len = (int)(cigar[k] >> 4);
                    if (op == 0) {
                        for (int i = 0; i < len; ++i) {
                            const int rc = rev ? RefSeqRev{&ix, rb, rlen}(y + i) : RefSeqFwd{&ix, rb}(y + i);
                            const int qc = rev ? SeqRev{query + qb, lq}(x + i) : SeqFwd{query + qb}(x + i);
                            if (qc != rc) { md.putw(u); md.putc_((rev ? "TGCAN" : "ACGTN")[rc]); ++n_mm; u = 0; }
                            else ++u;
                        }
                        x += len; y += len;
                    } else if (op == 2) {
                        if (k > 0 && k < n_cigar - 1) {
                            md.putw(u); md.putc_('^');
                            for (int i = 0; i < len; ++i) {
                                const int rc = rev ? RefSeqRev{&ix, rb, rlen}(y + i) : RefSeqFwd{&ix, rb}(y + i);
                                md.putc_((rev ? "TGCAN" : "ACGTN")[rc]);
                            }
                            u = 0; n_gap += len;
                        }
                        y += len;
                    } else if (op == 1) { x += len; n_gap += len; }
                }
                md.putw(u);
                NM = n_mm + n_gap;
            }
            a.NM = NM;
            int is_rev;
            i64 pos = dev_depos(ix, rb < l_pac ? rb : re - 1, &is_rev);
            a.is_rev = is_rev;
            const int cap = P.cig_stride;
            if (n_cigar > 0 && n_cigar <= cap) {                                       // squeeze out leading or trailing deletions
                if ((cigar[0] & 0xf) == 2) {
                    pos += cigar[0] >> 4;
                    --n_cigar;
                    for (int k = 0; k < n_cigar; ++k) cigar[k] = cigar[k + 1];
                } else if ((cigar[n_cigar - 1] & 0xf) == 2) --n_cigar;
            }
            if (qb != 0 || qe != l_seq) {                                              // add clipping to CIGAR
                const int clip5 = is_rev ? l_seq - qe : qb, clip3 = is_rev ? qb : l_seq - qe;
                if (n_cigar + 2 > cap) ovf = 1;
                else {
                    if (clip5) { for (int k = n_cigar; k > 0; --k) cigar[k] = cigar[k - 1]; cigar[0] = (u32)clip5 << 4 | 3; ++n_cigar; }
                    if (clip3) cigar[n_cigar++] = (u32)clip3 << 4 | 3;
                }
            }
            a.n_cigar = n_cigar; a.l_md = md.l;
            a.rid = dev_pos2rid(ix, pos);
            a.pos = pos - (a.rid >= 0 ? ix.ann_off[a.rid] : 0);
            a.score = ar.score; a.sub = ar.sub > ar.csub ? ar.sub : ar.csub;
            a.is_alt = ar.is_alt; a.alt_sc = ar.alt_sc;
            if (P.sel) {                                                               // mem_reg2sam, bwamem.c:1035-1038
                a.flag |= P.sel[g].flag;
                if (ar.secondary >= 0) a.sub = -1;
            }
            P.aln[g] = a;
        }
        __syncwarp();
    }
    if (ovf) atomicMax(P.overflow, ovf);
}

// ---- compaction of the fixed-stride arena into the output pools
__global__ void r2a_lengths_kernel(const b200_aln_t *aln, int n, i32 *n_cig, i32 *l_md)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    n_cig[g] = aln[g].n_cigar; l_md[g] = aln[g].l_md;
}
__global__ void r2a_gather_kernel(b200_aln_t *aln, int n, const u32 *cig, int cig_stride, const char *md, int md_stride,
                                  const i64 *cig_off, const i64 *md_off, u32 *cig_out, char *md_out, i64 cig_base, i64 md_base)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const int nc = aln[g].n_cigar, lm = aln[g].l_md;
    const u32 *cs = cig + (size_t)g * cig_stride;
    const char *ms = md + (size_t)g * md_stride;
    for (int k = 0; k < nc && k < cig_stride; ++k) cig_out[cig_off[g] + k] = cs[k];
    for (int k = 0; k < lm && k < md_stride; ++k) md_out[md_off[g] + k] = ms[k];
    aln[g].cigar_off = cig_base + cig_off[g];
    aln[g].md_off = md_base + md_off[g];
}

__global__ void pe_unmapped_fill_kernel(i32 *mate_reg, i32 *read_flag, int n)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) { mate_reg[r] = -1; read_flag[r] = (r & 1) ? 0x81 : 0x41; }
}

// regs_out / out_sel non-null: the worker2 steps of finalize.cuh run first on each chunk (chunks then hold whole reads)
static int64_t reg2aln_impl(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, const uint8_t *seqs, const int64_t *off,
                            const mem_alnreg_t *regs, const int32_t *cnt, b200_aln_t *out_aln, int64_t cap_cigar,
                            uint32_t *out_cigar, int64_t cap_md, char *out_md, int64_t totals[2],
                            mem_alnreg_t *regs_out, b200_sel_t *out_sel, int64_t n_processed,
                            r2a_chunk_fn sink = nullptr, void *sink_user = nullptr, const mem_pestat_t *pes = nullptr)
{
    const bool finalize = out_sel != nullptr || sink != nullptr;
    const bool pe = pes != nullptr;                        // pairs: reads 2i / 2i + 1 are mates, mem_sam_pe's decisions (finalize_pe.cuh)

    const bool resident = regs == nullptr;                 // reads and regions are those of the batch the context just aligned
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    if (resident) {
        B200_CHECK(cudaSetDevice(ctx->device));
        n_reads = ctx->n_reads;
    } else if (b200_batch_upload(ctx, n_reads, seqs, off)) return -1;
    if (pe && (!sink || (n_reads & 1) || (n_processed & 1))) { b200_set_error("paired-end worker2 needs the device consumer and whole pairs"); return -1; }
    cudaStream_t st = ctx->st;
    if (ctx->n_bytes && launch_pack_reads(ctx->d_seq.as<u8>(), ctx->n_bytes, st)) return -1;          // reads as nt4 (idempotent)
    // host arrays of the call live in page-locked buffers of the context (slots 12, 13): no 8 MB allocations per call from
    // every host thread (page faults, TLB shootdowns), and their uploads are real asynchronous copies
    i64 *roff = (i64 *)b200_ctx_pinned(ctx, 12, ((size_t)n_reads + 1) * 8);
    if (!roff) return -1;
    roff[0] = 0;
    int max_cnt = 0;
    for (int r = 0; r < n_reads; ++r) { roff[r + 1] = roff[r] + cnt[r]; max_cnt = std::max(max_cnt, (int)cnt[r]); }
    const i64 n_regs = roff[n_reads];
    if (resident && n_regs != ctx->total_regs) { b200_set_error("region counts do not match the resident batch"); return -1; }
    if (totals) totals[0] = totals[1] = 0;
    if (n_regs == 0 && sink) {                             // nothing aligned: the consumer still sees every read (unmapped records)
        DevBuf &zr = ctx->d_r2a[14];
        if (zr.ensure(((size_t)n_reads + 1) * 8)) return -1;
        B200_CHECK(cudaMemsetAsync(zr.p, 0, ((size_t)n_reads + 1) * 8, st));
        R2AChunk c;
        memset(&c, 0, sizeof(c));
        c.r_hi = n_reads; c.d_roff = zr.as<i64>();
        if (pe) {                                          // every read unmapped: mates unmapped too, flags 0x41 / 0x81 (no_pairing, bwamem_pair.c:395-416)
            DevBuf &d_mate = ctx->d_sam[16], &d_rflag = ctx->d_sam[17];
            if (d_mate.ensure((size_t)n_reads * 4 + 4) || d_rflag.ensure((size_t)n_reads * 4 + 4)) return -1;
            B200_LAUNCH(pe_unmapped_fill_kernel, (n_reads + 255) / 256, 256, 0, st, d_mate.as<i32>(), d_rflag.as<i32>(), n_reads);
            c.mate_reg = d_mate.as<i32>(); c.read_flag = d_rflag.as<i32>();
        }
        return sink(sink_user, c) ? -1 : 0;
    }
    if (n_regs == 0) return 0;
    if (n_regs >= 0x7fffffff) { b200_set_error("too many regions in one call"); return -1; }
    i32 *reg_read = (i32 *)b200_ctx_pinned(ctx, 13, (size_t)n_regs * 4 + 4);
    if (!reg_read) return -1;
    {
        i64 g = 0;
        for (int r = 0; r < n_reads; ++r) for (int k = 0; k < cnt[r]; ++k) reg_read[g++] = r;
    }
    const int max_len = std::max(1, ctx->max_len);
    const int w4 = (opt->w << 2) + 64;
    // regions with a query part longer than R2A_WARP_LQ go to the warp-per-region kernel (tunable r2a_warp = 0: never); the
    // lane-per-region passes then only see short regions, and their rows / traceback scratch are sized for those
    const bool use_warp = ctx->r2a_warp && max_len > R2A_WARP_LQ;
    const int ml_lane = use_warp ? R2A_WARP_LQ : max_len;
    const int cig_stride = 2 * max_len + 8, md_stride = 3 * max_len + 64, he_cap = ml_lane + 2;
    auto zcap_of = [&](int ml) { return (((i64)(std::min(ml, 2 * w4 + 1) + 7) / 8 * 4) * ((i64)ml + w4) + 15) / 16 * 16; };   // 4 bits per cell, rows padded to words
    const i64 z_cap = zcap_of(ml_lane);
    const int he_cap_w = max_len + 2;
    const i64 z_cap_w = zcap_of(max_len);
    const size_t per_reg = (size_t)cig_stride * 4 + md_stride + sizeof(b200_aln_t) + sizeof(mem_alnreg_t) + 32;
    const int chunk = (int)std::max<i64>(std::max(1024, 2 * max_cnt), std::min<i64>((i64)1 << 20, ((i64)6 << 30) / (i64)per_reg));
    const size_t per_lane = (size_t)z_cap + (size_t)he_cap * 8;
    // lane-private traceback scratch: short reads need ~50 KB per lane (all 8 blocks per SM fit in a few GB); long reads need
    // MBs per lane, and the number of lanes in flight is what bounds them, so they may take up to half of the free memory
    i64 z_budget = (i64)12 << 30;
    if ((i64)per_lane * ctx->n_sms * 8 * 64 > z_budget) {
        size_t mem_free = 0, mem_total = 0;
        if (cudaMemGetInfo(&mem_free, &mem_total) == cudaSuccess)
            z_budget = std::max<i64>(z_budget, std::min<i64>((i64)80 << 30, (i64)(mem_free + ctx->d_r2a[5].bytes + ctx->d_r2a[6].bytes) / 2));
    }
    const i64 lanes_fit = std::max<i64>(64, z_budget / (i64)per_lane);
    const int grid_max = (int)std::max<i64>(1, std::min<i64>((i64)ctx->n_sms * 8, lanes_fit / 64));
    int n_warps = 0;                                        // warps of reg2aln_warp_kernel: each needs a full-size traceback area
    if (use_warp) {
        size_t mem_free = 0, mem_total = 0;
        i64 budget = (i64)16 << 30;
        if (cudaMemGetInfo(&mem_free, &mem_total) == cudaSuccess) budget = std::min<i64>(budget, (i64)(mem_free + ctx->d_r2a[6].bytes) / 2);
        n_warps = (int)std::max<i64>(4, std::min<i64>((i64)ctx->n_sms * 8, budget / (z_cap_w + (i64)he_cap_w * 8))) & ~3;
    }
    // grow-only buffers kept in the context: a second call of the same shape allocates nothing
    DevBuf &d_regs = ctx->d_r2a[0], &d_rr = ctx->d_r2a[1], &d_aln = ctx->d_r2a[2], &d_cig = ctx->d_r2a[3], &d_md = ctx->d_r2a[4],
           &d_he = ctx->d_r2a[5], &d_z = ctx->d_r2a[6], &d_nc = ctx->d_r2a[7], &d_lm = ctx->d_r2a[8], &d_co = ctx->d_r2a[9],
           &d_mo = ctx->d_r2a[10], &d_tmp = ctx->d_r2a[11], &d_pool_c = ctx->d_r2a[12], &d_pool_m = ctx->d_r2a[13];
    if (d_regs.ensure((size_t)chunk * sizeof(mem_alnreg_t)) || d_rr.ensure((size_t)chunk * 4) || d_aln.ensure((size_t)chunk * sizeof(b200_aln_t)) ||
        d_cig.ensure((size_t)chunk * cig_stride * 4) || d_md.ensure((size_t)chunk * md_stride) ||
        d_he.ensure(std::max((size_t)grid_max * 64 * he_cap * 8, (size_t)n_warps * he_cap_w * 8)) ||
        d_z.ensure(std::max((size_t)grid_max * 64 * (size_t)z_cap, (size_t)n_warps * (size_t)z_cap_w)) ||
        d_nc.ensure((size_t)chunk * 4) || d_lm.ensure((size_t)chunk * 4) || d_co.ensure(((size_t)chunk + 1) * 8) ||
        d_mo.ensure(((size_t)chunk + 1) * 8) || d_tmp.ensure(((size_t)chunk / 1024 + 2) * 8) ||
        d_pool_c.ensure((size_t)chunk * cig_stride * 4) || d_pool_m.ensure((size_t)chunk * md_stride) || ctx->d_ctl.ensure(CTL_SLOTS * 8)) return -1;
    DevBuf &d_roff = ctx->d_r2a[14], &d_sel = ctx->d_r2a[15], &d_log = ctx->d_r2a[16], &d_dp = ctx->d_r2a[17];
    if (d_dp.ensure(((size_t)5 * chunk + 8) * 4)) return -1;
    if (finalize) {
        if (d_roff.ensure(((size_t)n_reads + 1) * 8) || d_sel.ensure((size_t)chunk * sizeof(b200_sel_t))) return -1;
        B200_CHECK(cudaMemcpyAsync(d_roff.p, roff, ((size_t)n_reads + 1) * 8, cudaMemcpyHostToDevice, st));
        const int want = 4 * max_len + 65536;                  // log(l), l <= max(qe - qb, re - rb); the kernel checks the bound
        if (ctx->log_cap < want) {
            std::vector<double> tab((size_t)want);
            for (int i = 0; i < want; ++i) tab[i] = std::log((double)i);                             // the host's libm, as mem_approx_mapq_se
            if (d_log.ensure((size_t)want * 8)) return -1;
            B200_CHECK(cudaMemcpyAsync(d_log.p, tab.data(), (size_t)want * 8, cudaMemcpyHostToDevice, st));
            B200_CHECK(cudaStreamSynchronize(st));
            ctx->log_cap = want;
        }
    }
    FinPeParams PE;
    memset(&PE, 0, sizeof(PE));
    if (pe) {                                              // per-orientation tables of the pairing score's transcendental term (finalize_pe.cuh)
        DevBuf &d_mate = ctx->d_sam[16], &d_rflag = ctx->d_sam[17], &d_vscr = ctx->d_sam[18];
        if (d_mate.ensure((size_t)n_reads * 4 + 4) || d_rflag.ensure((size_t)n_reads * 4 + 4) || d_vscr.ensure((size_t)chunk * 16 + 16)) return -1;
        memcpy(PE.pes, pes, 4 * sizeof(mem_pestat_t));
        for (int d = 0; d < 4; ++d) {
            PE.ptab[d] = nullptr; PE.ptab_n[d] = 0;
            if (pes[d].failed || pes[d].high < pes[d].low) continue;
            const i64 m = (i64)pes[d].high - pes[d].low + 1;
            if (m > ((i64)1 << 24)) { b200_set_error("paired-end: insert-size window of %lld values", (long long)m); return -1; }
            std::vector<double> tab((size_t)m);
            for (i64 k = 0; k < m; ++k) {                  // bwamem_pair.c:243-244, the host's libm as in the reference
                const i64 dist = pes[d].low + k;
                const double ns = (dist - pes[d].avg) / pes[d].std;
                tab[(size_t)k] = .721 * log(2. * erfc(fabs(ns) * M_SQRT1_2)) * opt->a;
            }
            DevBuf &dt = ctx->d_sam[19 + d];
            if (dt.ensure((size_t)m * 8)) return -1;
            B200_CHECK(cudaMemcpyAsync(dt.p, tab.data(), (size_t)m * 8, cudaMemcpyHostToDevice, st));
            B200_CHECK(cudaStreamSynchronize(st));          // tab is a local
            PE.ptab[d] = dt.as<double>(); PE.ptab_n[d] = (int)m;
        }
        PE.mate_reg = d_mate.as<i32>(); PE.read_flag = d_rflag.as<i32>(); PE.vscr = d_vscr.as<u64>();
        PE.l_pac = ctx->ix.l_pac; PE.ann_off = ctx->ix.ann_off;
    }
    i64 cig_total = 0, md_total = 0;
    bool fits = true;
    int r_lo = 0;
    const bool timing = getenv("B200_R2A_TIMING") != nullptr;        // per-phase CUDA-event times on stderr (diagnostics)
    cudaEvent_t tev[8];
    float tacc[7] = {0, 0, 0, 0, 0, 0, 0};
    if (timing) for (auto &e : tev) cudaEventCreate(&e);
    auto tmark = [&](int k) { if (timing) cudaEventRecord(tev[k], st); };
    for (i64 g0 = 0; g0 < n_regs;) {
        int n = (int)std::min<i64>(chunk, n_regs - g0), r_hi = r_lo;
        if (finalize) {
            const int step = pe ? 2 : 1;                   // paired-end chunks hold whole pairs
            while (r_hi + step <= n_reads && roff[r_hi + step] - g0 <= chunk) r_hi += step;
            n = (int)(roff[r_hi] - g0);
        }
        tmark(0);
        if (resident)      // a copy: marking is not idempotent (sub_n accumulates), the resident regions stay as worker1 left them
            B200_CHECK(cudaMemcpyAsync(d_regs.p, ctx->d_final.as<mem_alnreg_t>() + g0, (size_t)n * sizeof(mem_alnreg_t), cudaMemcpyDeviceToDevice, st));
        else B200_CHECK(cudaMemcpyAsync(d_regs.p, regs + g0, (size_t)n * sizeof(mem_alnreg_t), cudaMemcpyHostToDevice, st));
        B200_CHECK(cudaMemcpyAsync(d_rr.p, reg_read + g0, (size_t)n * 4, cudaMemcpyHostToDevice, st));
        B200_CHECK(cudaMemsetAsync(ctx->d_ctl.p, 0, CTL_SLOTS * 8, st));
        R2AParams P;
        memset(&P, 0, sizeof(P));
        P.ix = ctx->ix; P.opt = *opt; P.seq = ctx->d_seq.as<u8>(); P.off = ctx->d_off.as<i64>();
        P.regs = d_regs.as<mem_alnreg_t>(); P.reg_read = d_rr.as<i32>(); P.n_regs = n;
        P.aln = d_aln.as<b200_aln_t>(); P.cig = d_cig.as<u32>(); P.cig_stride = cig_stride; P.md = d_md.as<char>(); P.md_stride = md_stride;
        P.he = d_he.as<i32>(); P.he_cap = he_cap; P.z = d_z.as<u8>(); P.z_cap = z_cap;
        P.work_counter = (unsigned *)(ctx->d_ctl.as<u64>() + CTL_WORK); P.overflow = (int *)(ctx->d_ctl.as<u64>() + CTL_OVF2);
        P.dp_list = d_dp.as<i32>(); P.dp_stride = chunk; P.dp_n = (unsigned *)(d_dp.as<i32>() + (size_t)5 * chunk); P.pass = 0;
        P.warp_list = use_warp ? 1 : 0;
        B200_CHECK(cudaMemsetAsync(P.dp_n, 0, 32, st));
        tmark(1);
        if (finalize) {
            FinParams F;
            memset(&F, 0, sizeof(F));
            F.opt = *opt; F.regs = d_regs.as<mem_alnreg_t>(); F.roff = d_roff.as<i64>(); F.g0 = g0; F.r0 = r_lo; F.n_reads = r_hi - r_lo;
            F.id0 = n_processed; F.sel = d_sel.as<b200_sel_t>(); F.logtab = d_log.as<double>(); F.log_cap = ctx->log_cap;
            F.overflow = P.overflow;
            const int fgrid = std::max(1, std::min(ctx->n_sms * 8, (F.n_reads + 127) / 128));
            if (pe) { PE.F = F; B200_LAUNCH(finalize_pe_kernel, fgrid, 128, 0, st, PE); }
            else B200_LAUNCH(finalize_se_kernel, fgrid, 128, 0, st, F);
            P.sel = d_sel.as<b200_sel_t>();
        }
        const int grid = std::max(1, std::min(grid_max, (n + 63) / 64));
        const size_t he_bytes = (size_t)2 * he_cap * 64 * 4;
        P.q_words = ctx->r2a_small && max_len <= 256 ? (max_len + 7) / 8 : 0;
        const size_t q_bytes = (size_t)P.q_words * 64 * 4;
        P.he_smem = he_bytes + q_bytes <= 100 * 1024 ? 1 : 0;
        const size_t smem = (P.he_smem ? he_bytes : 0) + q_bytes;
#ifndef B200_EMU
        B200_CHECK(cudaFuncSetAttribute(reg2aln_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        B200_CHECK(cudaFuncSetAttribute(reg2aln_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
        tmark(2);
        {   // pass 0 does no DP: no shared-memory rows, so the registers (8 blocks per SM), not the rows (2), set its occupancy
            R2AParams P0 = P;
            P0.he_smem = 0; P0.q_words = 0;
            const int grid0 = std::max(1, std::min(ctx->n_sms * 14, (n + 63) / 64));       // 72 registers, no scratch: 14 blocks per SM
            B200_LAUNCH(reg2aln_kernel<0>, grid0, 64, 0, st, P0);
        }
        if (use_warp) {                                     // the long regions pass 0 listed: one per warp
            R2AParams W = P;
            W.he_cap = he_cap_w; W.z_cap = z_cap_w; W.he_smem = 0; W.q_words = 0;
            W.work_counter = (unsigned *)(ctx->d_ctl.as<u64>() + CTL_RETRY);
            B200_LAUNCH(reg2aln_warp_kernel, n_warps / 4, 128, 0, st, W);
        }
        tmark(3);
        P.pass = 1; P.work_counter = (unsigned *)(ctx->d_ctl.as<u64>() + CTL_WORK2);
        {   // pass 1 with a small row per lane: rings of up to HE_SMALL entries (w <= 23): 25 KB of shared memory per block instead of 78 KB at 150 bp
            const int HE_SMALL = 48;
            R2AParams Q = P;
            Q.dp_list2 = d_lm.as<i32>(); Q.dp_n2 = (unsigned *)(ctx->d_ctl.as<u64>() + CTL_TOTAL);      // d_lm is free until the lengths kernel
            if (he_cap > HE_SMALL && ctx->r2a_small) {
                Q.he_cap = HE_SMALL; Q.he_smem = 1;
                if (Q.q_words) B200_LAUNCH(reg2aln_kernel<1>, grid, 64, (size_t)2 * HE_SMALL * 64 * 4 + q_bytes, st, Q);
                else B200_LAUNCH(reg2aln_kernel<2>, grid, 64, (size_t)2 * HE_SMALL * 64 * 4, st, Q);
                Q.pass = 2; Q.he_cap = he_cap; Q.he_smem = P.he_smem;
                Q.work_counter = (unsigned *)(ctx->d_ctl.as<u64>() + CTL_ARENA_USED);
                if (Q.q_words) B200_LAUNCH(reg2aln_kernel<1>, grid, 64, smem, st, Q);
                else B200_LAUNCH(reg2aln_kernel<2>, grid, 64, smem, st, Q);
            } else if (Q.q_words) B200_LAUNCH(reg2aln_kernel<1>, grid, 64, smem, st, Q);
            else B200_LAUNCH(reg2aln_kernel<2>, grid, 64, smem, st, Q);
        }
        tmark(4);
        B200_LAUNCH(r2a_lengths_kernel, (n + 255) / 256, 256, 0, st, (const b200_aln_t *)d_aln.p, n, d_nc.as<i32>(), d_lm.as<i32>());
        B200_CHECK(cudaGetLastError());
        if (launch_exclusive_scan(d_nc.as<i32>(), d_co.as<i64>(), n, d_tmp.as<i64>(), st)) return -1;
        if (launch_exclusive_scan(d_lm.as<i32>(), d_mo.as<i64>(), n, d_tmp.as<i64>(), st)) return -1;
        B200_LAUNCH(r2a_gather_kernel, (n + 255) / 256, 256, 0, st, d_aln.as<b200_aln_t>(), n, (const u32 *)d_cig.p, cig_stride,
                    (const char *)d_md.p, md_stride, (const i64 *)d_co.p, (const i64 *)d_mo.p, d_pool_c.as<u32>(), d_pool_m.as<char>(),
                    cig_total, md_total);
        B200_CHECK(cudaGetLastError());
        tmark(5);
        i64 nc = 0, nm = 0; u64 ovf = 0;
        B200_CHECK(cudaMemcpyAsync(&nc, d_co.as<i64>() + n, 8, cudaMemcpyDeviceToHost, st));
        B200_CHECK(cudaMemcpyAsync(&nm, d_mo.as<i64>() + n, 8, cudaMemcpyDeviceToHost, st));
        B200_CHECK(cudaMemcpyAsync(&ovf, ctx->d_ctl.as<u64>() + CTL_OVF2, 8, cudaMemcpyDeviceToHost, st));
        B200_CHECK(cudaStreamSynchronize(st));
        if ((int)ovf) {
            b200_set_error("reg2aln: per-region %s capacity exceeded", (int)ovf == 3 ? "log table" : (int)ovf == 2 ? "traceback" : "CIGAR/MD");
            return -1;
        }
        if (sink) {                                        // results stay on the device: the consumer runs on this chunk now
            R2AChunk c;
            c.r_lo = r_lo; c.r_hi = r_hi; c.g0 = g0; c.n = n; c.regs = d_regs.as<mem_alnreg_t>(); c.aln = d_aln.as<b200_aln_t>();
            c.sel = d_sel.as<b200_sel_t>(); c.cig = d_pool_c.as<u32>(); c.md = d_pool_m.as<char>(); c.cig_base = cig_total; c.md_base = md_total;
            c.d_roff = d_roff.as<i64>();
            c.mate_reg = pe ? PE.mate_reg : nullptr; c.read_flag = pe ? PE.read_flag : nullptr;
            if (sink(sink_user, c)) return -1;
            cig_total += nc; md_total += nm;
            g0 += n; r_lo = r_hi;
            continue;
        }
        B200_CHECK(cudaMemcpyAsync(out_aln + g0, d_aln.p, (size_t)n * sizeof(b200_aln_t), cudaMemcpyDeviceToHost, st));
        if (finalize) {
            B200_CHECK(cudaMemcpyAsync(out_sel + g0, d_sel.p, (size_t)n * sizeof(b200_sel_t), cudaMemcpyDeviceToHost, st));
            if (regs_out) B200_CHECK(cudaMemcpyAsync(regs_out + g0, d_regs.p, (size_t)n * sizeof(mem_alnreg_t), cudaMemcpyDeviceToHost, st));
        }
        if (cig_total + nc > cap_cigar || md_total + nm > cap_md) fits = false;
        if (fits) {
            if (nc) B200_CHECK(cudaMemcpyAsync(out_cigar + cig_total, d_pool_c.p, (size_t)nc * 4, cudaMemcpyDeviceToHost, st));
            if (nm) B200_CHECK(cudaMemcpyAsync(out_md + md_total, d_pool_m.p, (size_t)nm, cudaMemcpyDeviceToHost, st));
        }
        tmark(6);
        B200_CHECK(cudaStreamSynchronize(st));
        if (timing) for (int k = 0; k < 6; ++k) { float ms = 0; cudaEventElapsedTime(&ms, tev[k], tev[k + 1]); tacc[k] += ms; }
        cig_total += nc; md_total += nm;
        g0 += n; r_lo = r_hi;
    }
    if (timing) {
        fprintf(stderr, "[b200 reg2aln] %lld regions: regs in %.2f ms, finalize %.2f, pass0 %.2f, pass1 %.2f, compaction %.2f, out %.2f\n",
                (long long)n_regs, tacc[0], tacc[1], tacc[2], tacc[3], tacc[4], tacc[5]);
        for (auto &e : tev) cudaEventDestroy(e);
    }
    if (totals) { totals[0] = cig_total; totals[1] = md_total; }
    if (!fits) { b200_set_error("reg2aln: output pools too small (need %lld CIGAR words, %lld MD bytes)", (long long)cig_total, (long long)md_total); return -2; }
    return n_regs;
}

extern "C" int64_t b200_reg2aln_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, const uint8_t *seqs, const int64_t *off,
                                      const mem_alnreg_t *regs, const int32_t *cnt, b200_aln_t *out_aln, int64_t cap_cigar,
                                      uint32_t *out_cigar, int64_t cap_md, char *out_md, int64_t totals[2])
{
    return reg2aln_impl(ctx, opt, n_reads, seqs, off, regs, cnt, out_aln, cap_cigar, out_cigar, cap_md, out_md, totals, nullptr, nullptr, 0);
}

extern "C" int64_t b200_reg2sam_se_batch(b200_ctx_t *ctx, const mem_opt_t *opt, int n_reads, int64_t n_processed, const uint8_t *seqs,
                                         const int64_t *off, mem_alnreg_t *regs, const int32_t *cnt, b200_sel_t *out_sel,
                                         b200_aln_t *out_aln, int64_t cap_cigar, uint32_t *out_cigar, int64_t cap_md, char *out_md,
                                         int64_t totals[2])
{
    if (!out_sel || !regs) { b200_set_error("b200_reg2sam_se_batch: regs and out_sel are required"); return -1; }
    return reg2aln_impl(ctx, opt, n_reads, seqs, off, regs, cnt, out_aln, cap_cigar, out_cigar, cap_md, out_md, totals, regs, out_sel, n_processed);
}

extern "C" int64_t b200_batch_reg2sam_se(b200_ctx_t *ctx, const mem_opt_t *opt, int64_t n_processed, const int32_t *cnt,
                                         mem_alnreg_t *regs_out, b200_sel_t *out_sel, b200_aln_t *out_aln, int64_t cap_cigar,
                                         uint32_t *out_cigar, int64_t cap_md, char *out_md, int64_t totals[2])
{
    if (!out_sel || !cnt) { b200_set_error("b200_batch_reg2sam_se: cnt and out_sel are required"); return -1; }
    return reg2aln_impl(ctx, opt, 0, nullptr, nullptr, nullptr, cnt, out_aln, cap_cigar, out_cigar, cap_md, out_md, totals, regs_out, out_sel, n_processed);
}

int64_t b200_batch_reg2sam_dev(b200_ctx *ctx, const mem_opt_t *opt, i64 n_processed, const i32 *cnt, r2a_chunk_fn fn, void *user,
                               const mem_pestat_t *pes)
{
    if (!fn || !cnt) { b200_set_error("b200_batch_reg2sam_dev: cnt and fn are required"); return -1; }
    return reg2aln_impl(ctx, opt, 0, nullptr, nullptr, nullptr, cnt, nullptr, 0, nullptr, 0, nullptr, nullptr, nullptr, nullptr, n_processed, fn, user, pes);
}

int launch_pestat_pairs(const mem_opt_t *opt, i64 l_pac, const mem_alnreg_t *regs, const i64 *roff, int n_pairs, u64 *out, cudaStream_t st)
{
    if (n_pairs <= 0) return 0;
    B200_LAUNCH(pestat_pairs_kernel, (n_pairs + 255) / 256, 256, 0, st, *opt, l_pac, regs, roff, n_pairs, out);
    B200_CHECK(cudaGetLastError());
    return 0;
}
