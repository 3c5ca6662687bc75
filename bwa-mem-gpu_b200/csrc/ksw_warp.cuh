// Warp-cooperative banded affine-gap extension: ksw_extend2 (ksw.c:380-479), bit-exact.
//
// Parallelisation.  The reference's DP depends on the previous row through H(i-1,j-1), E(i,j) AND through
// the adaptive band [beg,end) (shrunk to the non-zero span of the previous row, ksw.c:466-469), the
// zero-row break (ksw.c:454) and the z-drop test (ksw.c:458-464).  A skewed anti-diagonal wavefront would
// have to speculate on all three.  Instead one warp processes ONE ROW AT A TIME, 32 query columns per
// step, which keeps every row-level decision exactly where the CPU makes it:
//   * M(i,j), E(i+1,j) and the insertion seed t(i,j) = max(M(i,j)-oe_ins, 0) depend only on row i-1, so
//     all columns of a row are independent (ksw_extend2 derives E and F from M, not from H);
//   * F(i,j+1) = max(F(i,j)-e_ins, t(i,j)) is a max-plus prefix scan: with g(j) = t(i,j) + j*e_ins,
//     F(i,j+1) = max_{j'<=j} g(j') - j*e_ins.  Integer max/add are exact and associative, so a
//     5-step shuffle scan (+ a carry between 32-column chunks) reproduces the sequential F bit for bit;
//   * row maximum with the reference's tie rule (the LAST column attaining it, ksw.c:437-438) is a
//     redux.max followed by a redux.max over matching column indices;
//   * the new band is two ballots (first / last non-zero cell), the z-drop / zero-row exits are
//     warp-uniform branches.
// Cell arithmetic uses the DPX fused max/add instructions (__vimax3_s32, __viaddmax_s32_relu, ...).
// H and E rows (the reference's eh[] array, including its stale out-of-band entries, which the
// algorithm observes) live in a warp-private int32 array (shared memory for short queries).
#pragma once
#include "simt.h"

#define KSW_NEG_INF (-(1 << 29))

struct KswResult { int score, qle, tle, gtle, gscore, max_off; };

template <class QF, class TF>
__device__ __forceinline__ KswResult ksw_extend_warp(int qlen, QF qf, int tlen, TF tf, const int8_t *mat /* 25, any space */,
                                                     int o_del, int e_del, int o_ins, int e_ins, int w, int end_bonus,
                                                     int zdrop, int h0, i32 *H, i32 *E, u64 *cells)
{
    const unsigned lane = threadIdx.x & 31;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins;
    // first row (ksw.c:395-397) in closed form; E = 0 (calloc)
    for (int j = lane; j <= qlen; j += 32) {
        int v;
        if (j == 0) v = h0;
        else if (j == 1) v = h0 > oe_ins ? h0 - oe_ins : 0;
        else {
            int vprev = h0 - oe_ins - (j - 2) * e_ins;          // eh[j-1].h while the run is alive
            v = (h0 > oe_ins && vprev > e_ins) ? vprev - e_ins : 0;
        }
        H[j] = v; E[j] = 0;
    }
    // band clamp (ksw.c:399-407)
    int mx = 0;
    for (int k = 0; k < 25; ++k) mx = mx > mat[k] ? mx : mat[k];
    int max_ins = (int)((double)(qlen * mx + end_bonus - o_ins) / e_ins + 1.);
    max_ins = max_ins > 1 ? max_ins : 1;
    w = w < max_ins ? w : max_ins;
    int max_del = (int)((double)(qlen * mx + end_bonus - o_del) / e_del + 1.);
    max_del = max_del > 1 ? max_del : 1;
    w = w < max_del ? w : max_del;

    int max = h0, max_i = -1, max_j = -1, max_ie = -1, gscore = -1, max_off = 0;
    int beg = 0, end = qlen;
    u64 ncell = 0;
    int tcache = 0;
    __syncwarp();
    for (int i = 0; i < tlen; ++i) {
        if ((i & 31) == 0) tcache = (i + (int)lane < tlen) ? tf(i + (int)lane) : 0;      // 32 target bases per fetch
        const int tb = __shfl_sync(B200_FULL_MASK, tcache, i & 31);
        const int8_t *mrow = mat + tb * 5;
        if (beg < i - w) beg = i - w;
        if (end > i + w + 1) end = i + w + 1;
        if (end > qlen) end = qlen;
        int h1 = 0;
        if (beg == 0) { h1 = h0 - (o_del + e_del * (i + 1)); if (h1 < 0) h1 = 0; }
        int m = 0, mj = -1;
        int first_nz = 0x7fffffff, last_nz = -1;
        int hcarry = h1;                // H(i, base-1) for the chunk's first column
        int gcarry = KSW_NEG_INF;       // running max of g over previous chunks
        int jexit = beg;
        if (beg < end) {
            ncell += (u64)(end - beg);
            for (int base = beg; base < end; base += 32) {
                const int j = base + (int)lane;
                const bool act = j < end;
                int Mv = 0, e = 0;
                if (act) { Mv = H[j]; e = E[j]; }
                Mv = (act && Mv) ? Mv + mrow[qf(j)] : 0;                                   // ksw.c:433
                const int en = __viaddmax_s32_relu(e, -e_del, Mv - oe_del);                // E(i+1,j), ksw.c:439-442
                const int t = act ? __vimax_s32_relu(Mv - oe_ins, 0) : 0;
                int g = act ? t + j * e_ins : KSW_NEG_INF;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {                                         // inclusive max-scan of g
                    int o = __shfl_up_sync(B200_FULL_MASK, g, d);
                    if ((int)lane >= d) g = g > o ? g : o;
                }
                g = g > gcarry ? g : gcarry;
                int gex = __shfl_up_sync(B200_FULL_MASK, g, 1);
                if (lane == 0) gex = gcarry;
                const int f = (j == beg) ? 0 : gex - (j - 1) * e_ins;                      // F(i,j), ksw.c:444-447
                const int h = act ? __vimax3_s32(Mv, e, f) : 0;                            // ksw.c:434-435
                int hl = __shfl_up_sync(B200_FULL_MASK, h, 1);
                if (lane == 0) hl = hcarry;                                                // H(i,j-1) -> eh[j].h, ksw.c:432
                if (act) { H[j] = hl; E[j] = en; }
                // row max, last column on ties (ksw.c:437-438)
                const int cm = __reduce_max_sync(B200_FULL_MASK, h);
                if (cm >= m) {
                    const int cj = __reduce_max_sync(B200_FULL_MASK, (act && h == cm) ? j : -1);
                    m = cm; mj = cj;
                }
                // non-zero span of the updated cells (ksw.c:466-469)
                const unsigned nz = __ballot_sync(B200_FULL_MASK, act && (hl != 0 || en != 0));
                if (nz) {
                    if (first_nz == 0x7fffffff) first_nz = base + __ffs((int)nz) - 1;
                    last_nz = base + 31 - __clz((int)nz);
                }
                const int nact = end - base < 32 ? end - base : 32;
                hcarry = __shfl_sync(B200_FULL_MASK, h, nact - 1);
                gcarry = __shfl_sync(B200_FULL_MASK, g, nact - 1);
            }
            jexit = end;
        }
        if (lane == 0) { H[end] = hcarry; E[end] = 0; }                                    // ksw.c:449
        __syncwarp();
        if (jexit == qlen) {                                                               // ksw.c:450-453
            max_ie = gscore > hcarry ? max_ie : i;
            gscore = gscore > hcarry ? gscore : hcarry;
        }
        if (m == 0) break;                                                                 // ksw.c:454
        if (m > max) {
            max = m; max_i = i; max_j = mj;
            int off = mj - i; off = off < 0 ? -off : off;
            max_off = max_off > off ? max_off : off;
        } else if (zdrop > 0) {                                                            // ksw.c:458-464
            if (i - max_i > mj - max_j) {
                if (max - m - ((i - max_i) - (mj - max_j)) * e_del > zdrop) break;
            } else {
                if (max - m - ((mj - max_j) - (i - max_i)) * e_ins > zdrop) break;
            }
        }
        // next band
        const int nbeg = first_nz != 0x7fffffff ? first_nz : end;
        int jl;
        if (hcarry != 0) jl = end; else jl = last_nz >= nbeg ? last_nz : nbeg - 1;
        beg = nbeg;
        end = jl + 2 < qlen ? jl + 2 : qlen;
    }
    if (cells) *cells += ncell;
    KswResult r;
    r.score = max; r.qle = max_j + 1; r.tle = max_i + 1; r.gtle = max_ie + 1; r.gscore = gscore; r.max_off = max_off;
    return r;
}
