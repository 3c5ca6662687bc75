// ksw_align2 (ksw.c:343-365) = striped local Smith-Waterman ksw_u8 (ksw.c:111-230) / ksw_i16 (ksw.c:232-334) on a
// query profile (ksw_qinit, ksw.c:63-109), with the 2nd-best score outside the best hit's window, the optional
// reverse pass for start coordinates (KSW_XSTART) and the early stop (KSW_XSTOP).
//
// Bit-exactness here means reproducing the reference's VECTOR algorithm, not just "a" Smith-Waterman: E(i+1,j) is
// derived from H before the lazy-F correction, F is carried only inside a stripe segment in the main loop, and all
// arithmetic saturates (unsigned bytes, or signed-add / unsigned-sub 16-bit words).  So one SSE register of the
// reference = one GROUP OF 16 LANES here (8 used for 16-bit scores): lane k holds query positions j + k*slen, a
// byte shift across the register is a width-16 shuffle, movemask is a ballot, the horizontal max is a shuffle
// reduction.  Two independent problems run in one warp (sub-warp masks).  H0/H1/E/Hmax and the query codes live in
// shared memory, [segment index][lane] so that accesses are conflict-free.
#pragma once
#include "simt.h"

#define KSWA_XBYTE  0x10000
#define KSWA_XSTOP  0x20000
#define KSWA_XSUBO  0x40000
#define KSWA_XSTART 0x80000

struct KswAlnRes { int score, te, qe, score2, te2, tb, qb; };

struct KswaScratch {
    u16 *sm;        // group-private shared memory: 5 arrays x slen_cap x 16 lanes (H0, H1, E, Hmax, query codes)
    int slen_cap;
    u64 *b;         // group-private list of (row maximum << 32 | row) entries, ksw.c:205-213
    int b_cap;
};

// One pass of ksw_u8 / ksw_i16.  size: 1 = bytes, 2 = 16-bit.  qf(k): query code at position k (0 <= k < qlen),
// tf(i): target code.  All 16 lanes of the group call this together; the result is identical in all of them.
// *ovf is set when the problem does not fit the scratch (caller reports an error).
template <class QF, class TF>
__device__ inline KswAlnRes kswa_pass(int size, int qlen, QF qf, int tlen, TF tf, const int8_t *mat, int o_del, int e_del,
                                      int o_ins, int e_ins, int xtra, const KswaScratch &S, unsigned gmask, int glane, int *ovf)
{
    KswAlnRes r;
    r.score = 0; r.te = r.qe = r.score2 = r.te2 = r.tb = r.qb = -1;                      // g_defr, ksw.c:44
    const int p = size == 1 ? 16 : 8;                                                    // values per register
    const int slen = (qlen + p - 1) / p;                                                 // ksw.c:70
    if (slen > S.slen_cap) { *ovf = 1; return r; }
    const bool lane_on = glane < p;
    int shift = 127, mdiff = 0;                                                          // ksw.c:79-86
    for (int a = 0; a < 25; ++a) { if (mat[a] < (int8_t)shift) shift = mat[a]; if (mat[a] > (int8_t)mdiff) mdiff = mat[a]; }
    const int qmax = mdiff;
    shift = (256 - shift) & 0xff;
    u16 *H0 = S.sm + glane, *H1 = H0 + S.slen_cap * 16, *E = H1 + S.slen_cap * 16, *Hmax = E + S.slen_cap * 16,
        *QC = Hmax + S.slen_cap * 16;
    for (int j = 0; j < slen; ++j) {                                                     // ksw.c:135-139 + profile codes
        const int k = j + glane * slen;
        H0[j * 16] = 0; E[j * 16] = 0; Hmax[j * 16] = 0;
        QC[j * 16] = (u16)((lane_on && k < qlen) ? qf(k) : 255);
    }
    __syncwarp(gmask);
    const int minsc = (xtra & KSWA_XSUBO) ? (xtra & 0xffff) : 0x10000;
    const int endsc = (xtra & KSWA_XSTOP) ? (xtra & 0xffff) : 0x10000;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins;
    int te = -1, gmax = 0, n_b = 0, last_row = -2, last_max = 0;
    for (int i = 0; i < tlen; ++i) {
        const int8_t *mrow = mat + tf(i) * 5;
        int f = 0, mx = 0;
        int h = __shfl_up_sync(gmask, slen ? (int)H0[(slen - 1) * 16] : 0, 1, 16);       // ksw.c:144-145
        if (glane == 0) h = 0;
        for (int j = 0; j < slen; ++j) {                                                 // ksw.c:146-171
            const int qc = QC[j * 16];
            const int sc = qc < 5 ? mrow[qc] : 0;
            int e = E[j * 16];
            if (size == 1) {
                h = h + sc + shift; h = h > 255 ? 255 : h;                               // adds_epu8(h, S), S = score + shift
                h = h - shift; h = h < 0 ? 0 : h;                                        // subs_epu8(h, shift)
            } else {
                h = h + sc; h = h > 32767 ? 32767 : (h < -32768 ? -32768 : h);           // adds_epi16
            }
            h = h > e ? h : e;
            h = h > f ? h : f;
            mx = mx > h ? mx : h;
            H1[j * 16] = (u16)h;
            int t;
            e = e - e_del; e = e < 0 ? 0 : e; t = h - oe_del; t = t < 0 ? 0 : t; e = e > t ? e : t;
            E[j * 16] = (u16)e;
            f = f - e_ins; f = f < 0 ? 0 : f; t = h - oe_ins; t = t < 0 ? 0 : t; f = f > t ? f : t;
            h = H0[j * 16];
        }
        bool done = false;                                                               // lazy F, ksw.c:173-185 / 288-299
        for (int k = 0; k < 16 && !done; ++k) {
            f = __shfl_up_sync(gmask, f, 1, 16);
            if (glane == 0) f = 0;
            for (int j = 0; j < slen; ++j) {
                int hh = H1[j * 16];
                hh = hh > f ? hh : f;
                H1[j * 16] = (u16)hh;
                hh = hh - oe_ins; hh = hh < 0 ? 0 : hh;
                f = f - e_ins; f = f < 0 ? 0 : f;
                const unsigned gt = __ballot_sync(gmask, lane_on && f > hh);             // movemask(cmpeq(subs(f,h),0)) / cmpgt
                if ((gt & gmask) == 0) { done = true; break; }
            }
        }
        int imax = lane_on ? mx : 0;                                                     // __max_16 / __max_8
        for (int d = 8; d > 0; d >>= 1) { const int o = __shfl_xor_sync(gmask, imax, d, 16); imax = imax > o ? imax : o; }
        if (imax >= minsc) {                                                             // ksw.c:190-199
            // the list lives in global scratch (lane 0 writes); its length and last entry are tracked in registers
            if (n_b == 0 || last_row + 1 != i) {
                if (n_b < S.b_cap) { if (glane == 0) S.b[n_b] = (u64)imax << 32 | (u32)i; }
                else *ovf = 1;
                ++n_b; last_row = i; last_max = imax;
            } else if (last_max < imax) {
                if (glane == 0 && n_b <= S.b_cap) S.b[n_b - 1] = (u64)imax << 32 | (u32)i;
                last_row = i; last_max = imax;
            }
        }
        if (imax > gmax) {                                                               // ksw.c:200-206
            gmax = imax; te = i;
            for (int j = 0; j < slen; ++j) Hmax[j * 16] = H1[j * 16];
            if (size == 1 ? (gmax + shift >= 255 || gmax >= endsc) : (gmax >= endsc)) break;
        }
        u16 *tmp = H1; H1 = H0; H0 = tmp;
    }
    __syncwarp(gmask);
    r.score = size == 1 ? (gmax + shift < 255 ? gmax : 255) : gmax;
    r.te = te;
    if (size == 2) r.qe = -1;
    if (!(size == 1 && r.score == 255)) {                                                // ksw.c:210-228 / 319-331
        int best = -1, bpos = 0x7fffffff;
        if (lane_on)
            for (int j = 0; j < slen; ++j) {
                const int v = Hmax[j * 16], pos = j + glane * slen;
                if (v > best || (v == best && pos < bpos)) { best = v; bpos = pos; }
            }
        for (int d = 8; d > 0; d >>= 1) {
            const int ob = __shfl_xor_sync(gmask, best, d, 16), op = __shfl_xor_sync(gmask, bpos, d, 16);
            if (ob > best || (ob == best && op < bpos)) { best = ob; bpos = op; }
        }
        r.qe = bpos;
        if (n_b) {
            int sc2 = -1, te2 = -1;
            if (glane == 0) {
                const int win = (r.score + qmax - 1) / qmax;
                const int low = te - win, high = te + win;
                const int nb = n_b < S.b_cap ? n_b : S.b_cap;
                for (int k = 0; k < nb; ++k) {
                    const int e = (int)(u32)S.b[k];
                    if ((e < low || e > high) && (int)(S.b[k] >> 32) > sc2) { sc2 = (int)(S.b[k] >> 32); te2 = e; }
                }
            }
            r.score2 = __shfl_sync(gmask, sc2, 0, 16);
            r.te2 = __shfl_sync(gmask, te2, 0, 16);
        }
    }
    __syncwarp(gmask);
    return r;
}

// ksw_align2 (ksw.c:343-365).  Query and target come through accessors (plain byte arrays, the reverse complement
// of a read, bases decoded from the packed reference ...); the reverse pass of KSW_XSTART reads the reversed
// prefixes through index arithmetic instead of reversing buffers in place (ksw.c:358,361).
struct KswaFwd { const u8 *p; __device__ __forceinline__ int operator()(int i) const { return p[i]; } };
template <class Q> struct KswaRevQ { Q q; int last; __device__ __forceinline__ int operator()(int i) const { return q(last - i); } };
template <class T> struct KswaRevT { T t; int te; __device__ __forceinline__ int operator()(int i) const { return i <= te ? t(te - i) : t(i); } };

template <class Q, class T>
__device__ inline KswAlnRes kswa_align2_f(int qlen, Q query, int tlen, T target, const int8_t *mat, int o_del, int e_del,
                                          int o_ins, int e_ins, int xtra, const KswaScratch &S, unsigned gmask, int glane, int *ovf)
{
    const int size = (xtra & KSWA_XBYTE) ? 1 : 2;
    KswAlnRes r = kswa_pass(size, qlen, query, tlen, target, mat, o_del, e_del, o_ins, e_ins, xtra, S, gmask, glane, ovf);
    if ((xtra & KSWA_XSTART) == 0 || ((xtra & KSWA_XSUBO) && r.score < (xtra & 0xffff))) return r;
    const KswAlnRes rr = kswa_pass(size, r.qe + 1, KswaRevQ<Q>{query, r.qe}, tlen, KswaRevT<T>{target, r.te}, mat, o_del, e_del,
                                   o_ins, e_ins, KSWA_XSTOP | r.score, S, gmask, glane, ovf);
    if (r.score == rr.score) { r.tb = r.te - rr.te; r.qb = r.qe - rr.qe; }
    return r;
}

__device__ inline KswAlnRes kswa_align2(int qlen, const u8 *query, int tlen, const u8 *target, const int8_t *mat, int o_del,
                                        int e_del, int o_ins, int e_ins, int xtra, const KswaScratch &S, unsigned gmask,
                                        int glane, int *ovf)
{
    return kswa_align2_f(qlen, KswaFwd{query}, tlen, KswaFwd{target}, mat, o_del, e_del, o_ins, e_ins, xtra, S, gmask, glane, ovf);
}
