// mem_reg2aln for LONG regions: ksw_global2 (ksw.c:504-606) with one region per WARP, the band across the lanes.
//
// reg2aln_kernel maps one region to one lane, the right shape for 100-300 bp reads (a few dozen band columns, a few KB
// of traceback per lane).  A 5 kbp region with an inferred band of several hundred columns is millions of cells and MBs of
// traceback: only a few thousand lanes fit, each running a dependent scalar chain (the reference sketches the same
// decomposition in cuda/ksw_CUDA.cu:571, ksw_global_warp).  Here a warp owns the region: each row is cut into chunks of 32
// columns, one per lane.  Within a row only F depends on the previous column, and it is derived from M, not H
// (t = m - oe_ins, ksw.c:565-569), so F(i, j) = max_{k < j} (m_k - oe_ins - (j - 1 - k) e_ins) is a max-plus prefix scan:
// with g_k = m_k - oe_ins + k e_ins a 5-step shuffle scan (plus a carry between chunks) reproduces the sequential F bit for
// bit, and with it the direction bits (ties M > E > F; >= for H, > for gap continuation), so the CIGAR is the
// reference's, not merely an equivalent one.  Direction nibbles of eight neighbouring lanes are merged by shuffles into
// one 32-bit store in the layout of r2a_global2.  The backtrack, NM / MD and the clipping are done by lane 0.
#pragma once

template <class QF, class TF>
__device__ inline int r2a_global2_warp(int qlen, QF qf, int tlen, TF tf, const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins,
                                       int w, i32 *H, i32 *E, u8 *z, i64 z_cap, u32 *cigar, int cig_cap, int &n_cigar, int *ovf)
{
    const unsigned lane = threadIdx.x & 31;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins;
    const int n_col = qlen < 2 * w + 1 ? qlen : 2 * w + 1;
    const int row_words = (n_col + 7) >> 3;
    u32 *zw = (u32 *)z;
    n_cigar = 0;
    if ((i64)row_words * 4 * tlen > z_cap) { *ovf = 2; return R2A_MINUS_INF; }
    for (int j = (int)lane; j <= qlen; j += 32) {                     // first row (ksw.c:520-524)
        H[j] = j == 0 ? 0 : (j <= w ? -(o_ins + e_ins * j) : R2A_MINUS_INF);
        E[j] = R2A_MINUS_INF;
    }
    __syncwarp();
    for (int i = 0; i < tlen; ++i) {
        const int8_t *mrow = mat + tf(i) * 5;
        const int beg = i > w ? i - w : 0;
        const int end = i + w + 1 < qlen ? i + w + 1 : qlen;
        int f_carry = R2A_MINUS_INF;                                  // F entering the chunk's first column
        int h_carry = beg == 0 ? -(o_del + e_del * (i + 1)) : R2A_MINUS_INF;      // H(i, c0 - 1)
        u32 *zi = zw + (i64)i * row_words;
        for (int c0 = beg; c0 < end; c0 += 32) {
            const int j = c0 + (int)lane;
            const bool valid = j < end;
            int m = R2A_MINUS_INF, e = R2A_MINUS_INF;
            if (valid) { m = H[j] + mrow[qf(j)]; e = E[j]; }
            const int t_ins = m - oe_ins;
            // F of this column: the scan of g over the lanes below, and the carry from the chunks before
            int g = valid ? t_ins + (int)lane * e_ins : R2A_MINUS_INF;
            int incl = g;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(B200_FULL_MASK, incl, d); if ((int)lane >= d) incl = incl > o ? incl : o; }
            int excl = __shfl_up_sync(B200_FULL_MASK, incl, 1);
            if (lane == 0) excl = R2A_MINUS_INF;
            int f = f_carry - (int)lane * e_ins;
            { const int s = excl - ((int)lane - 1) * e_ins; if (lane > 0 && s > f) f = s; }
            const int all = __shfl_sync(B200_FULL_MASK, incl, 31);
            { int nf = f_carry - 32 * e_ins; const int s = all - 31 * e_ins; f_carry = nf > s ? nf : s; }
            u32 d = m >= e ? 0u : 1u;
            int h = m >= e ? m : e;
            d = h >= f ? d : 2u;
            h = h >= f ? h : f;
            int t = m - oe_del;
            e -= e_del;
            d |= e > t ? 4u : 0u;
            e = e > t ? e : t;
            d |= f - e_ins > t_ins ? 8u : 0u;
            int h_prev = __shfl_up_sync(B200_FULL_MASK, h, 1);
            if (lane == 0) h_prev = h_carry;
            const int n_valid = end - c0 < 32 ? end - c0 : 32;
            h_carry = __shfl_sync(B200_FULL_MASK, h, n_valid - 1);
            if (valid) { H[j] = h_prev; E[j] = e; }
            // eight nibbles -> one word (the chunk starts at a multiple of 32 columns from beg, so groups of 8 lanes = one word)
            u32 wd = valid ? d << ((lane & 7) << 2) : 0u;
            wd |= __shfl_xor_sync(B200_FULL_MASK, wd, 1);
            wd |= __shfl_xor_sync(B200_FULL_MASK, wd, 2);
            wd |= __shfl_xor_sync(B200_FULL_MASK, wd, 4);
            if ((lane & 7) == 0 && valid) zi[(j - beg) >> 3] = wd;
        }
        if (lane == 0) { H[end] = h_carry; E[end] = R2A_MINUS_INF; }   // column `end` (ksw.c:583)
        __syncwarp();
    }
    const int score = H[qlen];
    if (lane == 0) {   // backtrack (ksw.c:587-602)
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
    n_cigar = __shfl_sync(B200_FULL_MASK, n_cigar, 0);
    __syncwarp();
    return score;
}
