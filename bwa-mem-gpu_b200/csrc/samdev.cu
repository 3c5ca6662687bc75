// Device SAM formatter: see samdev.cuh.  Field order, tags and every special case follow mem_aln2sam (bwamem.c:832-956).
#include "samdev.cuh"
#include <algorithm>
#include <string.h>

#define SD_SOFTCLIP 0x200      /* MEM_F_SOFTCLIP, bwa.h:80 */
#define SD_REF_HDR 0x100       /* MEM_F_REF_HDR,  bwa.h:79 */
#define SD_XB 0x2000           /* MEM_F_XB,       bwa.h:84 */

namespace {

struct SdCount {                                          // pass 1: lengths only
    i64 l = 0;
    __device__ __forceinline__ void ch(char) { ++l; }
    __device__ __forceinline__ void lit(const char *, int n) { l += n; }
    __device__ __forceinline__ void mem(const char *, i64 n) { l += n; }
    __device__ __forceinline__ void bases(const u8 *, int, int n, bool) { l += n; }
    __device__ __forceinline__ void quals(const u8 *, int, int n, bool) { l += n; }
    __device__ __forceinline__ void tabs_to_spaces(i64) {}
};
struct SdWrite {                                          // pass 2: into the final text
    char *p; i64 l = 0;
    __device__ __forceinline__ explicit SdWrite(char *q) : p(q) {}
    __device__ __forceinline__ void ch(char c) { p[l++] = c; }
    __device__ __forceinline__ void lit(const char *z, int n) { for (int i = 0; i < n; ++i) p[l + i] = z[i]; l += n; }
    __device__ __forceinline__ void mem(const char *z, i64 n) { for (i64 i = 0; i < n; ++i) p[l + i] = z[i]; l += n; }
    // n bases starting at s[from] (forward) or ending at s[from] going down (reverse complement), as letters
    __device__ __forceinline__ void bases(const u8 *s, int from, int n, bool rev)
    {
        char *d = p + l;
        int i = 0;
        for (; i < n && ((size_t)(d + i) & 3); ++i) { const int c = rev ? s[from - i] : s[from + i]; d[i] = rev ? "TGCAN"[c > 4 ? 4 : c] : "ACGTN"[c > 4 ? 4 : c]; }
        for (; i + 4 <= n; i += 4) {                      // one 32-bit store per four letters
            u32 w = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int c = rev ? s[from - i - k] : s[from + i + k];
                w |= (u32)(u8)(rev ? "TGCAN"[c > 4 ? 4 : c] : "ACGTN"[c > 4 ? 4 : c]) << (8 * k);
            }
            *(u32 *)(d + i) = w;
        }
        for (; i < n; ++i) { const int c = rev ? s[from - i] : s[from + i]; d[i] = rev ? "TGCAN"[c > 4 ? 4 : c] : "ACGTN"[c > 4 ? 4 : c]; }
        l += n;
    }
    __device__ __forceinline__ void quals(const u8 *s, int from, int n, bool rev)
    {
        char *d = p + l;
        int i = 0;
        for (; i < n && ((size_t)(d + i) & 3); ++i) d[i] = (char)(rev ? s[from - i] : s[from + i]);
        for (; i + 4 <= n; i += 4) {
            u32 w = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) w |= (u32)(rev ? s[from - i - k] : s[from + i + k]) << (8 * k);
            *(u32 *)(d + i) = w;
        }
        for (; i < n; ++i) d[i] = (char)(rev ? s[from - i] : s[from + i]);
        l += n;
    }
    __device__ __forceinline__ void tabs_to_spaces(i64 from) { for (i64 i = from; i < l; ++i) if (p[i] == '\t') p[i] = ' '; }
};

template <class S>
__device__ __forceinline__ void sd_int(S &s, long long v)                   // kputw / kputl
{
    char buf[24]; int n = 0;
    unsigned long long x = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
    do { buf[n++] = (char)('0' + x % 10); x /= 10; } while (x);
    if (v < 0) s.ch('-');
    while (n) s.ch(buf[--n]);
}

// "%.3f" of (double)num / den for 0 < num, den < 2^31, as glibc prints it: the binary quotient (one correctly rounded
// division, as the reference's) rounded to three decimals, ties to even on the EXACT binary value.  d = M * 2^E with a
// 53-bit M, so d * 1000 = (M * 1000) * 2^E is exact in 64-bit integer arithmetic.
template <class S>
__device__ void sd_fixed3(S &s, int num, int den)
{
#ifdef B200_EMU
    volatile double dv = (double)num / (double)den; const double d = dv;
    unsigned long long bits; memcpy(&bits, &d, 8);
#else
    const double d = __ddiv_rn((double)num, (double)den);
    const unsigned long long bits = (unsigned long long)__double_as_longlong(d);
#endif
    const int be = (int)(bits >> 52 & 0x7ff);
    unsigned long long M = bits & 0xfffffffffffffull;
    int E;
    if (be == 0) E = -1074; else { M |= 1ull << 52; E = be - 1075; }      // d = M * 2^E
    unsigned long long q;                                                  // round(d * 1000)
    if (E >= 0) q = (M * 1000ull) << E;                                    // d < 2^31: no overflow
    else {
        const int sh = -E;
        const unsigned long long P = M * 1000ull;                          // < 2^63
        if (sh >= 64) q = 0;
        else {
            q = P >> sh;
            const unsigned long long rem = P & ((1ull << sh) - 1), half = 1ull << (sh - 1);
            if (rem > half || (rem == half && (q & 1))) ++q;
        }
    }
    sd_int(s, (long long)(q / 1000));
    s.ch('.');
    const int f = (int)(q % 1000);
    s.ch((char)('0' + f / 100)); s.ch((char)('0' + f / 10 % 10)); s.ch((char)('0' + f % 10));
}

struct SdAln {                                            // mem_aln_t as the text needs it
    i64 pos; int rid, flag, is_rev, is_alt, n_cigar, NM, score, sub, alt_sc, l_md, mapq;
    const u32 *cig; const char *md;
};

__device__ __forceinline__ SdAln sd_unmapped()
{
    SdAln a; a.pos = -1; a.rid = -1; a.flag = 0x4; a.is_rev = a.is_alt = a.n_cigar = a.NM = a.score = a.sub = a.alt_sc = a.l_md = a.mapq = 0;
    a.cig = nullptr; a.md = nullptr;
    return a;
}
__device__ __forceinline__ SdAln sd_load(const SamDevParams &P, i64 g)     // region g of the chunk
{
    const b200_aln_t t = P.aln[g];
    SdAln a;
    a.pos = t.pos; a.rid = t.rid; a.flag = t.flag; a.is_rev = t.is_rev; a.is_alt = t.is_alt; a.n_cigar = t.n_cigar; a.NM = t.NM;
    a.score = t.score; a.sub = t.sub; a.alt_sc = t.alt_sc; a.l_md = t.l_md; a.mapq = P.sel[g].mapq;
    a.cig = P.cig + (t.cigar_off - P.cig_base); a.md = P.md + (t.md_off - P.md_base);
    return a;
}

template <class S>
__device__ __forceinline__ void sd_name(S &s, const SamDevParams &P, int rid)
{
    s.mem(P.ann_names + P.ann_name_off[rid], P.ann_name_off[rid + 1] - P.ann_name_off[rid]);
}

template <class S>
__device__ void sd_cigar(S &s, const mem_opt_t &opt, const SdAln &p, int which)     // add_cigar, bwamem.c:819-830
{
    if (!p.n_cigar) { s.ch('*'); return; }
    for (int i = 0; i < p.n_cigar; ++i) {
        int c = (int)(p.cig[i] & 0xf);
        if (!(opt.flag & SD_SOFTCLIP) && !p.is_alt && (c == 3 || c == 4)) c = which ? 4 : 3;
        sd_int(s, p.cig[i] >> 4); s.ch("MIDSH"[c]);
    }
}

__device__ __forceinline__ int sd_rlen(const SdAln &a)                       // get_rlen, bwamem.c:808-817
{
    int l = 0;
    for (int k = 0; k < a.n_cigar; ++k) { const int op = (int)(a.cig[k] & 0xf); if (op == 0 || op == 2) l += (int)(a.cig[k] >> 4); }
    return l;
}

// one record: the read's `which`-th printed region (chunk-relative region g), or the unmapped record (g < 0)
template <class S>
__device__ void sd_record(S &s, const SamDevParams &P, int r /* read of the batch */, i64 b /* first region of the read */, int n,
                          i64 g, int which, int extra_flag, bool has_mate, const SdAln &mate_in)
{
    const mem_opt_t &opt = P.opt;
    SdAln p = g >= 0 ? sd_load(P, g) : sd_unmapped();
    p.flag |= extra_flag;
    SdAln m = mate_in;
    const int l_seq = (int)(P.off[r + 1] - P.off[r]);
    const u8 *seq = P.seq + P.off[r], *qual = P.qual ? P.qual + P.off[r] : nullptr;
    // flags (bwamem.c:838-848)
    if (has_mate) p.flag |= 0x1;
    if (p.rid < 0) p.flag |= 0x4;
    if (has_mate && m.rid < 0) p.flag |= 0x8;
    if (p.rid < 0 && has_mate && m.rid >= 0) { p.rid = m.rid; p.pos = m.pos; p.is_rev = m.is_rev; p.n_cigar = 0; }
    if (has_mate && m.rid < 0 && p.rid >= 0) { m.rid = p.rid; m.pos = p.pos; m.is_rev = p.is_rev; m.n_cigar = 0; }
    if (p.is_rev) p.flag |= 0x10;
    if (has_mate && m.is_rev) p.flag |= 0x20;
    const int flag = p.flag;
    s.mem(P.names + P.name_off[r], P.name_off[r + 1] - P.name_off[r]); s.ch('\t');
    sd_int(s, (flag & 0xffff) | (flag & 0x10000 ? 0x100 : 0)); s.ch('\t');
    if (p.rid >= 0) {
        sd_name(s, P, p.rid); s.ch('\t');
        sd_int(s, p.pos + 1); s.ch('\t');
        sd_int(s, p.mapq); s.ch('\t');
        sd_cigar(s, opt, p, which);
    } else s.lit("*\t0\t0\t*", 7);
    s.ch('\t');
    if (has_mate && m.rid >= 0) {                          // mate position (bwamem.c:866-879)
        if (p.rid == m.rid) s.ch('='); else sd_name(s, P, m.rid);
        s.ch('\t');
        sd_int(s, m.pos + 1); s.ch('\t');
        if (p.rid == m.rid) {
            const i64 p0 = p.pos + (p.is_rev ? sd_rlen(p) - 1 : 0);
            const i64 p1 = m.pos + (m.is_rev ? sd_rlen(m) - 1 : 0);
            if (m.n_cigar == 0 || p.n_cigar == 0) s.ch('0');
            else sd_int(s, -(p0 - p1 + (p0 > p1 ? 1 : p0 < p1 ? -1 : 0)));
        } else s.ch('0');
    } else s.lit("*\t0\t0", 5);
    s.ch('\t');
    if (flag & 0x100) s.lit("*\t*", 3);
    else {
        int qb = 0, qe = l_seq;
        if (p.n_cigar && which && !(opt.flag & SD_SOFTCLIP) && !p.is_alt) {
            const int c0 = (int)(p.cig[0] & 0xf), c1 = (int)(p.cig[p.n_cigar - 1] & 0xf);
            if (!p.is_rev) {
                if (c0 == 4 || c0 == 3) qb += (int)(p.cig[0] >> 4);
                if (c1 == 4 || c1 == 3) qe -= (int)(p.cig[p.n_cigar - 1] >> 4);
            } else {
                if (c0 == 4 || c0 == 3) qe -= (int)(p.cig[0] >> 4);
                if (c1 == 4 || c1 == 3) qb += (int)(p.cig[p.n_cigar - 1] >> 4);
            }
        }
        const int mlen = qe - qb > 0 ? qe - qb : 0;
        if (!p.is_rev) {
            s.bases(seq, qb, mlen, false); s.ch('\t');
            if (qual) s.quals(qual, qb, mlen, false); else s.ch('*');
        } else {
            s.bases(seq, qe - 1, mlen, true); s.ch('\t');
            if (qual) s.quals(qual, qe - 1, mlen, true); else s.ch('*');
        }
    }
    if (p.n_cigar) {
        s.lit("\tNM:i:", 6); sd_int(s, p.NM);
        s.lit("\tMD:Z:", 6); s.mem(p.md, p.l_md);
    }
    if (has_mate && m.n_cigar) { s.lit("\tMC:Z:", 6); sd_cigar(s, opt, m, which); }
    if (p.score >= 0) { s.lit("\tAS:i:", 6); sd_int(s, p.score); }
    if (p.sub >= 0) { s.lit("\tXS:i:", 6); sd_int(s, p.sub); }
    if (P.rg && P.rg_len) { s.lit("\tRG:Z:", 6); s.mem(P.rg, P.rg_len); }
    if (!(flag & 0x100)) {
        // other records of the read that are not flagged 0x100 (the flags a record carries in `list`: before this call's own bits)
        bool any = false;
        int wi = 0;
        for (int k = 0; k < n && !any; ++k) {
            if (!(P.sel[b + k].kind & 1)) continue;
            if (wi != which && !((P.aln[b + k].flag | extra_flag) & 0x100)) any = true;
            ++wi;
        }
        if (any) {
            s.lit("\tSA:Z:", 6);
            wi = 0;
            for (int k = 0; k < n; ++k) {
                if (!(P.sel[b + k].kind & 1)) continue;
                const int me = wi++;
                if (me == which) continue;
                const SdAln q = sd_load(P, b + k);
                if ((q.flag | extra_flag) & 0x100) continue;
                sd_name(s, P, q.rid); s.ch(',');
                sd_int(s, q.pos + 1); s.ch(',');
                s.ch("+-"[q.is_rev]); s.ch(',');
                for (int c = 0; c < q.n_cigar; ++c) { sd_int(s, q.cig[c] >> 4); s.ch("MIDSH"[q.cig[c] & 0xf]); }
                s.ch(','); sd_int(s, q.mapq);
                s.ch(','); sd_int(s, q.NM);
                s.ch(';');
            }
        }
        if (p.alt_sc > 0) { s.lit("\tpa:f:", 6); sd_fixed3(s, p.score, p.alt_sc); }
    }
    if (g >= 0 && n > 1) {                                 // XA / XB: the entries mem_gen_alt appends to this record's string, in region order
        bool first = true;
        const int me = (int)(g - b);
        for (int i = 0; i < n; ++i) {
            const b200_sel_t e = P.sel[b + i];
            if (!(e.kind & 2) || e.xa_of != me) continue;
            if (first) { s.lit((opt.flag & SD_XB) ? "\tXB:Z:" : "\tXA:Z:", 6); first = false; }
            const SdAln t = sd_load(P, b + i);
            sd_name(s, P, t.rid); s.ch(','); s.ch("+-"[t.is_rev]); sd_int(s, t.pos + 1); s.ch(',');
            for (int c = 0; c < t.n_cigar; ++c) { sd_int(s, t.cig[c] >> 4); s.ch("MIDSHN"[t.cig[c] & 0xf]); }
            s.ch(','); sd_int(s, t.NM);
            if (opt.flag & SD_XB) { s.ch(','); sd_int(s, t.score); }
            s.ch(';');
        }
    }
    if (P.comments) {
        const i64 c0 = P.comment_off[r], c1 = P.comment_off[r + 1];
        if (c1 > c0) { s.ch('\t'); s.mem(P.comments + c0, c1 - c0); }
    }
    if ((opt.flag & SD_REF_HDR) && p.rid >= 0 && P.ann_annos) {
        const int a0 = P.ann_anno_off[p.rid], a1 = P.ann_anno_off[p.rid + 1];
        if (a1 > a0) {
            s.lit("\tXR:Z:", 6);
            const i64 t = s.l;
            s.mem(P.ann_annos + a0, a1 - a0);
            s.tabs_to_spaces(t);
        }
    }
    s.ch('\n');
}

template <class S>
__device__ void sd_read(S &s, const SamDevParams &P, int rc /* read of the chunk */)
{
    const int r = P.r_lo + rc;
    const i64 b = P.roff[r] - P.g0;
    const int n = (int)(P.roff[r + 1] - P.roff[r]);
    bool has_mate = false;
    SdAln mate = sd_unmapped();
    if (P.mate_reg) {
        has_mate = true;
        const i32 mg = P.mate_reg[r];
        if (mg >= 0) { mate = sd_load(P, mg); }
    }
    int n_rec = 0;
    for (int k = 0; k < n; ++k) n_rec += P.sel[b + k].kind & 1;
    if (n_rec == 0) {
        const int xf = P.read_flag ? P.read_flag[r] : 0;
        sd_record(s, P, r, b, 0, -1, 0, xf, has_mate, mate);
        return;
    }
    int which = 0;
    for (int k = 0; k < n; ++k) {
        if (!(P.sel[b + k].kind & 1)) continue;
        sd_record(s, P, r, b, n, b + k, which, 0, has_mate, mate);
        ++which;
    }
}

__global__ void __launch_bounds__(128) sam_count_kernel(SamDevParams P)
{
    for (int rc = blockIdx.x * blockDim.x + threadIdx.x; rc < P.n_reads; rc += gridDim.x * blockDim.x) {
        SdCount c;
        sd_read(c, P, rc);
        P.len[rc] = (i32)c.l;
        if (c.l > 0x7fffffff) atomicMax(P.overflow, 1);
    }
}

__global__ void __launch_bounds__(128) sam_write_kernel(SamDevParams P)
{
    for (int rc = blockIdx.x * blockDim.x + threadIdx.x; rc < P.n_reads; rc += gridDim.x * blockDim.x) {
        SdWrite w(P.out + P.pos[rc]);
        sd_read(w, P, rc);
        if (w.l != P.pos[rc + 1] - P.pos[rc]) atomicMax(P.overflow, 2);       // the two passes disagree: a bug, never silent
    }
}

}  // namespace

int launch_sam_count(const SamDevParams &P, int n_sms, cudaStream_t st)
{
    if (P.n_reads <= 0) return 0;
    const int grid = std::max(1, std::min(n_sms * 16, (P.n_reads + 127) / 128));
    B200_LAUNCH(sam_count_kernel, grid, 128, 0, st, P);
    B200_CHECK(cudaGetLastError());
    return 0;
}

int launch_sam_write(const SamDevParams &P, int n_sms, cudaStream_t st)
{
    if (P.n_reads <= 0) return 0;
    const int grid = std::max(1, std::min(n_sms * 16, (P.n_reads + 127) / 128));
    B200_LAUNCH(sam_write_kernel, grid, 128, 0, st, P);
    B200_CHECK(cudaGetLastError());
    return 0;
}
