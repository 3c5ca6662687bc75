// Device FM-index: Occ blocks, bidirectional extension, SA lookup.
//
// Follows the arithmetic of the reference's bwt_occ4 / bwt_2occ4 (bwt.c:169-220), bwt_extend
// (bwt.c:262-275), bwt_invPsi / bwt_sa (bwt.c:53-59,86-96) -- counts must be identical -- but the
// mechanics are GPU-first: the 64-byte checkpoint block (4 x u64 counts + 128 bases of 2-bit BWT,
// layout from bwtindex.c:149-172) is fetched with four 128-bit read-only loads from a 64-byte-aligned
// address (one DRAM sector pair), and the per-base byte table (cnt_table) is replaced by popcounts
// on the two bit-planes of each 32-base word.
#pragma once
#include "simt.h"

struct DevIndex {
    const uint4 *bwt;        // 4 x uint4 per 64-byte block
    u64 primary, seq_len;
    u64 L2[5];
    const u64 *sa;           // sampled SA (every sa_intv-th row)
    u64 n_sa;
    int sa_intv;
    const u32 *sa_full32;    // optional full SA (seq_len+1 entries) when seq_len < 2^32
    const u64 *sa_full64;    // optional full SA otherwise (transient: b200_pack_full_sa turns it into sa_full40)
    const u32 *sa_full40;    // full SA as 5 bytes per row, little-endian, rows packed back to back (seq_len < 2^40); row 0 is special-cased
    const u8 *pac;           // 2-bit packed forward reference
    i64 l_pac;
    int n_seqs;
    const i64 *ann_off;      // contig offsets (n_seqs)
    const i32 *ann_len;
    const i32 *ann_alt;
    // re-laid Occ index for the seeding kernel (built once on the device, see occ32.cuh)
    const uint4 *occ32;      // 2 x uint4 per 64-row block
    u64 sb_cnt[8][4];        // counts at the start of each 2^32-row superblock (rows < 2^35)
    u64 sb_gt[8][4];         // sb_gt[s][c] = sum of sb_cnt[s][d] over d > c (filled by b200_index_finish)
    // bi-intervals of all strings of length 1..kmer_k (see kmer_tab.cuh); 0 = no table
    const uint4 *kmer_tab;   // entry of string W (length k): kmer_tab[(kmer_off[k] + W) * quads], quads = 1 (u32) / 2 (u64)
    int kmer_k;
    u32 kmer_off[16];        // depth <= 15: string codes and the virtual-entry masks of smem3.cu are 32-bit
};

struct Intv { u64 x0, x1, x2, info; };   // == bwtintv_t (bwt.h:62)

// -------------------------------------------------------------------------------------------------
// one Occ block in registers
struct OccBlock { uint4 c01, c23, b0, b1; };

__device__ __forceinline__ OccBlock load_block(const uint4 *bwt, u64 k /* $-adjusted row */)
{
    const uint4 *p = bwt + ((k >> 7) << 2);
    OccBlock r;
    r.c01 = __ldg(p); r.c23 = __ldg(p + 1); r.b0 = __ldg(p + 2); r.b1 = __ldg(p + 3);
    return r;
}

__device__ __forceinline__ u64 u64_of(u32 lo, u32 hi) { return (u64)hi << 32 | lo; }

// counts of A,C,G,T among bases [0, off] (inclusive) of the block + its checkpoint
__device__ __forceinline__ void block_occ4(const OccBlock &b, int off /* 0..127 */, u64 cnt[4])
{
    // the reference reads uint32 words p[0..7], base j of a word at bits 30-2j (bwt_B0, bwt.h:80);
    // pair them as 32-base words with base 0 at the top.
    u64 w[4] = { u64_of(b.b0.y, b.b0.x), u64_of(b.b0.w, b.b0.z), u64_of(b.b1.y, b.b1.x), u64_of(b.b1.w, b.b1.z) };
    int nA = 0, nC = 0, nG = 0, nT = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        int nb = off + 1 - 32 * j;                 // bases of this word that are counted
        nb = nb < 0 ? 0 : (nb > 32 ? 32 : nb);
        u64 m = nb == 0 ? 0ull : ((~0ull) << (64 - 2 * nb)) & 0x5555555555555555ull;
        u64 lo = w[j] & m, hi = (w[j] >> 1) & m;
        int t = __popcll(hi & lo), ph = __popcll(hi), pl = __popcll(lo);
        nT += t; nG += ph - t; nC += pl - t; nA += nb - ph - pl + t;
    }
    cnt[0] = u64_of(b.c01.x, b.c01.y) + nA;
    cnt[1] = u64_of(b.c01.z, b.c01.w) + nC;
    cnt[2] = u64_of(b.c23.x, b.c23.y) + nG;
    cnt[3] = u64_of(b.c23.z, b.c23.w) + nT;
}

// bwt_occ4 (bwt.c:169-186)
__device__ __forceinline__ void dev_occ4(const DevIndex &ix, u64 k, u64 cnt[4])
{
    if (k == (u64)-1) { cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0; return; }
    k -= (k >= ix.primary);
    OccBlock b = load_block(ix.bwt, k);
    block_occ4(b, (int)(k & 127), cnt);
}

// bwt_2occ4 (bwt.c:189-220): same results as two bwt_occ4 calls; one block fetch when both rows
// fall into the same 128-base block.  *two_blocks reports which case (for the traffic counters).
__device__ __forceinline__ void dev_2occ4(const DevIndex &ix, u64 k, u64 l, u64 ck[4], u64 cl[4], int *two_blocks)
{
    if (k == (u64)-1 || l == (u64)-1) {
        dev_occ4(ix, k, ck); dev_occ4(ix, l, cl);
        if (two_blocks) *two_blocks = 1;
        return;
    }
    u64 k2 = k - (k >= ix.primary), l2 = l - (l >= ix.primary);
    OccBlock bk = load_block(ix.bwt, k2);
    block_occ4(bk, (int)(k2 & 127), ck);
    if ((k2 >> 7) == (l2 >> 7)) {
        block_occ4(bk, (int)(l2 & 127), cl);
        if (two_blocks) *two_blocks = 0;
    } else {
        OccBlock bl = load_block(ix.bwt, l2);
        block_occ4(bl, (int)(l2 & 127), cl);
        if (two_blocks) *two_blocks = 1;
    }
}

// bwt_extend (bwt.c:262-275).  ok[c] for c = 0..3
__device__ __forceinline__ void dev_extend(const DevIndex &ix, u64 x0, u64 x1, u64 x2, int is_back,
                                           u64 o0[4], u64 o1[4], u64 o2[4], int *two_blocks)
{
    u64 tk[4], tl[4];
    u64 a = is_back ? x0 : x1;      // x[!is_back]
    u64 b = is_back ? x1 : x0;      // x[is_back]
    dev_2occ4(ix, a - 1, a - 1 + x2, tk, tl, two_blocks);
    u64 na[4], nb[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) { na[c] = ix.L2[c] + 1 + tk[c]; o2[c] = tl[c] - tk[c]; }
    nb[3] = b + (a <= ix.primary && a + x2 - 1 >= ix.primary);
    nb[2] = nb[3] + o2[3];
    nb[1] = nb[2] + o2[2];
    nb[0] = nb[1] + o2[1];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        o0[c] = is_back ? na[c] : nb[c];
        o1[c] = is_back ? nb[c] : na[c];
    }
}

// bwt_set_intv (bwt.h:82)
__device__ __forceinline__ void dev_set_intv(const DevIndex &ix, int c, u64 &x0, u64 &x1, u64 &x2)
{
    x0 = ix.L2[c] + 1; x2 = ix.L2[c + 1] - ix.L2[c]; x1 = ix.L2[3 - c] + 1;
}

// bwt_occ (bwt.c:107-129) for one symbol, from a block in registers
__device__ __forceinline__ u64 block_occ1(const OccBlock &b, int off, int c)
{
    u64 cnt[4];
    block_occ4(b, off, cnt);
    return cnt[c];
}

// one LF step: bwt_invPsi (bwt.c:53-59)
__device__ __forceinline__ u64 dev_invpsi(const DevIndex &ix, u64 k)
{
    if (k == ix.primary) return 0;
    u64 x = k - (k > ix.primary);
    OccBlock b = load_block(ix.bwt, x);
    int off = (int)(x & 127);
    u32 wd[8] = { b.b0.x, b.b0.y, b.b0.z, b.b0.w, b.b1.x, b.b1.y, b.b1.z, b.b1.w };
    u32 word = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) if ((off >> 4) == j) word = wd[j];
    int c = (word >> ((~off & 15) << 1)) & 3;                       // bwt_B0 (bwt.h:80)
    // bwt_occ(bwt, k, c): k == seq_len and k == -1 cannot happen for a valid row 0 < k <= seq_len
    // other than k == seq_len, handled like the reference (bwt.c:112)
    u64 occ;
    if (k == ix.seq_len) occ = ix.L2[c + 1] - ix.L2[c];
    else {
        u64 kk = k - (k >= ix.primary);
        if ((kk >> 7) == (x >> 7)) occ = block_occ1(b, (int)(kk & 127), c);
        else { OccBlock b2 = load_block(ix.bwt, kk); occ = block_occ1(b2, (int)(kk & 127), c); }
    }
    return ix.L2[c] + occ;
}

// bwt_sa (bwt.c:86-96): walk to the next sampled row.  *steps gets the number of LF steps.
__device__ __forceinline__ u64 dev_sa_walk(const DevIndex &ix, u64 k, int *steps)
{
    u64 sa = 0, mask = (u64)ix.sa_intv - 1;
    while (k & mask) { ++sa; k = dev_invpsi(ix, k); }
    if (steps) *steps = (int)sa;
    return sa + ix.sa[k / ix.sa_intv];
}

__device__ __forceinline__ u64 dev_sa(const DevIndex &ix, u64 k, int *steps)
{
    if (ix.sa_full32) { if (steps) *steps = 0; return k == 0 ? (u64)-1 : (u64)ix.sa_full32[k]; }
    if (ix.sa_full40) {                                             // 40 bits at byte 5k: always inside two aligned words
        if (steps) *steps = 0;
        if (k == 0) return (u64)-1;                                 // bwt->sa[0] = (bwtint_t)-1 (bwt.c:71)
        const u64 b = k * 5;
        const u32 *w = ix.sa_full40 + (b >> 2);
        return (((u64)__ldg(w + 1) << 32 | __ldg(w)) >> ((int)(b & 3) * 8)) & 0xffffffffffull;
    }
    if (ix.sa_full64) { if (steps) *steps = 0; return ix.sa_full64[k]; }
    return dev_sa_walk(ix, k, steps);
}

// ---- reference sequence access: bns_depos (bntseq.h:87), bns_pos2rid (bntseq.c:354), bns_intv2rid (:370)
__device__ __forceinline__ int dev_pos2rid(const DevIndex &ix, i64 pos_f)
{
    if (pos_f >= ix.l_pac) return -1;
    int left = 0, mid = 0, right = ix.n_seqs;
    while (left < right) {
        mid = (left + right) >> 1;
        if (pos_f >= ix.ann_off[mid]) {
            if (mid == ix.n_seqs - 1) break;
            if (pos_f < ix.ann_off[mid + 1]) break;
            left = mid + 1;
        } else right = mid;
    }
    return mid;
}
__device__ __forceinline__ i64 dev_depos(const DevIndex &ix, i64 pos, int *is_rev)
{
    *is_rev = pos >= ix.l_pac;
    return *is_rev ? (ix.l_pac << 1) - 1 - pos : pos;
}
__device__ __forceinline__ int dev_intv2rid(const DevIndex &ix, i64 rb, i64 re)
{
    if (rb < ix.l_pac && re > ix.l_pac) return -2;
    int is_rev;
    int rid_b = dev_pos2rid(ix, dev_depos(ix, rb, &is_rev));
    int rid_e = rb < re ? dev_pos2rid(ix, dev_depos(ix, re - 1, &is_rev)) : rid_b;
    return rid_b == rid_e ? rid_b : -1;
}
// base at position p of the forward+reverse-complement text (bns_get_seq, bntseq.c:403-424)
__device__ __forceinline__ int dev_ref_base(const DevIndex &ix, i64 p)
{
    if (p >= ix.l_pac) {
        i64 f = (ix.l_pac << 1) - 1 - p;
        return 3 - ((ix.pac[f >> 2] >> ((~f & 3) << 1)) & 3);
    }
    return (ix.pac[p >> 2] >> ((~p & 3) << 1)) & 3;
}
