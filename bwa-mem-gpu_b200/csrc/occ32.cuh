// "occ32": the Occ/BWT index re-laid for one-child bidirectional extension on the GPU.
//
// The reference block (bwtindex.c:149-172, bwt.h:73-80) is 64 bytes per 128 rows: 4 x u64 checkpoint counts +
// 32 bytes of interleaved 2-bit symbols, and bwt_extend (bwt.c:262-275) derives all four children from eight
// popcount-table sweeps.  Every caller on the seeding path (bwt_smem1a bwt.c:308-341, bwt_seed_strategy1
// bwt.c:366-375) uses exactly ONE child ok[c], for which only two numbers per interval end are needed:
// occ(c, k) and the number of symbols greater than c before k.  occ32 stores, per 64 rows, ONE 32-byte DRAM
// sector:  u32 cnt[4] (A,C,G,T before the block, relative to the 2^32-row superblock) + the low-bit plane and
// the high-bit plane of the 64 symbols as two u64.  With planes, "== c" and "> c" are single boolean
// expressions, so one end costs two 64-bit popcounts instead of a byte-table sweep, and a lookup touches
// 32 bytes (64 when both ends fall in different blocks) instead of 64-128.  Only the resulting interval
// values must equal the reference's -- they do (tests: occ32 child == bwt_extend child, bit for bit).
#pragma once
#include "fmindex.cuh"

struct Occ32Blk { uint4 cnt; uint4 pl; };     // pl = {lo.lo32, lo.hi32, hi.lo32, hi.hi32}

__device__ __forceinline__ Occ32Blk occ32_load(const uint4 *occ32, u64 blk)
{
    const uint4 *p = occ32 + (blk << 1);
    Occ32Blk b;
    b.cnt = __ldg(p); b.pl = __ldg(p + 1);
    return b;
}

// Both sectors of one extension in ONE issue group: the four 128-bit loads are emitted back to back so that the
// extension costs one memory round trip (ptxas otherwise sinks the second sector's loads below the first
// sector's arithmetic, which serialises two round trips).  When bl == bk the second pair hits the sector in flight.
__device__ __forceinline__ void occ32_load2p(const uint4 *pk, const uint4 *pl, Occ32Blk &K, Occ32Blk &L)
{
#ifdef B200_EMU
    K.cnt = pk[0]; K.pl = pk[1]; L.cnt = pl[0]; L.pl = pl[1];
#else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u64 p, %16, %17;\n\t"
                 "ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%16];\n\t"
                 "@p ld.global.nc.v8.u32 {%8,%9,%10,%11,%12,%13,%14,%15}, [%17];\n\t}"
                 : "=r"(K.cnt.x), "=r"(K.cnt.y), "=r"(K.cnt.z), "=r"(K.cnt.w), "=r"(K.pl.x), "=r"(K.pl.y), "=r"(K.pl.z), "=r"(K.pl.w),
                   "=r"(L.cnt.x), "=r"(L.cnt.y), "=r"(L.cnt.z), "=r"(L.cnt.w), "=r"(L.pl.x), "=r"(L.pl.y), "=r"(L.pl.z), "=r"(L.pl.w)
                 : "l"(pk), "l"(pl));
    if (pk == pl) L = K;
#endif
}
__device__ __forceinline__ void occ32_load2(const uint4 *occ32, u64 bk, u64 bl, Occ32Blk &K, Occ32Blk &L)
{
    occ32_load2p(occ32 + (bk << 1), occ32 + (bl << 1), K, L);
}

// (# of symbol c, # of symbols > c) among rows [block start, block start + off] of a block, plus checkpoint
__device__ __forceinline__ void occ32_pair(const Occ32Blk &b, int off, int c, u32 &n_eq, u32 &n_gt)
{
    const u32 m1 = (u32)-(c >> 1), m0 = (u32)-(c & 1);             // all-ones when the bit of c is set
    // 64-bit inclusive mask of rows 0..off (off = 63 shifts the single zero bit out: all ones)
    const u64 m64 = ~(~1ull << off);
    const u32 mlo = (u32)m64, mhi = (u32)(m64 >> 32);
    const u32 lo0 = b.pl.x, lo1 = b.pl.y, hi0 = b.pl.z, hi1 = b.pl.w;
    const u32 e0 = ~(hi0 ^ m1) & ~(lo0 ^ m0) & mlo, e1 = ~(hi1 ^ m1) & ~(lo1 ^ m0) & mhi;
    const u32 w0 = lo0 & ~m0, w1 = lo1 & ~m0;
    const u32 g0 = (m1 ? (hi0 & w0) : (hi0 | w0)) & mlo, g1 = (m1 ? (hi1 & w1) : (hi1 | w1)) & mhi;
    const u32 c0 = b.cnt.x, c1 = b.cnt.y, c2 = b.cnt.z, c3 = b.cnt.w;
    const u32 ceq = c & 2 ? (c & 1 ? c3 : c2) : (c & 1 ? c1 : c0);
    const u32 cgt = c & 2 ? (c & 1 ? 0u : c3) : (c & 1 ? c2 + c3 : c1 + c2 + c3);
    n_eq = ceq + __popc(e0) + __popc(e1);
    n_gt = cgt + __popc(g0) + __popc(g1);
}

// Child c of a bi-interval: ok[c] of bwt_extend (bwt.c:262-275).  PT = u32 when seq_len < 2^32 - 2, else u64.
// Split in two so that a caller can issue the two sector requests of the NEXT extension at the bottom of its
// loop and consume them at the top (the requests of all lanes are then in flight together and each extension
// costs one memory round trip).  *two_ref = 1 when the two interval ends fall into different 128-row blocks of
// the REFERENCE layout (what bwt_2occ4 would have fetched; the algorithmic-traffic counter of the roofline).
template <class PT>
__device__ __forceinline__ void occ32_issue(const DevIndex &ix, PT x0, PT x1, PT x2, int is_back, Occ32Blk &Bk, Occ32Blk &Bl)
{
    const PT primary = (PT)ix.primary;
    const PT a = is_back ? x0 : x1;
    const PT k = a - 1, l = k + x2;                                   // bwt.c:265
    const PT kp = k - (PT)(k >= primary), lp = l - (PT)(l >= primary); // bwt.c:192-193
    occ32_load2(ix.occ32, (u64)kp >> 6, (u64)lp >> 6, Bk, Bl);
}

// The two sector addresses of an extension without the loads: a caller whose lanes take their data from different
// tables (smem3.cu: k-mer table entry or index sectors) forms the addresses per lane and then issues ONE load group for
// the whole warp -- loads into the same registers from two sides of a branch would wait for each other.
template <class PT>
__device__ __forceinline__ void occ32_addr(const DevIndex &ix, PT x0, PT x1, PT x2, int is_back, const uint4 *&pk, const uint4 *&pl)
{
    const PT primary = (PT)ix.primary;
    const PT a = is_back ? x0 : x1;
    const PT k = a - 1, l = k + x2;
    const PT kp = k - (PT)(k >= primary), lp = l - (PT)(l >= primary);
    pk = ix.occ32 + (((u64)kp >> 6) << 1); pl = ix.occ32 + (((u64)lp >> 6) << 1);
}

// 1 when the reference's bwt_2occ4 (bwt.c:189-220) would have read two 128-row blocks for this extension
template <class PT>
__device__ __forceinline__ int occ32_two_ref(const DevIndex &ix, PT x0, PT x1, PT x2, int is_back)
{
    const PT primary = (PT)ix.primary;
    const PT a = is_back ? x0 : x1;
    const PT k = a - 1, l = k + x2;
    const PT kp = k - (PT)(k >= primary), lp = l - (PT)(l >= primary);
    return (int)(((u64)kp >> 7) != ((u64)lp >> 7));
}

template <class PT>
__device__ __forceinline__ void occ32_finish(const DevIndex &ix, PT x0, PT x1, PT x2, int is_back, int c,
                                             const Occ32Blk &Bk, const Occ32Blk &Bl, PT &n0, PT &n1, PT &n2, int &two_ref)
{
    const PT primary = (PT)ix.primary;
    const PT a = is_back ? x0 : x1, b = is_back ? x1 : x0;
    const PT k = a - 1, l = k + x2;
    const PT kp = k - (PT)(k >= primary), lp = l - (PT)(l >= primary);
    u32 ek, gk, el, gl;
    occ32_pair(Bk, (int)(kp & 63), c, ek, gk);
    occ32_pair(Bl, (int)(lp & 63), c, el, gl);
    two_ref = (int)(((u64)kp >> 7) != ((u64)lp >> 7));
    PT tk = ek, tl = el, sk = gk, sl = gl;
    if (sizeof(PT) == 8) {                                            // superblock bases (rows >= 2^32)
        const int s0 = (int)((u64)kp >> 32), s1 = (int)((u64)lp >> 32);
        tk += (PT)ix.sb_cnt[s0][c]; tl += (PT)ix.sb_cnt[s1][c]; sk += (PT)ix.sb_gt[s0][c]; sl += (PT)ix.sb_gt[s1][c];
    }
    n2 = tl - tk;                                                     // bwt.c:270
    const PT na = (PT)ix.L2[c] + 1 + tk;                              // bwt.c:269
    const PT nb = b + (PT)(a <= primary && a + x2 - 1 >= primary) + (sl - sk);   // bwt.c:271-274
    n0 = is_back ? na : nb;
    n1 = is_back ? nb : na;
}

template <class PT>
__device__ __forceinline__ void occ32_child(const DevIndex &ix, PT x0, PT x1, PT x2, int is_back, int c,
                                            PT &n0, PT &n1, PT &n2, int &two_ref)
{
    Occ32Blk Bk, Bl;
    occ32_issue<PT>(ix, x0, x1, x2, is_back, Bk, Bl);
    occ32_finish<PT>(ix, x0, x1, x2, is_back, c, Bk, Bl, n0, n1, n2, two_ref);
}
