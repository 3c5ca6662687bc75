// Stage 1, fast path: mem_collect_intv (bwamem.c:137-185) = three seeding passes built from bwt_smem1a
// (bwt.c:289-351) and bwt_seed_strategy1 (bwt.c:358-379), one read per lane over the occ32 index.
//
// Why this shape.  Per 150-bp read the reference issues ~700 DEPENDENT bwt_extend calls; each is one or two
// random index sectors followed by a little integer work.  The batch holds millions of independent reads, so
// the machine is filled across reads (one read per lane, persistent lanes pulling reads from an atomic
// counter) and the per-lane code is a flattened state machine with exactly ONE extension site per iteration:
// whatever phase the 32 reads of a warp are in (forward sweep, lock-step backward sweep, re-seeding,
// LAST-like sweep) all lanes meet at the gather, issue their sector loads together and do the same popcount
// arithmetic.  Everything around the gather is kept uniform and small:
//   * one register triple holds "the interval to extend next" in every phase and one index holds "the query
//     position to extend with", so forming the request is the same three instructions for all phases;
//   * only the child ok[c] that the algorithm consumes is computed (occ32.cuh);
//   * extensions of short strings (up to K bases, K = depth of the k-mer interval table, kmer_tab.cu) are answered by the
//     table; whatever a lane reads -- two index sectors, or the 32-byte sector holding a 16-byte table record -- all lanes
//     issue it through ONE load group (loads on the two sides of a branch into the same registers wait for each other);
//   * the read is staged once in SHARED memory as 4-bit codes ([word][thread] interleaved), so the per-step
//     base fetch is a conflict-free LDS instead of a global byte load;
//   * the forward change-point list lives in SHARED memory as a ring of the last LCAP entries ([slot][thread]
//     interleaved 16/32-byte records, conflict-free 128-bit accesses); older entries spill to a per-thread
//     global area; the backward sweep compacts the list IN PLACE (survivors overwrite consumed entries);
//   * list entries whose string is still inside the table's range are BITS, not records ("virtual" entries, see vm below),
//     and all of them at one backward position are settled in one iteration from the size-change mask of the window's
//     depth-K record (bwt.c:339 needs only "same size as the next longer entry?");
//   * positions are 32-bit when seq_len < 2^32 (template PT), 64-bit otherwise;
//   * chain restarts (pass-1 continuation, pass-3 restart) are inlined after the consume step; only pass
//     changes, pass-2 candidate scans and read changes take the slow path;
//   * the bwt_extend / two-block counters of the roofline are a template flag (STATS), off in production launches.
// The emitted interval set equals the reference's; records are determined by (start,end), so after the sort
// by `info` (bwamem.c:184) the order is unique.
#include "b200_internal.h"
#include "occ32.cuh"

#define S3_THREADS 128

enum { S3_TASK = 0, S3_FWD = 1, S3_BWD = 2, S3_P3 = 3, S3_FIN = 4 };

// List entry / k-mer table entry formats.  PT = position type of the kernel's arithmetic.
//   Ent<u32>  : seq_len < 2^32, one 16-byte record (x0, x1, x2, end)
//   Ent<u64>  : any seq_len, two 16-byte records
//   Ent33     : 2^32 <= seq_len < 2^33 (the 3.1 Gbp configs: 6.2 G rows), ONE 16-byte record: the low words of x0, x1, x2
//               and `end` (< 2^29) with the three bit-32s in its top bits.  Half the shared memory of Ent<u64>: 6 resident
//               blocks per SM instead of 3 on the human-sized index, and one LDG.128 per k-mer table hit instead of two.
template <class PT> struct Ent;
template <> struct Ent<u32> {
    typedef u32 PT;
    static constexpr int kQuads = 1, kMinBlocks = 0;          // 0 = no bound (ptxas picks 80 registers, no spill)
    static __device__ __forceinline__ void st(uint4 *p, int stride, u32 x0, u32 x1, u32 x2, u32 end)
    { p[0] = make_uint4(x0, x1, x2, end); (void)stride; }
    static __device__ __forceinline__ void ld(const uint4 *p, int stride, u32 &x0, u32 &x1, u32 &x2, u32 &end)
    { const uint4 v = p[0]; x0 = v.x; x1 = v.y; x2 = v.z; end = v.w; (void)stride; }
    static __device__ __forceinline__ void tab(const uint4 &a, const uint4 &, u32 &x0, u32 &x1, u32 &x2) { x0 = a.x; x1 = a.y; x2 = a.z; }
    static __device__ __forceinline__ u32 tabmask(const uint4 &a, const uint4 &) { return a.w; }      // prefix size-change mask (kmer_tab.cu)
};
template <> struct Ent<u64> {
    typedef u64 PT;
    static constexpr int kQuads = 2, kMinBlocks = 0;
    static __device__ __forceinline__ void st(uint4 *p, int stride, u64 x0, u64 x1, u64 x2, u32 end)
    {
        p[0] = make_uint4((u32)x0, (u32)(x0 >> 32), (u32)x1, (u32)(x1 >> 32));
        p[stride] = make_uint4((u32)x2, (u32)(x2 >> 32), end, 0);
    }
    static __device__ __forceinline__ void ld(const uint4 *p, int stride, u64 &x0, u64 &x1, u64 &x2, u32 &end)
    {
        const uint4 a = p[0], b = p[stride];
        x0 = (u64)a.y << 32 | a.x; x1 = (u64)a.w << 32 | a.z; x2 = (u64)b.y << 32 | b.x; end = b.z;
    }
    static __device__ __forceinline__ void tab(const uint4 &a, const uint4 &b, u64 &x0, u64 &x1, u64 &x2)
    { x0 = (u64)a.y << 32 | a.x; x1 = (u64)a.w << 32 | a.z; x2 = (u64)b.y << 32 | b.x; }
    static __device__ __forceinline__ u32 tabmask(const uint4 &, const uint4 &b) { return b.z; }
};
struct Ent33 {
    typedef u64 PT;
#ifndef S3_MINB33
#define S3_MINB33 6
#endif
    static constexpr int kQuads = 1, kMinBlocks = S3_MINB33;          // 92 registers unbounded; 80 (6 resident blocks per SM) measured faster
    static __device__ __forceinline__ uint4 pack(u64 x0, u64 x1, u64 x2, u32 end)
    { return make_uint4((u32)x0, (u32)x1, (u32)x2, end | (u32)(x0 >> 32) << 31 | ((u32)(x1 >> 32) & 1u) << 30 | ((u32)(x2 >> 32) & 1u) << 29); }
    static __device__ __forceinline__ void unpack(const uint4 &v, u64 &x0, u64 &x1, u64 &x2, u32 &end)
    { x0 = (u64)(v.w >> 31) << 32 | v.x; x1 = (u64)(v.w >> 30 & 1u) << 32 | v.y; x2 = (u64)(v.w >> 29 & 1u) << 32 | v.z; end = v.w & 0x1fffffffu; }
    static __device__ __forceinline__ void st(uint4 *p, int stride, u64 x0, u64 x1, u64 x2, u32 end) { p[0] = pack(x0, x1, x2, end); (void)stride; }
    static __device__ __forceinline__ void ld(const uint4 *p, int stride, u64 &x0, u64 &x1, u64 &x2, u32 &end) { unpack(p[0], x0, x1, x2, end); (void)stride; }
    static __device__ __forceinline__ void tab(const uint4 &a, const uint4 &, u64 &x0, u64 &x1, u64 &x2) { u32 e; unpack(a, x0, x1, x2, e); }
    static __device__ __forceinline__ u32 tabmask(const uint4 &a, const uint4 &) { return a.w & 0xffffu; }
};

__device__ __forceinline__ void s3_emit(Intv *out, int &n_out, int cap_out, u64 x0, u64 x1, u64 x2, int start, int end)
{
    if (n_out < cap_out) {
        uint4 *q = (uint4 *)&out[n_out];                   // streaming stores: keep the index resident in L2
        __stcs(q, make_uint4((u32)x0, (u32)(x0 >> 32), (u32)x1, (u32)(x1 >> 32)));
        __stcs(q + 1, make_uint4((u32)x2, (u32)(x2 >> 32), (u32)end, (u32)start));
    }
    ++n_out;
}

// ASCII / nt4 bytes -> nt4 in place (nst_nt4_table, bntseq.c:40-57; mem_align1_core, bwamem.c:1067-1068) and a
// 4-bit packed copy, `stride` words per read, for the seeding kernel.
__global__ void pack4_kernel(u8 *seq, const i64 *off, int n_reads, u32 *pk, int stride)
{
    const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const i64 o = off[r];
    const int len = (int)(off[r + 1] - o);
    u32 *dst = pk + (size_t)r * stride;
    for (int w = (int)lane; w < stride; w += 32) {
        u32 v = 0;
        for (int k = 0; k < 8; ++k) {
            const int i = w * 8 + k;
            u32 code = 4;
            if (i < len) {
                u8 c = seq[o + i];
                if (c >= 4) {
                    const u8 u = c & 0xdf;
                    c = u == 'A' ? 0 : u == 'C' ? 1 : u == 'G' ? 2 : u == 'T' ? 3 : (c == '-' ? 5 : 4);
                    seq[o + i] = c;
                }
                code = c < 4 ? c : 4;
            }
            v |= code << (4 * k);
        }
        dst[w] = v;
    }
}

// QGLOB: the read is too long to stage in shared memory (long reads): its 4-bit codes are fetched from the packed
// copy in global memory instead (consecutive positions of a lane share a 32-byte sector, so these are L1 hits).
// STATS: count the bwt_extend calls and how many of them the reference's bwt_2occ4 would have served from two blocks (the
// roofline's algorithmic traffic, a property of the input).  About a tenth of the kernel's instructions, so production
// launches leave it out (SmemParams::count_stats, tunable seed_stats); bench.py counts in its warm-up pass.
template <class E, int LCAP, bool SPLIT, bool QGLOB, bool STATS>                // SPLIT: pass 3 ran in smem_p3_kernel
__global__ void __launch_bounds__(S3_THREADS, E::kMinBlocks) smem_fast3_kernel(SmemParams P)
{
    B200_DYN_SMEM(s3_raw);
    typedef typename E::PT PT;
    constexpr int EQ = E::kQuads;
    uint4 *L = (uint4 *)s3_raw + threadIdx.x;                            // slot s: L[s * EQ * S3_THREADS]
    u32 *Q = (u32 *)((uint4 *)s3_raw + LCAP * EQ * S3_THREADS) + threadIdx.x;   // word w: Q[w * S3_THREADS]
    const u32 *Qg = P.pk;                                                // QGLOB: packed codes of the current read
    const size_t gtid = (size_t)blockIdx.x * S3_THREADS + threadIdx.x;
    uint4 *SP = (uint4 *)P.lists + gtid * (size_t)P.spill_cap * EQ;      // spilled entry e: SP[e * EQ]
    const DevIndex &ix = P.ix;
    const int min_seed_len = P.min_seed_len;
    const PT max_mem_intv = (PT)P.max_mem_intv;
    u64 n_ext = 0, n_two = 0;

    Intv *out = nullptr;
    int len = 0, r = -1, n_out = 0;
    int st = S3_FIN, pass = 0, x = 0, k2 = 0, old_n = 0;
    int i = -1, sx = 0, nfw = 0, base = 0, nprev = 0, ncurr = 0, j = 0, last_start = 0, ret = 0;
    PT min_intv = 1, k0 = 0, k1 = 0, kk2 = 0, last_x2 = 0;              // (k0,k1,kk2): the interval to extend next
    u32 pend = 0;
    // k-mer table (kmer_tab.cu): strings of up to KT bases are answered by one table load instead of two index sectors.
    // wf / flen: code and length of the current forward chain; wst: its first KT bases, left-aligned; wb: the KT bases
    // that start at the backward sweep's position i + 1, left-aligned (first base most significant everywhere).
    const int KT = ix.kmer_k;
    const uint4 *const ktab = ix.kmer_tab;
    u32 wf = 0, wst = 0, wb = 0;
    int flen = 0;
    int ptab = 0;
    // "Virtual" list entries.  A change point of the forward sweep whose string is shorter than KT bases is extended
    // backwards through the k-mer table for its first KT - length steps: those extensions are looked up by STRING, the
    // entry's own interval is never read, and it cannot be emitted (an emitted match has >= min_seed_len > KT bases,
    // bwamem.c:150).  Such an entry is therefore only a bit: bit L of `vm` = the entry with end sx + L.  In the forward
    // sweep vm collects them; in the backward sweep vm holds those of the current position still to be visited (after
    // the real entries, longest first), vnext the survivors for the next position, and an entry whose child reaches KT
    // bases is written to the real list (it is shorter than every real entry, so the order of bwt.c:326-341 is kept).
    // Virtual entries still take a list index in the forward sweep (they are the first ones), so that the slots below the
    // real entries exist for the ones that become real.
    // vtwo: bit L = the reference's bwt_2occ4 would read two blocks for that entry's next extension (traffic counter).
    const int KTV = KT < min_seed_len ? KT : 0;
    u32 vm = 0, vnext = 0, vtwo = 0;

#define S3_QGET(pos) (((QGLOB ? __ldg(Qg + ((pos) >> 3)) : Q[((pos) >> 3) * S3_THREADS]) >> (((pos) & 7) << 2)) & 7u)
#define S3_SLOT(e) (L + ((e) % LCAP) * EQ * S3_THREADS)
#define S3_PUSH(endv)                                                                                   \
    do {                                                                                                \
        if (nfw - LCAP >= (int)__popc(vm)) {               /* evict entry nfw - LCAP (a real one) to the spill area */ \
            PT t0, t1, t2; u32 te;                                                                      \
            E::ld(S3_SLOT(nfw), S3_THREADS, t0, t1, t2, te);                                            \
            if (nfw - LCAP < P.spill_cap) E::st(SP + (size_t)(nfw - LCAP) * EQ, 1, t0, t1, t2, te);     \
        }                                                                                               \
        E::st(S3_SLOT(nfw), S3_THREADS, k0, k1, kk2, (u32)(endv));                                      \
        ++nfw;                                                                                          \
    } while (0)
#define S3_LOAD_ENTRY(jj)                                                                               \
    do {                                                                                                \
        const int e_ = base - (jj);                                                                     \
        if ((jj) < LCAP) E::ld(S3_SLOT(e_), S3_THREADS, k0, k1, kk2, pend);                             \
        else E::ld(SP + (size_t)e_ * EQ, 1, k0, k1, kk2, pend);                                         \
    } while (0)
#define S3_STORE_ENTRY(jj, a0, a1, a2, ev)                                                              \
    do {                                                                                                \
        const int e_ = base - (jj);                                                                     \
        if ((jj) < LCAP) E::st(S3_SLOT(e_), S3_THREADS, a0, a1, a2, ev);                                \
        else E::st(SP + (size_t)e_ * EQ, 1, a0, a1, a2, ev);                                            \
    } while (0)
#define S3_NEXT_ENTRY()                                /* real entry j, else the longest virtual one left */ \
    do {                                                                                                \
        if (j < nprev) S3_LOAD_ENTRY(j);                                                                \
        else { const int lb_ = 31 - __clz((int)vm); vm ^= 1u << lb_; pend = (u32)(sx + lb_); }          \
    } while (0)
#define S3_PUSH_FWD(endv)                                                                               \
    do {                                                                                                \
        if (flen < KTV) {                                  /* takes list index nfw, stores nothing */    \
            vm |= 1u << flen; ++nfw;                                                                    \
            if (STATS) vtwo = (vtwo & ~(1u << flen)) | (u32)occ32_two_ref<PT>(ix, k0, k1, kk2, 1) << flen; \
        } else S3_PUSH(endv);                                                                           \
    } while (0)
#define S3_START_BWD()                                                                                  \
    do {                                                                                                \
        ret = i; base = nfw - 1; nprev = nfw - (int)__popc(vm); ncurr = 0; j = 0; i = sx - 1;           \
        last_start = 0x7fffffff; st = S3_BWD; wb = wst; vnext = 0;                                      \
        S3_NEXT_ENTRY();                                                                                \
    } while (0)
#define S3_SET_INTV(b0)                                                                                 \
    do {                                                                                                \
        k0 = (PT)ix.L2[b0] + 1; kk2 = (PT)(ix.L2[(b0) + 1] - ix.L2[b0]); k1 = (PT)ix.L2[3 - (b0)] + 1;  \
        wf = (u32)(b0); flen = 1; wst = KT > 0 ? (u32)(b0) << (2 * KT - 2) : 0u;                        \
    } while (0)

    Occ32Blk Bk, Bl;                                                        // sectors of the pending extension
    bool pending = false;
    int is_back = 0, c = 0;
    for (;;) {
      if (pending) {
        // ------------------------------------------------------------ the single gather site: consume the sectors
        PT n0, n1, n2; int two;
        if (ptab) {                                                         // table entry: the child's bi-interval itself
            if (EQ == 1) { const uint4 v = (ptab & 2) ? Bk.pl : Bk.cnt; E::tab(v, v, n0, n1, n2); }
            else E::tab(Bk.cnt, Bk.pl, n0, n1, n2);
            two = 0;
            if (STATS) {
                if (is_back && j >= nprev) two = (int)(vtwo >> ((int)pend - sx) & 1u); // virtual entry: counted when it was made
                else two = occ32_two_ref<PT>(ix, k0, k1, kk2, is_back);
            }
        } else occ32_finish<PT>(ix, k0, k1, kk2, is_back, c, Bk, Bl, n0, n1, n2, two);
        if (STATS) { ++n_ext; n_two += two; }
        // ------------------------------------------------------------ consume
        bool adv = false;                                                   // forward phases: continue with the child
        int restart = -1;                                                   // >= 0: start a new forward chain of this kind at x
        if (st == S3_FWD) {                                                 // bwt.c:308-315
            adv = true;
            if (n2 != kk2) {
                if (nfw - LCAP >= P.spill_cap) { atomicMax(P.overflow + 1, 1); st = S3_FIN; i = -1; adv = false; }
                else {
                    S3_PUSH_FWD(i);
                    if (n2 < min_intv) { S3_START_BWD(); adv = false; }
                }
            }
        } else if (!SPLIT && st == S3_P3) {                                 // bwt.c:369-375
            if (n2 < max_mem_intv && i - x >= min_seed_len) {
                if (n2 > 0) s3_emit(out, n_out, P.cap_out, n0, n1, n2, x, i + 1);   // bwamem.c:174
                x = i + 1; restart = S3_P3;
            } else adv = true;
        } else {                                                            // S3_BWD, bwt.c:331-341
            const bool none = ncurr == 0 && vnext == 0;                     // no longer match survived at this position
            if (n2 < min_intv) {
                if (none && i + 1 < last_start) {
                    last_start = i + 1;
                    if ((int)pend - (i + 1) >= min_seed_len) s3_emit(out, n_out, P.cap_out, k0, k1, kk2, i + 1, (int)pend);
                }
            } else if (none || n2 != last_x2) {
                last_x2 = n2;
                if ((int)pend - i < KTV) {                                  // the child is a table string again: stays a bit
                    const int lb = (int)pend - sx;
                    vnext |= 1u << lb;
                    if (STATS) vtwo = (vtwo & ~(1u << lb)) | (u32)occ32_two_ref<PT>(ix, n0, n1, n2, 1) << lb;
                } else {
                    S3_STORE_ENTRY(ncurr, n0, n1, n2, pend);                // ncurr <= j: slot already consumed
                    ++ncurr;
                }
            }
            if (!STATS && (ptab & 4) && n2 >= min_intv) {
                // The longest virtual entry lives, so every shorter one does (sizes grow as strings shrink), and each is
                // kept exactly when its size differs from its predecessor's (bwt.c:339): a set bit of the window's
                // change mask between the two lengths.  Child length of entry L is L + (sx - i).
                const u32 cm = (EQ == 1 ? E::tabmask((ptab & 8) ? Bl.pl : Bl.cnt, Bl.pl) : E::tabmask(Bl.cnt, Bl.pl)) >> (sx - i);
                // Entry L owns the lengths [L, next longer entry): it is kept when cm has a bit there.  Branch-free: in
                // bit-reversed order an entry's bit sits on top of its range, and adding all-ones over the non-entry
                // positions carries any cm bit of a range up into the entry's position, where the carry stops.
                const u32 rv = __brev(vm), rc = __brev(cm & ((1u << ((int)pend - sx)) - 1u)), ns = ~rv;
                vnext |= __brev((((rc & ns) + ns) & rv) | (rc & rv));
                vm = 0;
            }
            ++j;
            if (j >= nprev && vm == 0) {                                    // bwt.c:343-344
                if (ncurr == 0 && vnext == 0) {
                    if (pass == 1) { x = ret; restart = S3_FWD; } else { ++k2; st = S3_TASK; i = -1; }
                } else {                                                    // next backward position: base c joins the window
                    nprev = ncurr; ncurr = 0; j = 0; vm = vnext; vnext = 0; --i;
                    if (KT > 0) wb = ((u32)c << (2 * KT - 2)) | (wb >> 2);
                }
            }
            if (restart < 0 && st == S3_BWD) S3_NEXT_ENTRY();
        }
        if (adv) {
            k0 = n0; k1 = n1; kk2 = n2; ++i;
            wf = (wf << 2) | (u32)(3 - c); ++flen;                          // forward phases extend with base 3 - c
            if (flen <= KT) wst = wf << (2 * (KT - flen));
        }
        if (restart >= 0) {                                                 // inline restart of a forward chain at x
            u32 b0 = 4;
            if (x < len) b0 = S3_QGET(x);
            if (b0 < 4) {
                S3_SET_INTV((int)b0);
                i = x + 1; st = restart;
                if (restart == S3_FWD) { sx = x; min_intv = 1; nfw = 0; vm = 0; }
            } else { st = S3_TASK; i = -1; }                                // N or end of read: slow path decides
        }
      }
        // ------------------------------------------------------------ request: base at position i, or slow path
        u32 b = 4;
        if ((unsigned)i < (unsigned)len) b = S3_QGET(i);
        while (b > 3) {
            if (st == S3_FWD) {                                             // read end or N: bwt.c:316-318, 321
                if (nfw - LCAP >= P.spill_cap) { atomicMax(P.overflow + 1, 1); st = S3_FIN; i = -1; continue; }
                S3_PUSH_FWD(i);
                S3_START_BWD();
            } else if (st == S3_BWD) {                                      // i < 0 or N: the longest entry is emitted, sweep ends
                if ((int)pend - (i + 1) >= min_seed_len) s3_emit(out, n_out, P.cap_out, k0, k1, kk2, i + 1, (int)pend);
                if (pass == 1) x = ret; else ++k2;
                st = S3_TASK; i = -1;
            } else if (!SPLIT && st == S3_P3) {                             // bwt.c:366-378
                if (i >= len) { st = S3_FIN; i = -1; }
                else { x = i + 1; st = S3_TASK; i = -1; }
            } else if (st == S3_TASK) {
                if (pass == 1) {                                            // bwamem.c:144-154
                    while (x < len && S3_QGET(x) > 3) ++x;
                    if (x >= len) { pass = 2; old_n = n_out < P.cap_out ? n_out : P.cap_out; k2 = 0; continue; }
                    sx = x; min_intv = 1;
                } else if (pass == 2) {                                     // bwamem.c:156-165
                    bool found = false;
                    while (k2 < old_n) {
                        const uint4 v1 = ((const uint4 *)&out[k2])[1];
                        const u64 sz = (u64)v1.y << 32 | v1.x;
                        const int s0 = (int)v1.w, e0 = (int)v1.z;
                        if (e0 - s0 < P.split_len || sz > (u64)P.split_width) { ++k2; continue; }
                        sx = (s0 + e0) >> 1; min_intv = (PT)sz + 1; found = true;
                        break;
                    }
                    if (!found) {
                        if (SPLIT) { st = S3_FIN; continue; }
                        pass = 3; x = 0; continue;
                    }
                } else if (!SPLIT) {                                        // bwamem.c:167-182
                    if (max_mem_intv == 0) { st = S3_FIN; continue; }
                    while (x < len && S3_QGET(x) > 3) ++x;
                    if (x >= len) { st = S3_FIN; continue; }
                    const int b0 = (int)S3_QGET(x);
                    S3_SET_INTV(b0);
                    i = x + 1; st = S3_P3;
                    b = 4; if (i < len) b = S3_QGET(i);
                    continue;
                }
                {                                                           // bwt.c:301-304
                    const int b0 = (int)S3_QGET(sx);
                    S3_SET_INTV(b0);
                    i = sx + 1; nfw = 0; vm = 0; st = S3_FWD;
                }
            } else {                                                        // S3_FIN: finish this read, take the next
                if (r >= 0) {
                    int n = n_out;
                    if (SPLIT && len >= min_seed_len) {                // append the third pass's intervals (stored at the end)
                        const int n3 = P.out3_n[r];
                        if (n + n3 > P.cap_out) { atomicMax(P.overflow, n + n3); n = n > P.cap_out ? P.cap_out : n; }
                        else if (n + n3 <= P.cap_out - n3) {                // destination below the stored block
                            for (int a = 0; a < n3; ++a) {
                                const uint4 *src = (const uint4 *)&out[P.cap_out - 1 - a];
                                const uint4 v0 = src[0], v1 = src[1];
                                ((uint4 *)&out[n + a])[0] = v0; ((uint4 *)&out[n + a])[1] = v1;
                            }
                            n += n3;
                        } else {                                            // they overlap: reverse the block in place, move it down
                            for (int a = 0, z = n3 - 1; a < z; ++a, --z) {
                                uint4 *pa = (uint4 *)&out[P.cap_out - 1 - a], *pz = (uint4 *)&out[P.cap_out - 1 - z];
                                const uint4 a0 = pa[0], a1 = pa[1], z0 = pz[0], z1 = pz[1];
                                pa[0] = z0; pa[1] = z1; pz[0] = a0; pz[1] = a1;
                            }
                            for (int a = 0; a < n3; ++a) {                  // now element a sits at cap_out - n3 + a >= n + a
                                const uint4 *src = (const uint4 *)&out[P.cap_out - n3 + a];
                                const uint4 v0 = src[0], v1 = src[1];
                                ((uint4 *)&out[n + a])[0] = v0; ((uint4 *)&out[n + a])[1] = v1;
                            }
                            n += n3;
                        }
                    }
                    if (n > P.cap_out) { atomicMax(P.overflow, n); n = P.cap_out; }
                    for (int a = 1; a < n; ++a) {                           // sort by info (bwamem.c:184)
                        const uint4 v0 = ((uint4 *)&out[a])[0], v1 = ((uint4 *)&out[a])[1];
                        const u64 key = (u64)v1.w << 32 | v1.z;
                        int bb = a;
                        while (bb > 0) {
                            const uint4 u1 = ((uint4 *)&out[bb - 1])[1];
                            if (((u64)u1.w << 32 | u1.z) <= key) break;
                            ((uint4 *)&out[bb])[0] = ((uint4 *)&out[bb - 1])[0]; ((uint4 *)&out[bb])[1] = u1;
                            --bb;
                        }
                        if (bb != a) { ((uint4 *)&out[bb])[0] = v0; ((uint4 *)&out[bb])[1] = v1; }
                    }
                    P.out_n[r] = n;
                }
                const unsigned nr = atomicAdd(P.work_counter, 1u);
                if (nr >= (unsigned)P.n_reads) { r = -1; break; }
                r = (int)nr;
                len = (int)(P.off[r + 1] - P.off[r]);
                out = P.out + (size_t)r * P.cap_out;
                n_out = 0; i = -1;
                if (len < min_seed_len) continue;                           // bwamem.c:267 (FIN again -> out_n = 0)
                if (QGLOB) Qg = P.pk + (size_t)r * P.pk_stride;
                else {                                                      // stage the packed read
                    const uint4 *src = (const uint4 *)(P.pk + (size_t)r * P.pk_stride);
                    const int nq = (len + 31) >> 5;
                    for (int w = 0; w < nq; ++w) {
                        const uint4 v = __ldcs(src + w);
                        Q[(4 * w) * S3_THREADS] = v.x; Q[(4 * w + 1) * S3_THREADS] = v.y;
                        Q[(4 * w + 2) * S3_THREADS] = v.z; Q[(4 * w + 3) * S3_THREADS] = v.w;
                    }
                }
                pass = 1; x = 0; st = S3_TASK;
                continue;
            }
            b = 4;
            if ((unsigned)i < (unsigned)len) b = S3_QGET(i);
        }
        if (r < 0) break;                                                   // no reads left for this lane
        is_back = st == S3_BWD;
        c = is_back ? (int)b : 3 - (int)b;
        {
            const int cur = is_back ? (int)pend - (i + 1) : flen;           // length of the string whose interval is extended
            // One load group for the whole warp whatever the lanes read: a 16-byte table record is fetched as the 32-byte
            // sector that holds it (ptab = 1 / 2: low / high half), two-quad records as one "sector".
            const uint4 *pk, *pl;
            ptab = cur < KT;
            if (ptab) {                                                     // child = string of cur + 1 <= KT bases: table
                const u32 wn = is_back ? (b << (2 * cur)) | (cur ? wb >> (2 * (KT - cur)) : 0u) : (wf << 2) | b;
                const size_t e = (size_t)(ix.kmer_off[cur + 1] + wn) * EQ;
                if (EQ == 1) { ptab = 1 | (int)(e & 1) << 1; pk = ktab + (e & ~(size_t)1); } else pk = ktab + e;
                pl = pk;
                // a virtual entry with other virtual entries behind it, and the KT bases from i on all inside the forward
                // sweep's range: fetch the window's depth-KT record too, its mask settles the others (consume step)
                if (!STATS && is_back && j >= nprev && vm != 0 && ret - i >= KT) {
                    const size_t ew = (size_t)(ix.kmer_off[KT] + ((b << (2 * KT - 2)) | (wb >> 2))) * EQ;
                    if (EQ == 1) { ptab |= 4 | (int)(ew & 1) << 3; pl = ktab + (ew & ~(size_t)1); } else { ptab |= 4; pl = ktab + ew; }
                }
            } else occ32_addr<PT>(ix, k0, k1, kk2, is_back, pk, pl);
            occ32_load2p(pk, pl, Bk, Bl);                                   // requests in flight across the back-edge
        }
        pending = true;
    }
    if (STATS && n_ext) { atomicAdd(&P.stats[CNT_EXTEND], (unsigned long long)n_ext); atomicAdd(&P.stats[CNT_TWO_BLOCK], (unsigned long long)n_two); }
}

// Third seeding pass on its own (bwamem.c:167-182 around bwt_seed_strategy1, bwt.c:358-379): forward-only sweeps that
// restart right after every seed, so every lane runs the same short loop -- no lists, no backward phase.  One read per
// lane (persistent), requests issued at the bottom of the loop and consumed at the top like in the main kernel.
template <class E, bool QGLOB, bool STATS>
__global__ void __launch_bounds__(S3_THREADS) smem_p3_kernel(SmemParams P)
{
    B200_DYN_SMEM(p3_raw);
    typedef typename E::PT PT;
    constexpr int EQ = E::kQuads;
    u32 *Q = (u32 *)p3_raw + threadIdx.x;                                   // word w: Q[w * S3_THREADS]
    const u32 *Qg = P.pk;
    const DevIndex &ix = P.ix;
    const int min_seed_len = P.min_seed_len;
    const PT max_mem_intv = (PT)P.max_mem_intv;
    const int KT = ix.kmer_k;
    const uint4 *const ktab = ix.kmer_tab;
    u64 n_ext = 0, n_two = 0;
    Intv *out = nullptr;
    int len = 0, r = -1, n_out = 0, x = 0, i = 0, flen = 0, c = 0;
    PT k0 = 0, k1 = 0, kk2 = 0;
    u32 wf = 0;
    Occ32Blk Bk, Bl;
    bool pending = false;
    int ptab = 0;
    for (;;) {
        if (pending) {
            PT n0, n1, n2; int two;
            if (ptab) {
                if (EQ == 1) { const uint4 v = ptab == 2 ? Bk.pl : Bk.cnt; E::tab(v, v, n0, n1, n2); }
                else E::tab(Bk.cnt, Bk.pl, n0, n1, n2);
                two = STATS ? occ32_two_ref<PT>(ix, k0, k1, kk2, 0) : 0;
            } else occ32_finish<PT>(ix, k0, k1, kk2, 0, c, Bk, Bl, n0, n1, n2, two);
            if (STATS) { ++n_ext; n_two += two; }
            if (n2 < max_mem_intv && i - x >= min_seed_len) {               // bwt.c:369-373
                if (n2 > 0) {                                               // bwamem.c:174
                    if (n_out < P.cap_out) {
                        uint4 *q = (uint4 *)&out[P.cap_out - 1 - n_out];
                        __stcs(q, make_uint4((u32)n0, (u32)((u64)n0 >> 32), (u32)n1, (u32)((u64)n1 >> 32)));
                        __stcs(q + 1, make_uint4((u32)n2, (u32)((u64)n2 >> 32), (u32)(i + 1), (u32)x));
                    }
                    ++n_out;
                }
                x = i + 1; flen = 0;                                        // next sweep starts after the seed
            } else { k0 = n0; k1 = n1; kk2 = n2; wf = (wf << 2) | (u32)(3 - c); ++flen; ++i; }
            pending = false;
        }
        // ---- next request: continue the sweep, start a new one, or take the next read
        u32 b = 4;
        for (;;) {
            if (flen > 0) {                                                 // inside a sweep: base at i
                if (i < len) { b = S3_QGET(i); if (b < 4) break; x = i + 1; flen = 0; continue; }   // N: bwt.c:376
                x = len; flen = 0; continue;                                // read end: bwt.c:378
            }
            while (x < len && S3_QGET(x) > 3) ++x;                          // bwamem.c:169-181
            if (r >= 0 && x < len) {
                const int b0 = (int)S3_QGET(x);
                k0 = (PT)ix.L2[b0] + 1; kk2 = (PT)(ix.L2[b0 + 1] - ix.L2[b0]); k1 = (PT)ix.L2[3 - b0] + 1;   // bwt_set_intv
                wf = (u32)b0; flen = 1; i = x + 1;
                continue;
            }
            if (r >= 0) {
                if (n_out > P.cap_out) { atomicMax(P.overflow, n_out); n_out = P.cap_out; }
                P.out3_n[r] = n_out;
            }
            const unsigned nr = atomicAdd(P.work_counter, 1u);
            if (nr >= (unsigned)P.n_reads) { r = -1; break; }
            r = (int)nr;
            len = (int)(P.off[r + 1] - P.off[r]);
            out = P.out + (size_t)r * P.cap_out;
            n_out = 0; x = 0; flen = 0;
            if (len < min_seed_len) { x = len; continue; }                  // bwamem.c:267: no seeding at all
            if (QGLOB) { Qg = P.pk + (size_t)r * P.pk_stride; continue; }
            const uint4 *src = (const uint4 *)(P.pk + (size_t)r * P.pk_stride);
            const int nq = (len + 31) >> 5;
            for (int w = 0; w < nq; ++w) {
                const uint4 v = __ldcs(src + w);
                Q[(4 * w) * S3_THREADS] = v.x; Q[(4 * w + 1) * S3_THREADS] = v.y;
                Q[(4 * w + 2) * S3_THREADS] = v.z; Q[(4 * w + 3) * S3_THREADS] = v.w;
            }
        }
        if (r < 0) break;
        c = 3 - (int)b;
        const uint4 *pk, *pl;                                               // one load group per warp, as in the main kernel
        ptab = flen < KT;
        if (ptab) {
            const size_t e = (size_t)(ix.kmer_off[flen + 1] + ((wf << 2) | b)) * EQ;
            if (EQ == 1) { ptab = 1 + (int)(e & 1); pk = ktab + (e & ~(size_t)1); } else pk = ktab + e;
            pl = pk;
        } else occ32_addr<PT>(ix, k0, k1, kk2, 0, pk, pl);
        occ32_load2p(pk, pl, Bk, Bl);
        pending = true;
    }
    if (STATS && n_ext) { atomicAdd(&P.stats[CNT_EXTEND], (unsigned long long)n_ext); atomicAdd(&P.stats[CNT_TWO_BLOCK], (unsigned long long)n_two); }
}

template <class E, bool QGLOB, bool STATS>
static int launch_p3_ts(const SmemParams &p, int n_sms, cudaStream_t st)
{
    const size_t smem = QGLOB ? 0 : (size_t)p.pk_stride * 4 * S3_THREADS;
    int per_sm = (int)std::max<size_t>(1, std::min<size_t>(12, (200 * 1024) / (smem + 1024)));
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute((smem_p3_kernel<E, QGLOB, STATS>), cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    int occ = 0;                                                            // resident blocks per SM (registers + shared memory)
    B200_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (smem_p3_kernel<E, QGLOB, STATS>), S3_THREADS, smem));
    if (occ > 0) per_sm = std::min(per_sm, occ);
#endif
    const int grid = std::max(1, std::min((p.n_reads + S3_THREADS - 1) / S3_THREADS, n_sms * per_sm));
    B200_LAUNCH((smem_p3_kernel<E, QGLOB, STATS>), grid, S3_THREADS, smem, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}

template <class E, bool QGLOB>
static int launch_p3_t(const SmemParams &p, int n_sms, cudaStream_t st)
{
    return p.count_stats ? launch_p3_ts<E, QGLOB, true>(p, n_sms, st) : launch_p3_ts<E, QGLOB, false>(p, n_sms, st);
}

// wide: 0 = 32-bit positions, 1 = 64-bit positions with two-quad records, 2 = 64-bit positions < 2^33 with packed records
int launch_smem_p3(const SmemParams &p, int n_sms, int wide, int qglob, cudaStream_t st)
{
    if (qglob) return wide == 2 ? launch_p3_t<Ent33, true>(p, n_sms, st) : wide ? launch_p3_t<Ent<u64>, true>(p, n_sms, st) : launch_p3_t<Ent<u32>, true>(p, n_sms, st);
    return wide == 2 ? launch_p3_t<Ent33, false>(p, n_sms, st) : wide ? launch_p3_t<Ent<u64>, false>(p, n_sms, st) : launch_p3_t<Ent<u32>, false>(p, n_sms, st);
}

int launch_pack4(u8 *seq, const i64 *off, int n_reads, u32 *pk, int stride, cudaStream_t st)
{
    if (n_reads <= 0) return 0;
    B200_LAUNCH(pack4_kernel, (int)(((i64)n_reads * 32 + 255) / 256), 256, 0, st, seq, off, n_reads, pk, stride);
    B200_CHECK(cudaGetLastError());
    return 0;
}

size_t smem_fast3_bytes(int wide, int lcap, int pk_stride)             // pk_stride = 0: read not staged (QGLOB)
{
    return (size_t)lcap * (wide == 1 ? 32 : 16) * S3_THREADS + (size_t)pk_stride * 4 * S3_THREADS;
}

template <class E, int LCAP, bool SPLIT, bool QGLOB, bool STATS>
static int launch_fast3_tss(const SmemParams &p, int grid, size_t smem, cudaStream_t st)
{
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute((smem_fast3_kernel<E, LCAP, SPLIT, QGLOB, STATS>), cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
#endif
    B200_LAUNCH((smem_fast3_kernel<E, LCAP, SPLIT, QGLOB, STATS>), grid, S3_THREADS, smem, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}

template <class E, int LCAP, bool SPLIT, bool QGLOB>
static int launch_fast3_ts(const SmemParams &p, int grid, size_t smem, cudaStream_t st)
{
    return p.count_stats ? launch_fast3_tss<E, LCAP, SPLIT, QGLOB, true>(p, grid, smem, st) : launch_fast3_tss<E, LCAP, SPLIT, QGLOB, false>(p, grid, smem, st);
}

template <class E>
static int launch_fast3_e(const SmemParams &p, int grid, int wide, int lcap, int qglob, cudaStream_t st)
{
    if (qglob) {                                                            // long reads: lcap 16, read in global memory
        const size_t smem = smem_fast3_bytes(wide, 16, 0);
        return p.split_p3 ? launch_fast3_ts<E, 16, true, true>(p, grid, smem, st) : launch_fast3_ts<E, 16, false, true>(p, grid, smem, st);
    }
    const size_t smem = smem_fast3_bytes(wide, lcap, p.pk_stride);
    if (lcap == 8) return p.split_p3 ? launch_fast3_ts<E, 8, true, false>(p, grid, smem, st) : launch_fast3_ts<E, 8, false, false>(p, grid, smem, st);
    return p.split_p3 ? launch_fast3_ts<E, 16, true, false>(p, grid, smem, st) : launch_fast3_ts<E, 16, false, false>(p, grid, smem, st);
}

int launch_smem_fast3(const SmemParams &p, int grid, int wide, int lcap, int qglob, cudaStream_t st)
{
    if (wide == 2) return launch_fast3_e<Ent33>(p, grid, wide, lcap, qglob, st);
    if (wide) return launch_fast3_e<Ent<u64> >(p, grid, wide, lcap, qglob, st);
    return launch_fast3_e<Ent<u32> >(p, grid, wide, lcap, qglob, st);
}
