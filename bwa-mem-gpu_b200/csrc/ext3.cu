// Stage 3 DP kernel for short reads, 8-bit scores: ksw_extend2 (ksw.c:380-479) with TWO columns per instruction.
//
// Same round structure and job format as ext2.cu (ext_plan_kernel / ext_finish_kernel are shared); this kernel
// replaces ext_dp_kernel when every score of a read fits one byte (l_query * a <= 255, e.g. 150-250 bp reads at a = 1).
//
// In the reference's inner loop (ksw.c:424-448) a cell depends on its row only through F (a running max-plus scan of
// values that come from the PREVIOUS row) and through the bookkeeping of the row maximum; everything else is
// column-parallel.  So two adjacent columns are evaluated together in one 32-bit register of two 16-bit halves with
// the DPX 16x2 instructions (VIADDMNMX.S16x2[.RELU], VIMNMX3.S16x2, VIMNMX.U16x2):
//   M'  = min(H + s, 64 H)                 == "M ? M + s : 0" wherever it is observable (a negative M acts like 0)
//   E'  = max(E - e_del, M' - oe_del, 0)    one add + one fused add-max-relu
//   F   : (F_j, F_j+1) from the carried F_j, then the carry F_j+2, two fused add-max
//   H   = max3(M', E, F)
//   row maximum with "last column wins ties" (ksw.c:437-438) = unsigned max of the keys (H << 8 | column)
// The H/E row (including the stale out-of-band entries the algorithm observes later) lives in shared memory as one
// byte each per column, eight columns = one 16-byte word per lane, laid out [column group][lane]: a hot slot is ONE
// 128-bit load, one 64-bit load of the eight pre-computed score selectors (prmt picks the two substitution scores of a
// column pair out of the target base's matrix row), and one 128-bit store.  Columns outside [beg, end) of the row are
// masked on input (so they are neutral for F) and on output (so stale entries survive); the store mask includes column
// `end`, which makes ksw.c:449 (eh[end].h = h1, eh[end].e = 0) part of the same store.  The zero-trimming of the next
// row's [beg, end) (ksw.c:466-469) re-reads the row ends from shared memory in the control slot instead of tracking
// non-zero columns per cell.
#include "b200_internal.h"
#include "extend_common.cuh"
#include "ext_plan.cuh"

#define D8_THREADS 64

__device__ __forceinline__ u32 d8_prmt(u32 a, u32 b, u32 s)
{
#ifdef B200_EMU
    return emu_prmt(a, b, s);
#else
    u32 r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(s));
    return r;
#endif
}

__device__ __forceinline__ ExtJob *d8_job_slot(const Ext2Params &P, int parity, int bin, unsigned slot)
{
    if (bin == 1) return P.jobs_mid[parity] + slot;
    return bin == 0 ? P.jobs[parity] + slot : P.jobs[parity] + (P.jobs_cap - 1 - (int)slot);
}

struct D8Lane {
    ExtJob *job;
    const u8 *qbase; int qstep;                         // column j of the current side is qbase[j * qstep]
    i64 fpos; int fstep, comp; u32 pacw, pacw_tag;      // forward-strand coordinate of the current target row
    int side, t, a_score, sc0, prev;
    int qlen, tlen, w, h0;                              // ksw_extend2 arguments
    int i, beg, end, g, gend, h1i;                      // row, its column range, current / last column group, H(i, beg-1)
    u32 fc, hc, key;                                    // F carry (high half), H of the previous pair, packed row-maximum keys
    u32 srow_lo, srow_hi;                               // matrix row of the current target base (5 signed bytes)
    int max, max_i, max_j, max_ie, gscore, max_off;
};

__device__ __forceinline__ int d8_target_base(const DevIndex &ix, D8Lane &L)
{
    const u32 tag = (u32)(L.fpos >> 4);
    if (tag != L.pacw_tag) { L.pacw = ((const u32 *)ix.pac)[L.fpos >> 4]; L.pacw_tag = tag; }
    const int f = (int)L.fpos;
    const int sh = (((f >> 2) & 3) << 3) + ((~f & 3) << 1);
    return (int)((L.pacw >> sh) & 3) ^ L.comp;
}

struct D8Const { u32 ne_del2, noe_del2, noe_ins2, ne_ins2, ne_ins_hi; };

// two columns (j, j+1): w = their H/E bytes [h_j, e_j, h_j+1, e_j+1], im / sm = input / store masks, qsel = prmt selector
// of the two substitution scores, col2 = (j, j+1)
template <bool MASKED>
__device__ __forceinline__ u32 d8_pair(u32 w, u32 im, u32 sm, u32 qsel, u32 col2, const D8Lane &L, const D8Const &K,
                                       u32 &fc, u32 &hc, u32 &key)
{
    const u32 wm = MASKED ? w & im : w;
    const u32 H2 = d8_prmt(wm, 0u, 0x4240u), E2 = d8_prmt(wm, 0u, 0x4341u);
    const u32 sc2 = d8_prmt(L.srow_lo, L.srow_hi, qsel);
    const u32 Mp = __viaddmin_s16x2(H2, sc2, H2 << 6);                                  // ksw.c:433
    const u32 en2 = __viaddmax_s16x2_relu(Mp, K.noe_del2, __vadd2(E2, K.ne_del2));      // ksw.c:439-442
    const u32 T2 = __viaddmax_s16x2_relu(Mp, K.noe_ins2, 0u);                           // ksw.c:444
    const u32 A = d8_prmt(fc, 0u, 0x3232u);
    const u32 F2 = __viaddmax_s16x2(A, K.ne_ins_hi, T2 << 16);                          // (F_j, F_j+1), ksw.c:445-447
    fc = __viaddmax_s16x2(F2, K.ne_ins2, T2);                                           // high half: F_j+2
    const u32 h2 = __vimax3_s16x2(Mp, E2, F2);                                          // ksw.c:434-435
    key = __vmaxu2(key, MASKED ? ((h2 << 8) + col2) & im : (h2 << 8) + col2);           // ksw.c:437-438
    const u32 Hs2 = d8_prmt(hc, h2, 0x5432u);                                           // (H(i,j-1), H(i,j)), ksw.c:432
    hc = h2;
    const u32 nw = en2 * 256u + Hs2;
    return MASKED ? (nw & sm) | (w & ~sm) : nw;
}

// bit c set when column c of the 16-byte group word holds a non-zero h or e
__device__ __forceinline__ unsigned d8_nz8(const uint4 w)
{
    unsigned m = 0;
    const u32 v[4] = { w.x, w.y, w.z, w.w };
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        if (v[p] & 0x0000ffffu) m |= 1u << (2 * p);
        if (v[p] & 0xffff0000u) m |= 2u << (2 * p);
    }
    return m;
}

// ROWSYNC: a lane runs all column groups of its row before the warp meets in the control slot, so the per-row
// bookkeeping executes once per row for (nearly) all 32 lanes instead of once per 8-column slot for the few lanes that
// happen to finish a row in that slot; lanes with shorter rows idle until the longest row of the warp is done (jobs are
// binned by flank length, which bounds that).  ROWSYNC = false is the one-slot-per-iteration state machine.
// `phased` (with ROWSYNC): the warp moves through its jobs in phases -- left sides, left band retries, right sides, right
// retries, next 32 jobs -- a lane whose next DP belongs to a later phase parks until no lane is mid-DP.  Together with the
// job order (same flank shape next to each other) the rows a warp executes in one step then have the same number of
// column groups, instead of a short left flank in one lane waiting for a long right flank in another.
// job t of a round, counted over the three bins in the order the planning kernel reads them back
__device__ __forceinline__ bool d8_locate(const Ext2Params &P, int round, unsigned t, int &bin, unsigned &slot)
{
    const unsigned *nj = P.njobs + 8 * round;
    bin = 0; slot = t;
    if (slot >= nj[0]) { slot -= nj[0]; bin = 2; if (slot >= nj[2]) { slot -= nj[2]; bin = 1; if (slot >= nj[1]) return false; } }
    return true;
}

// TAIL: the last rounds of a chunk hold few jobs each (the k-th extended seed of the few reads that have one) and cost a
// planning launch, a host round trip and up to three nearly empty DP launches per round.  In tail mode ONE launch finishes
// the chunk: a lane takes a job of the current round (from any bin; rows are sized for the longest flank) and, when it is
// done, plans the next seed of that read itself (ext_plan_read) and runs it, until the read has no seed left.
template <bool ROWSYNC, bool TAIL>
__global__ void __launch_bounds__(D8_THREADS) ext_dp8_kernel(Ext2Params P, int round, int bin, int n_groups, int phased)
{
    B200_DYN_SMEM(d8_raw);
    __shared__ u32 s_rows[5 * 2];                          // matrix row of each target base, 8 bytes (5 used)
    __shared__ uint4 s_pref[9];                            // s_pref[k]: the k lowest columns of a group selected
    uint4 *HE = (uint4 *)d8_raw + threadIdx.x;             // group g of this lane: HE[g * D8_THREADS]
    uint2 *QS = (uint2 *)(d8_raw + (size_t)n_groups * D8_THREADS * 16) + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    const unsigned wbase = threadIdx.x & ~31u;
    if (threadIdx.x < 10) {
        u32 v = 0;
        for (int k = 0; k < 4; ++k) {
            const int col = (threadIdx.x & 1) * 4 + k;
            if (col < 5) v |= (u32)(u8)P.opt.mat[(threadIdx.x >> 1) * 5 + col] << (8 * k);
        }
        s_rows[threadIdx.x] = v;
    }
    if (threadIdx.x < 9) {
        u32 v[4];
        for (int p = 0; p < 4; ++p) v[p] = (2 * p < (int)threadIdx.x ? 0x0000ffffu : 0u) | (2 * p + 1 < (int)threadIdx.x ? 0xffff0000u : 0u);
        s_pref[threadIdx.x] = make_uint4(v[0], v[1], v[2], v[3]);
    }
    __syncthreads();
    const DevIndex &ix = P.ix;
    const mem_opt_t &opt = P.opt;
    const int o_del = opt.o_del, e_del = opt.e_del, o_ins = opt.o_ins, e_ins = opt.e_ins;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins, zdrop = opt.zdrop;
    D8Const K;
    K.ne_del2 = ((u32)(-e_del) & 0xffffu) * 0x10001u; K.noe_del2 = ((u32)(-oe_del) & 0xffffu) * 0x10001u;
    K.noe_ins2 = ((u32)(-oe_ins) & 0xffffu) * 0x10001u; K.ne_ins2 = ((u32)(-e_ins) & 0xffffu) * 0x10001u;
    K.ne_ins_hi = ((u32)(-e_ins) & 0xffffu) << 16;
    const unsigned njobs = TAIL ? P.njobs[8 * round] + P.njobs[8 * round + 1] + P.njobs[8 * round + 2] : P.njobs[8 * round + bin];
    unsigned *work_counter = P.njobs + 8 * round + 4 + (TAIL ? 3 : bin);
    ExtJob *my_job = TAIL ? P.tail_jobs + ((size_t)blockIdx.x * D8_THREADS + threadIdx.x) : nullptr;     // the jobs this lane plans itself
    bool have_done = false;
    const i64 l_pac = ix.l_pac;
    u64 n_ksw = 0, n_cells = 0;
    unsigned ncell32 = 0;

    D8Lane L;
    memset(&L, 0, sizeof(L));
    bool active = false, exhausted = false, need_init = false, parked = false;

    for (;;) {
        // ---------------------------------------------------------------- hot slot: one group of 8 columns
        for (bool first = true; active && !parked && L.g <= L.gend && (ROWSYNC || first); first = false) {
            const int c0 = L.g << 3;
            uint4 *hp = HE + L.g * D8_THREADS;
            const uint4 w = *hp;
            const uint2 q = QS[L.g * D8_THREADS];
            const u32 cb = (u32)c0 * 0x10001u + 0x00010000u;
            uint4 o;
            // a group wholly inside [beg, end) needs no masks; the warp takes the unmasked path only when that holds for every
            // lane still in its row (both paths give the same values for such a group)
            bool edge = L.beg > c0 || L.end < c0 + 8;
#ifndef B200_EMU
            if (ROWSYNC) edge = __any_sync(__activemask(), edge);
#endif
            if (edge) {
                int lo = L.beg - c0; lo = lo < 0 ? 0 : lo;
                int hi = L.end - c0, hs = hi + 1;
                hi = hi > 8 ? 8 : hi; hs = hs > 8 ? 8 : hs;
                const uint4 plo = s_pref[lo], phi = s_pref[hi], phs = s_pref[hs];
                o.x = d8_pair<true>(w.x, phi.x & ~plo.x, phs.x & ~plo.x, q.x, cb, L, K, L.fc, L.hc, L.key);
                o.y = d8_pair<true>(w.y, phi.y & ~plo.y, phs.y & ~plo.y, q.x >> 16, cb + 0x00020002u, L, K, L.fc, L.hc, L.key);
                o.z = d8_pair<true>(w.z, phi.z & ~plo.z, phs.z & ~plo.z, q.y, cb + 0x00040004u, L, K, L.fc, L.hc, L.key);
                o.w = d8_pair<true>(w.w, phi.w & ~plo.w, phs.w & ~plo.w, q.y >> 16, cb + 0x00060006u, L, K, L.fc, L.hc, L.key);
            } else {
                o.x = d8_pair<false>(w.x, 0u, 0u, q.x, cb, L, K, L.fc, L.hc, L.key);
                o.y = d8_pair<false>(w.y, 0u, 0u, q.x >> 16, cb + 0x00020002u, L, K, L.fc, L.hc, L.key);
                o.z = d8_pair<false>(w.z, 0u, 0u, q.y, cb + 0x00040004u, L, K, L.fc, L.hc, L.key);
                o.w = d8_pair<false>(w.w, 0u, 0u, q.y >> 16, cb + 0x00060006u, L, K, L.fc, L.hc, L.key);
            }
            *hp = o;
            ++L.g;
        }
        if (ROWSYNC) __syncwarp();
        // ---------------------------------------------------------------- control slot
        bool ksw_fin = false;
        if (active && !parked && L.g > L.gend) {                                          // row i finished
            int m = 0, mj = -1, h1 = L.h1i;
            if (L.beg < L.end) {
                const u32 klo = L.key & 0xffffu, khi = L.key >> 16;
                const u32 km = klo > khi ? klo : khi;
                m = (int)(km >> 8); mj = (int)(km & 0xffu);
                h1 = ((const u8 *)(HE + (L.end >> 3) * D8_THREADS))[(L.end & 7) * 2];     // eh[end].h, ksw.c:449
            }
            const int jexit = L.beg < L.end ? L.end : L.beg;
            if (jexit == L.qlen) {                                                        // ksw.c:450-453
                L.max_ie = L.gscore > h1 ? L.max_ie : L.i;
                L.gscore = L.gscore > h1 ? L.gscore : h1;
            }
            if (m == 0) ksw_fin = true;                                                   // ksw.c:454
            else {
                if (m > L.max) {
                    L.max = m; L.max_i = L.i; L.max_j = mj;
                    int off = mj - L.i; off = off < 0 ? -off : off;
                    L.max_off = L.max_off > off ? L.max_off : off;
                } else if (zdrop > 0) {                                                   // ksw.c:458-464
                    if (L.i - L.max_i > mj - L.max_j) {
                        if (L.max - m - ((L.i - L.max_i) - (mj - L.max_j)) * e_del > zdrop) ksw_fin = true;
                    } else {
                        if (L.max - m - ((mj - L.max_j) - (L.i - L.max_i)) * e_ins > zdrop) ksw_fin = true;
                    }
                }
                if (!ksw_fin) {                                                           // ksw.c:466-469 on the stored row
                    int nb = L.end;
                    for (int g = L.beg >> 3; (g << 3) < L.end; ++g) {
                        const int c0 = g << 3;
                        unsigned nz = d8_nz8(HE[g * D8_THREADS]);
                        const int a = L.beg - c0, b = L.end - c0;                         // columns [a, b) of the group
                        if (a > 0) nz &= ~((1u << a) - 1u);
                        if (b < 8) nz &= (1u << b) - 1u;
                        if (nz) { nb = c0 + __ffs((int)nz) - 1; break; }
                    }
                    int jl = nb - 1;
                    for (int g = L.end >> 3; g >= 0 && (g << 3) + 7 >= nb; --g) {
                        const int c0 = g << 3;
                        unsigned nz = d8_nz8(HE[g * D8_THREADS]);
                        const int a = nb - c0, b = L.end - c0 + 1;                        // columns [a, b) = [nb, end]
                        if (a > 0) nz &= ~((1u << a) - 1u);
                        if (b < 8) nz &= (1u << b) - 1u;
                        if (nz) { jl = c0 + 31 - __clz((int)nz); break; }
                    }
                    L.beg = nb;
                    L.end = jl + 2 < L.qlen ? jl + 2 : L.qlen;
                    if (++L.i >= L.tlen) ksw_fin = true;
                    else {                                                                // next row, ksw.c:411-423
                        L.fpos += L.fstep;
                        { const int tb = d8_target_base(ix, L); L.srow_lo = s_rows[2 * tb]; L.srow_hi = s_rows[2 * tb + 1]; }
                        if (L.beg < L.i - L.w) L.beg = L.i - L.w;
                        if (L.end > L.i + L.w + 1) L.end = L.i + L.w + 1;
                        if (L.end > L.qlen) L.end = L.qlen;
                        L.h1i = 0;
                        if (L.beg == 0) { L.h1i = L.h0 - (o_del + e_del * (L.i + 1)); if (L.h1i < 0) L.h1i = 0; }
                        L.fc = 0; L.key = 0; L.hc = (u32)L.h1i << 16;
                        L.g = L.beg >> 3; L.gend = L.end >> 3;
                        if (L.end > L.beg) ncell32 += (unsigned)(L.end - L.beg);
                        else L.g = L.gend + 1;                                            // empty row: m == 0 ends the DP next slot
                    }
                }
            }
        }
        // job / side / retry transitions; loops only when a DP has no rows (tlen <= 0)
        const bool may_fetch = !phased || !__any_sync(B200_FULL_MASK, active);              // phased: the warp takes 32 jobs at a time
        while ((active && ksw_fin) || (!active && (!exhausted || (TAIL && have_done)) && may_fetch)) {
            int next_side = -1, next_t = 0;
            if (!active) {                                                                // fetch a job
                ExtJob *nj = nullptr;
                if (TAIL && have_done) {                                                  // the read of the job just finished: its next seed, if any
                    have_done = false;
                    ExtJob tmp;
                    if (ext_plan_read(P, L.job->read, L.job, false, &tmp)) { *my_job = tmp; nj = my_job; }
                }
                if (!nj) {
                    const unsigned slot = atomicAdd(work_counter, 1u);
                    if (slot >= njobs) { exhausted = true; break; }
                    if (TAIL) {
                        int jb; unsigned js;
                        d8_locate(P, round, slot, jb, js);
                        nj = d8_job_slot(P, round & 1, jb, js);
                    } else nj = d8_job_slot(P, round & 1, bin, P.perm ? P.perm[(size_t)bin * P.jobs_cap + slot] : slot);
                }
                L.job = nj;
                active = true;
                L.a_score = -1;
                if (L.job->qbeg) { next_side = 0; next_t = 0; L.prev = -1; }
                else {                                                                    // bwamem.c:751
                    L.a_score = L.job->slen * opt.a;
                    L.job->truesc = L.a_score; L.job->qb = 0; L.job->rb = L.job->rbeg; L.job->w = opt.w;
                    next_side = 1;
                }
            } else {                                                                      // a ksw_extend2 finished
                ksw_fin = false;
                const int score = L.max, qle = L.max_j + 1, tle = L.max_i + 1, gtle = L.max_ie + 1;
                const int aw = opt.w << L.t;
                L.a_score = score;
                const bool again = !(score == L.prev || L.max_off < (aw >> 1) + (aw >> 2)) && L.t == 0;   // bwamem.c:733-741
                ExtJob *J = L.job;
                if (L.side == 0) {
                    if (again) { next_side = 0; next_t = 1; L.prev = score; }
                    else {
                        if (L.gscore <= 0 || L.gscore <= score - opt.pen_clip5) {         // bwamem.c:743-749
                            J->qb = J->qbeg - qle; J->rb = J->rbeg - tle; J->truesc = score;
                        } else { J->qb = 0; J->rb = J->rbeg - gtle; J->truesc = L.gscore; }
                        J->w = aw;
                        next_side = 1;
                    }
                } else {
                    if (again) { next_side = 1; next_t = 1; L.prev = score; }
                    else {
                        const int qe = J->qbeg + J->slen;
                        const i64 re = J->rbeg + J->slen;
                        if (L.gscore <= 0 || L.gscore <= score - opt.pen_clip3) {         // bwamem.c:771-777
                            J->qe = qe + qle; J->re = re + tle; J->truesc += score - L.sc0;
                        } else { J->qe = J->lq; J->re = re + gtle; J->truesc += L.gscore - L.sc0; }
                        J->w = J->w > aw ? J->w : aw;
                        J->score = score;
                        active = false;                                                   // job done
                        if (TAIL) have_done = true;
                        n_cells += ncell32; ncell32 = 0;
                        continue;
                    }
                }
            }
            if (next_side == 1 && next_t == 0) {     // entering the right side
                ExtJob *J = L.job;
                if (J->qbeg + J->slen == J->lq) {                                         // bwamem.c:778
                    J->qe = J->lq; J->re = J->rbeg + J->slen; J->score = L.a_score;
                    active = false;
                    if (TAIL) have_done = true;
                    n_cells += ncell32; ncell32 = 0;
                    continue;
                }
                L.sc0 = L.a_score; L.prev = L.a_score;
            }
            {   // set up ksw_extend2 for (next_side, next_t)
                ExtJob *J = L.job;
                L.side = next_side; L.t = next_t;
                i64 p0;
                const u8 *rq = P.seq + P.off[J->read];
                if (next_side == 0) {                                                     // bwamem.c:722-735
                    L.qlen = J->qbeg; L.tlen = (int)(J->rbeg - J->rmax0); L.h0 = J->slen * opt.a;
                    L.w = next_t ? J->wl1 : J->wl0;
                    L.qbase = rq + J->qbeg - 1; L.qstep = -1;
                    p0 = J->rbeg - 1;
                } else {                                                                  // bwamem.c:753-765
                    const int qe = J->qbeg + J->slen;
                    L.qlen = J->lq - qe; L.tlen = (int)(J->rmax1 - (J->rbeg + J->slen)); L.h0 = L.sc0;
                    L.w = next_t ? J->wr1 : J->wr0;
                    L.qbase = rq + qe; L.qstep = 1;
                    p0 = J->rbeg + J->slen;
                }
                const int dir = next_side == 0 ? -1 : 1;
                if (p0 >= l_pac) { L.fpos = (l_pac << 1) - 1 - p0; L.fstep = -dir; L.comp = 3; }
                else { L.fpos = p0; L.fstep = dir; L.comp = 0; }
                L.max = L.h0; L.max_i = L.max_j = L.max_ie = -1; L.gscore = -1; L.max_off = 0;
                L.beg = 0; L.end = L.qlen; L.i = 0;
                ++n_ksw;
                if (L.tlen <= 0) { ksw_fin = true; continue; }                            // no rows: result is the initial state
                L.pacw_tag = 0xffffffffu;
                { const int tb = d8_target_base(ix, L); L.srow_lo = s_rows[2 * tb]; L.srow_hi = s_rows[2 * tb + 1]; }
                if (L.end > L.w + 1) L.end = L.w + 1;                                     // row 0: beg = 0 >= -w
                if (L.end > L.qlen) L.end = L.qlen;
                L.h1i = L.h0 - (o_del + e_del); if (L.h1i < 0) L.h1i = 0;
                L.fc = 0; L.key = 0; L.hc = (u32)L.h1i << 16;
                L.g = 0; L.gend = L.end >> 3;
                if (L.end > 0) ncell32 += (unsigned)L.end; else L.g = L.gend + 1;
                need_init = true;
            }
            break;
        }
        // ---------------------------------------------------------------- cooperative first-row initialisation (ksw.c:395-397)
        unsigned need = __ballot_sync(B200_FULL_MASK, need_init);
        while (need) {
            const int src = __ffs((int)need) - 1;
            need &= need - 1;
            const int qlen = __shfl_sync(B200_FULL_MASK, L.qlen, src);
            const int h0 = __shfl_sync(B200_FULL_MASK, L.h0, src);
            const int qstep = __shfl_sync(B200_FULL_MASK, L.qstep, src);
            const unsigned long long qb = __shfl_sync(B200_FULL_MASK, (unsigned long long)(size_t)L.qbase, src);
            const u8 *qp = (const u8 *)(size_t)qb;
            uint4 *he_src = (uint4 *)d8_raw + wbase + src;
            uint2 *qs_src = (uint2 *)(d8_raw + (size_t)n_groups * D8_THREADS * 16) + wbase + src;
            const int jtop = ((qlen >> 3) + 1) << 3;                                      // whole groups up to the one holding column qlen
            for (int j = (int)lane; j < jtop; j += 32) {
                int v = 0;
                if (j == 0) v = h0;
                else if (j == 1) v = h0 > oe_ins ? h0 - oe_ins : 0;
                else if (j <= qlen) {
                    const int vprev = h0 - oe_ins - (j - 2) * e_ins;
                    v = (h0 > oe_ins && vprev > e_ins) ? vprev - e_ins : 0;
                }
                const u32 qv = j < qlen ? (u32)qp[j * qstep] : 0u;
                ((u16 *)(he_src + (j >> 3) * D8_THREADS))[j & 7] = (u16)v;                 // h = v, e = 0
                ((u8 *)(qs_src + (j >> 3) * D8_THREADS))[j & 7] = (u8)(qv * 0x11u + 0x80u);
            }
        }
        if (phased) {                                                                       // open the gate for the earliest waiting phase
            if (need_init) parked = true;
            if (!__any_sync(B200_FULL_MASK, active && !parked)) {
                const int ph = parked ? L.side * 2 + L.t : 99;
                const int pmin = __reduce_min_sync(B200_FULL_MASK, ph);
                if (ph == pmin) parked = false;
            }
        }
        need_init = false;
        __syncwarp();
        if (__all_sync(B200_FULL_MASK, exhausted && !active && !(TAIL && have_done))) break;
    }
    if (n_ksw) {
        atomicAdd(&P.stats[CNT_KSW], (unsigned long long)n_ksw);
        atomicAdd(&P.stats[CNT_CELLS], (unsigned long long)n_cells);
    }
}

// ---- job order: counting sort of every bin's job slots by ExtJob::key, descending (longest flanks first, so the tail of a
// launch is short work), so that the lanes of a warp hold extensions of the same shape and run their rows in step
#define JS_KEYS 1024
__device__ __forceinline__ bool js_locate(const Ext2Params &P, int round, unsigned t, int &bin, unsigned &slot)
{
    const unsigned *nj = P.njobs + 8 * round;
    bin = 0; slot = t;
    if (slot >= nj[0]) { slot -= nj[0]; bin = 2; if (slot >= nj[2]) { slot -= nj[2]; bin = 1; if (slot >= nj[1]) return false; } }
    return true;
}
__global__ void job_hist_kernel(Ext2Params P, int round, int n_jobs)
{
    const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    int bin; unsigned slot;
    if ((int)t >= n_jobs || !js_locate(P, round, t, bin, slot)) return;
    atomicAdd(&P.sort_cnt[bin * JS_KEYS + (d8_job_slot(P, round & 1, bin, slot)->key & (JS_KEYS - 1))], 1u);
}
__global__ void job_scan_kernel(Ext2Params P)                // one block of 3 warps: counts -> first position of each key
{
    const int bin = threadIdx.x >> 5, lane = threadIdx.x & 31;
    u32 *c = P.sort_cnt + bin * JS_KEYS;
    unsigned run = 0;
    for (int base = JS_KEYS - 32; base >= 0; base -= 32) {   // descending keys; 32 keys per step, lane 0 takes the highest
        const int k = base + 31 - lane;
        const unsigned v = c[k];
        unsigned inc = v;
        for (int d = 1; d < 32; d <<= 1) { const unsigned o = __shfl_up_sync(B200_FULL_MASK, inc, d); if (lane >= d) inc += o; }
        c[k] = run + inc - v;
        run += __shfl_sync(B200_FULL_MASK, inc, 31);
    }
}
__global__ void job_scatter_kernel(Ext2Params P, int round, int n_jobs)
{
    const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    int bin; unsigned slot;
    if ((int)t >= n_jobs || !js_locate(P, round, t, bin, slot)) return;
    const unsigned pos = atomicAdd(&P.sort_cnt[bin * JS_KEYS + (d8_job_slot(P, round & 1, bin, slot)->key & (JS_KEYS - 1))], 1u);
    P.perm[(size_t)bin * P.jobs_cap + pos] = slot;
}
int launch_ext2_sort(const Ext2Params &p, int round, int n_jobs, int n_sms, cudaStream_t st)
{
    B200_CHECK(cudaMemsetAsync(p.sort_cnt, 0, 3 * JS_KEYS * 4, st));
    B200_LAUNCH(job_hist_kernel, (n_jobs + 255) / 256, 256, 0, st, p, round, n_jobs);
    B200_LAUNCH(job_scan_kernel, 1, 96, 0, st, p);
    B200_LAUNCH(job_scatter_kernel, (n_jobs + 255) / 256, 256, 0, st, p, round, n_jobs);
    B200_CHECK(cudaGetLastError());
    (void)n_sms;
    return 0;
}

size_t ext2_dp8_smem(int n_groups) { return (size_t)n_groups * D8_THREADS * (16 + 8); }

int launch_ext2_dp8(const Ext2Params &p, int round, int bin, int n_groups, int grid, int rowsync, int phased, cudaStream_t st)
{
    const size_t smem = ext2_dp8_smem(n_groups);
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute((ext_dp8_kernel<true, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    B200_CHECK(cudaFuncSetAttribute((ext_dp8_kernel<false, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
    if (rowsync) B200_LAUNCH((ext_dp8_kernel<true, false>), grid, D8_THREADS, smem, st, p, round, bin, n_groups, phased);
    else B200_LAUNCH((ext_dp8_kernel<false, false>), grid, D8_THREADS, smem, st, p, round, bin, n_groups, 0);
    B200_CHECK(cudaGetLastError());
    return 0;
}

// the rest of the chunk in one launch: every job of `round` and every later seed of their reads (p.tail_jobs: one slot per lane)
int launch_ext2_dp8_tail(const Ext2Params &p, int round, int n_groups, int grid, cudaStream_t st)
{
    const size_t smem = ext2_dp8_smem(n_groups);
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute((ext_dp8_kernel<true, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
    B200_LAUNCH((ext_dp8_kernel<true, true>), grid, D8_THREADS, smem, st, p, round, 2, n_groups, 0);
    B200_CHECK(cudaGetLastError());
    return 0;
}
