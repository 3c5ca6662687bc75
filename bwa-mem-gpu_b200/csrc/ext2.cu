// Stage 3, short-read fast path: mem_chain2aln (bwamem.c:639-793) as a round-based pipeline.
//
// The reference decides seed by seed whether to extend (the containment test against every region already
// produced for the read, bwamem.c:678-713), so per read the extensions form a dependent sequence -- but the
// extension of ONE seed (left ksw_extend2, right ksw_extend2 seeded with the left score, optional wider-band
// retries, bwamem.c:722-778) is a pure function of the seed and its chain window.  The batch is therefore
// advanced in rounds:
//   ext_plan_kernel   one thread per read: consume the finished job (seedcov, bwamem.c:781-790), walk the
//                     chains/seeds in the reference's order, skip contained seeds, emit the next JOB
//   ext_dp_kernel     one JOB PER LANE: both sides + retries of one seed, bit-exact ksw_extend2 (ksw.c:380-479)
//   ext_finish_kernel one thread per read: mem_sort_dedup_patch (bwamem.c:444-496) + is_alt (bwamem.c:1091-1095)
// Round k holds the k-th extended seed of every read (typically 1-2 rounds carry almost all work).
//
// DP kernel.  A warp runs 32 independent DPs in lock step over a flattened cell loop (8 cells, then a control
// slot for row ends / job changes), which costs ~25 instructions per cell instead of a warp-wide scan per
// 32-column chunk.  The H/E row of the reference (eh[], including the stale out-of-band entries the algorithm
// observes) is staged in SHARED memory, one 32-bit word per column = h:14 | e:14 | query base:3, laid out
// [column][thread] so that lanes at different columns never conflict; the row of a new DP is initialised
// cooperatively by the whole warp (coalesced query read).  Cell arithmetic uses the DPX fused ops
// (__vimax3_s32, __viaddmax_s32_relu); z-drop / zero-row / end-of-target exits are per-lane state changes and
// warp votes (__ballot_sync/__all_sync) drive initialisation and termination.  Used when every score fits
// 14 bits (max_len * a < 16384) and the row fits the shared-memory budget; otherwise the general warp-per-read
// kernel (extend.cu) runs.
#include "b200_internal.h"
#include "extend_common.cuh"

#define DP_THREADS 64
#define DP_HMASK 0x3fffu
#define DP_QMASK 0x70000000u

__device__ __forceinline__ ExtJob *job_slot(const Ext2Params &P, int parity, int bin, unsigned slot)
{
    if (bin == 1) return P.jobs_mid[parity] + slot;
    return bin == 0 ? P.jobs[parity] + slot : P.jobs[parity] + (P.jobs_cap - 1 - (int)slot);
}

// ------------------------------------------------------------------------------------------------ planning
#include "ext_plan.cuh"

__global__ void __launch_bounds__(128) ext_plan_kernel(Ext2Params P, int round)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    int r;
    const ExtJob *done = nullptr;
    if (round == 0) { if (t >= P.n_reads) return; r = t; }
    else {
        const unsigned *nj = P.njobs + 8 * (round - 1);
        unsigned u = (unsigned)t;
        int bin = 0;
        if (u >= nj[0]) { u -= nj[0]; bin = 2; if (u >= nj[2]) { u -= nj[2]; bin = 1; if (u >= nj[1]) return; } }
        done = job_slot(P, (round + 1) & 1, bin, u);
        r = done->read;
    }
    ExtJob j;
    if (!ext_plan_read(P, r, done, round == 0, &j)) return;
    const int ql = j.qbeg, qr = j.lq - (j.qbeg + j.slen);
    const int flank = ql > qr ? ql : qr;
    const int bin = flank <= P.bin_max[0] ? 0 : flank <= P.bin_max[1] ? 1 : 2;
    const unsigned slot = atomicAdd(&P.njobs[8 * round + bin], 1u);
    *job_slot(P, round & 1, bin, slot) = j;
}

// ------------------------------------------------------------------------------------------------ DP
struct DpLane {
    // job
    ExtJob *job;
    const u8 *qbase;          // first query base of the current side; column j is qbase[j * qstep]
    int qstep;
    i64 fpos; int fstep, comp; u32 pacw, pacw_tag;     // forward-strand coordinate of the current target row
    int side, t, a_score, sc0, prev;
    // ksw_extend2 state (ksw.c:380-479)
    int qlen, tlen, w, h0;
    int i, j, beg, end, h1, f, m, mj, first_nz, last_nz;
    u32 srow_lo, srow_hi;                 // substitution scores of the current target base vs query base 0..7
    int max, max_i, max_j, max_ie, gscore, max_off;
};

__device__ __forceinline__ int dp_target_base(const DevIndex &ix, DpLane &L)
{
    const u32 tag = (u32)(L.fpos >> 4);
    if (tag != L.pacw_tag) { L.pacw = ((const u32 *)ix.pac)[L.fpos >> 4]; L.pacw_tag = tag; }
    const int f = (int)L.fpos;
    const int sh = (((f >> 2) & 3) << 3) + ((~f & 3) << 1);
    return (int)((L.pacw >> sh) & 3) ^ L.comp;
}

// signed byte q of the 8-byte matrix row {hi:lo}: one PRMT (byte q into the low byte, its sign bit replicated
// into the three upper bytes).  The __byte_perm intrinsic documents only 3 selector bits, so prmt is spelled out.
__device__ __forceinline__ int dp_score(u32 lo, u32 hi, u32 q)
{
#ifdef B200_EMU
    return (int)(int8_t)(((u64)hi << 32 | lo) >> (8 * q));
#else
    u32 r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(q * 0x1111u + 0x8880u));
    return (int)r;
#endif
}

// rowsync / phased: the scheduling of ext3.cu (a lane finishes its row before the warp's control slot; warp-wide phases over
// 32 shape-ordered jobs), here as run-time flags.
__global__ void __launch_bounds__(DP_THREADS) ext_dp_kernel(Ext2Params P, int round, int bin, int rowsync, int phased)
{
    B200_DYN_SMEM(dp_raw);
    __shared__ u32 s_rows[5 * 2];                          // matrix row of each target base, 8 bytes (5 used)
    u32 *he = (u32 *)dp_raw + threadIdx.x;                 // column j of this lane: he[j * DP_THREADS]
    const unsigned lane = threadIdx.x & 31;
    u32 *he_warp = (u32 *)dp_raw + (threadIdx.x & ~31u);
    if (threadIdx.x < 10) {
        u32 v = 0;
        for (int k = 0; k < 4; ++k) {
            const int col = (threadIdx.x & 1) * 4 + k;
            if (col < 5) v |= (u32)(u8)P.opt.mat[(threadIdx.x >> 1) * 5 + col] << (8 * k);
        }
        s_rows[threadIdx.x] = v;
    }
    __syncthreads();
    const DevIndex &ix = P.ix;
    const mem_opt_t &opt = P.opt;
    const int o_del = opt.o_del, e_del = opt.e_del, o_ins = opt.o_ins, e_ins = opt.e_ins;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins, zdrop = opt.zdrop;
    const unsigned njobs = P.njobs[8 * round + bin];
    unsigned *work_counter = P.njobs + 8 * round + 4 + bin;
    const i64 l_pac = ix.l_pac;
    u64 n_ksw = 0, n_cells = 0;
    unsigned ncell32 = 0;                  // cells of the current job (flushed into n_cells when it ends)

    DpLane L;
    memset(&L, 0, sizeof(L));
    bool active = false, exhausted = false, need_init = false, parked = false;

    for (;;) {
        // ---------------------------------------------------------------- hot slot: up to 8 cells of the current row
        // The 8 H/E words are fetched first (independent LDS, the row is padded by 8 columns), the cells are then
        // evaluated from registers; the substitution score comes from the target base's matrix row held in two
        // registers (byte select with sign replication), so a cell has no memory latency on its dependency chain.
        for (bool first = true; active && !parked && (rowsync ? L.j < L.end : first); first = false) {
            u32 *p = he + L.j * DP_THREADS;
            u32 w8[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) w8[u] = p[u * DP_THREADS];
            const int rem = L.end - L.j;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (u < rem) {
                    const u32 wv = w8[u];
                    const int jj = L.j + u;
                    int M = (int)(wv & DP_HMASK);
                    const int e = (int)((wv >> 14) & DP_HMASK);
                    const int sc = dp_score(L.srow_lo, L.srow_hi, wv >> 28);
                    M = M ? M + sc : 0;                                                   // ksw.c:433
                    const int h = __vimax3_s32(M, e, L.f);                                // ksw.c:434-435
                    const int en = __viaddmax_s32_relu(e, -e_del, M - oe_del);            // ksw.c:439-442
                    p[u * DP_THREADS] = (wv & DP_QMASK) | (u32)L.h1 | ((u32)en << 14);    // ksw.c:432,442
                    if (L.h1 | en) { L.last_nz = jj; if (L.first_nz > jj) L.first_nz = jj; }
                    L.h1 = h;
                    if (h >= L.m) { L.m = h; L.mj = jj; }                                 // ksw.c:437-438 (last column on ties)
                    L.f = __viaddmax_s32_relu(L.f, -e_ins, M - oe_ins);                   // ksw.c:444-447
                }
            }
            if (rem > 0) L.j += rem < 8 ? rem : 8;
        }
        if (rowsync) __syncwarp();
        // ---------------------------------------------------------------- control slot
        bool ksw_fin = false;
        if (active && !parked && L.j >= L.end) {                                          // row i finished
            const int jexit = L.beg < L.end ? L.end : L.beg;
            u32 *p = he + L.end * DP_THREADS;
            *p = (*p & DP_QMASK) | (u32)L.h1;                                             // ksw.c:449
            if (jexit == L.qlen) {                                                        // ksw.c:450-453
                L.max_ie = L.gscore > L.h1 ? L.max_ie : L.i;
                L.gscore = L.gscore > L.h1 ? L.gscore : L.h1;
            }
            if (L.m == 0) ksw_fin = true;                                                 // ksw.c:454
            else {
                if (L.m > L.max) {
                    L.max = L.m; L.max_i = L.i; L.max_j = L.mj;
                    int off = L.mj - L.i; off = off < 0 ? -off : off;
                    L.max_off = L.max_off > off ? L.max_off : off;
                } else if (zdrop > 0) {                                                   // ksw.c:458-464
                    if (L.i - L.max_i > L.mj - L.max_j) {
                        if (L.max - L.m - ((L.i - L.max_i) - (L.mj - L.max_j)) * e_del > zdrop) ksw_fin = true;
                    } else {
                        if (L.max - L.m - ((L.mj - L.max_j) - (L.i - L.max_i)) * e_ins > zdrop) ksw_fin = true;
                    }
                }
                if (!ksw_fin) {                                                           // ksw.c:466-469
                    const int nbeg = L.first_nz != 0x7fffffff ? L.first_nz : L.end;
                    int jl;
                    if (L.h1 != 0) jl = L.end; else jl = L.last_nz >= nbeg ? L.last_nz : nbeg - 1;
                    L.beg = nbeg;
                    L.end = jl + 2 < L.qlen ? jl + 2 : L.qlen;
                    if (++L.i >= L.tlen) ksw_fin = true;
                    else {                                                                // next row, ksw.c:411-423
                        L.fpos += L.fstep;
                        { const int tb = dp_target_base(ix, L); L.srow_lo = s_rows[2 * tb]; L.srow_hi = s_rows[2 * tb + 1]; }
                        if (L.beg < L.i - L.w) L.beg = L.i - L.w;
                        if (L.end > L.i + L.w + 1) L.end = L.i + L.w + 1;
                        if (L.end > L.qlen) L.end = L.qlen;
                        L.h1 = 0;
                        if (L.beg == 0) { L.h1 = L.h0 - (o_del + e_del * (L.i + 1)); if (L.h1 < 0) L.h1 = 0; }
                        L.f = 0; L.m = 0; L.mj = -1; L.first_nz = 0x7fffffff; L.last_nz = -1; L.j = L.beg;
                        if (L.end > L.beg) ncell32 += (unsigned)(L.end - L.beg);
                    }
                }
            }
        }
        // job / side / retry transitions; loops only when a DP has no rows (tlen <= 0)
        const bool may_fetch = !phased || !__any_sync(B200_FULL_MASK, active);              // phased: the warp takes 32 jobs at a time
        while ((active && ksw_fin) || (!active && !exhausted && may_fetch)) {
            int next_side = -1, next_t = 0;
            if (!active) {                                                                // fetch a job
                const unsigned slot = atomicAdd(work_counter, 1u);
                if (slot >= njobs) { exhausted = true; break; }
                L.job = job_slot(P, round & 1, bin, P.perm ? P.perm[(size_t)bin * P.jobs_cap + slot] : slot);
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
                    n_cells += ncell32; ncell32 = 0;
                    continue;
                }
                L.sc0 = L.a_score; L.prev = L.a_score;
            }
            // set up ksw_extend2 for (next_side, next_t)
            {
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
                { const int tb = dp_target_base(ix, L); L.srow_lo = s_rows[2 * tb]; L.srow_hi = s_rows[2 * tb + 1]; }
                if (L.end > L.w + 1) L.end = L.w + 1;                                     // row 0: beg = 0 >= -w
                if (L.end > L.qlen) L.end = L.qlen;
                L.h1 = L.h0 - (o_del + e_del); if (L.h1 < 0) L.h1 = 0;
                L.f = 0; L.m = 0; L.mj = -1; L.first_nz = 0x7fffffff; L.last_nz = -1; L.j = 0;
                if (L.end > 0) ncell32 += (unsigned)L.end;
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
            u32 *col = he_warp + src;
            for (int j = (int)lane; j <= qlen; j += 32) {
                int v;
                if (j == 0) v = h0;
                else if (j == 1) v = h0 > oe_ins ? h0 - oe_ins : 0;
                else {
                    const int vprev = h0 - oe_ins - (j - 2) * e_ins;
                    v = (h0 > oe_ins && vprev > e_ins) ? vprev - e_ins : 0;
                }
                const u32 qv = j < qlen ? (u32)qp[j * qstep] : 0u;
                col[j * DP_THREADS] = (u32)v | (qv << 28);
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
        if (__all_sync(B200_FULL_MASK, exhausted && !active)) break;
    }
    if (n_ksw) {
        atomicAdd(&P.stats[CNT_KSW], (unsigned long long)n_ksw);
        atomicAdd(&P.stats[CNT_CELLS], (unsigned long long)n_cells);
    }
}

// ------------------------------------------------------------------------------------------------ finish
__global__ void __launch_bounds__(128) ext_finish_kernel(Ext2Params P)
{
    const size_t gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    i32 *H = P.eh + gt * 2 * (size_t)P.eh_cap, *E = H + P.eh_cap;
    for (;;) {
        const unsigned r = atomicAdd(P.work_counter, 1u);
        if (r >= (unsigned)P.n_reads) break;
        mem_alnreg_t *av = P.regs + P.chn_off[r];
        int nav = P.rstate[r].nav;
        if (nav > 1) nav = sort_dedup_patch(P.ix, P.opt, P.seq + P.off[r], nav, av, H, E);   // bwamem.c:1083
        for (int i = 0; i < nav; ++i)
            if (av[i].rid >= 0 && P.ix.ann_alt[av[i].rid]) av[i].is_alt = 1;                 // bwamem.c:1091-1095
        P.reg_off[r] = P.chn_off[r];
        P.reg_n[r] = nav;
    }
}

int launch_ext2_plan(const Ext2Params &p, int round, int n_threads, cudaStream_t st)
{
    if (n_threads <= 0) return 0;
    B200_LAUNCH(ext_plan_kernel, (n_threads + 127) / 128, 128, 0, st, p, round);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_ext2_dp(const Ext2Params &p, int round, int bin, int he_cols, int grid, int rowsync, int phased, cudaStream_t st)
{
    const size_t smem = (size_t)he_cols * DP_THREADS * 4;
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute(ext_dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
    B200_LAUNCH(ext_dp_kernel, grid, DP_THREADS, smem, st, p, round, bin, rowsync, rowsync ? phased : 0);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_ext2_finish(const Ext2Params &p, int grid, cudaStream_t st)
{
    B200_LAUNCH(ext_finish_kernel, grid, 128, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}
