// Stage 1: SMEM seeding (mem_collect_intv, bwamem.c:137-185; bwt_smem1a, bwt.c:289-351;
// bwt_seed_strategy1, bwt.c:358-379).
//
// Design.  The reference walks each read through three nested-loop passes whose innermost operation
// is always one bwt_extend (= one or two random 64-byte Occ-block reads, ~460 per 150-bp read).  On the
// GPU that dependent gather is the whole cost, so every thread runs a *flattened state machine* for one
// read with exactly ONE bwt_extend call site per iteration: whatever phase the 32 reads of a warp are
// in (forward pass, lock-step backward pass, re-seeding, LAST-like sweep), all lanes reconverge at the
// gather and issue their block loads together; the control logic between two gathers is a few dozen
// ALU instructions.  Threads are persistent and pull the next read from an atomic work counter as soon
// as they finish one, so lanes never idle while the batch has work left.
//
// Result equivalence: the emitted interval set is the reference's; intervals are fully determined by
// their (start,end) substring, so equal sort keys mean identical records and the final order after the
// sort by `info` (bwamem.c:184) is unique.
#include "b200_internal.h"

#define ST_NEXT 0
#define ST_FWD 1
#define ST_BWD 2
#define ST_P3 3
#define ST_DONE 4

__device__ __forceinline__ void store_intv(Intv *p, u64 x0, u64 x1, u64 x2, u64 info)
{
    uint4 *q = (uint4 *)p;
    q[0] = make_uint4((u32)x0, (u32)(x0 >> 32), (u32)x1, (u32)(x1 >> 32));
    q[1] = make_uint4((u32)x2, (u32)(x2 >> 32), (u32)info, (u32)(info >> 32));
}
__device__ __forceinline__ Intv load_intv(const Intv *p)
{
    const uint4 *q = (const uint4 *)p;
    uint4 a = q[0], b = q[1];
    Intv r;
    r.x0 = (u64)a.y << 32 | a.x; r.x1 = (u64)a.w << 32 | a.z;
    r.x2 = (u64)b.y << 32 | b.x; r.info = (u64)b.w << 32 | b.z;
    return r;
}

// One read (collect = 1: all three passes, length filter) or one bwt_smem1a call (collect = 0).
// Returns the number of intervals written to out (cap_out max; *ovf set when exceeded).
__device__ __forceinline__ int smem_machine(const DevIndex &ix, const u8 *q, int len, int collect,
                                            int x_single, int min_intv_single, u64 max_intv,
                                            int min_seed_len, int split_len, int split_width, u64 max_mem_intv,
                                            Intv *LA, Intv *LB, int list_cap,
                                            Intv *out, int cap_out, int *ret_single, int *ovf,
                                            u64 &n_ext, u64 &n_two)
{
    int n_out = 0, st = ST_NEXT, pass = collect ? 1 : 0;
    int x = 0, k2 = 0, old_n = 0, sx = 0;
    int i = 0, j = 0, min_intv = 1, nfw = 0, nprev = 0, ncurr = 0, last_start = 0x7fffffff, ret = 0;
    int ik_end = 0;
    u64 ik0 = 0, ik1 = 0, ik2 = 0;
    Intv *prev = LA, *curr = LB;
    const int min_len = collect ? min_seed_len : 0;
    int single_started = 0;

    for (;;) {
        u64 r0 = 0, r1 = 0, r2 = 0; int rback = 0;
        Intv pj; pj.x0 = pj.x1 = pj.x2 = pj.info = 0;
        bool need = false;
        // ---- advance the control state until the next bwt_extend is needed
        while (!need) {
            if (st == ST_NEXT) {
                bool begin = false;
                if (pass == 0) {
                    if (single_started) { st = ST_DONE; continue; }
                    single_started = 1;
                    sx = x_single; min_intv = min_intv_single < 1 ? 1 : min_intv_single;
                    if (q[sx] > 3) { ret = sx + 1; st = ST_DONE; continue; }       // bwt.c:296
                    begin = true;
                } else if (pass == 1) {                                              // bwamem.c:144-154
                    while (x < len && q[x] > 3) ++x;
                    if (x >= len) { pass = 2; old_n = n_out < cap_out ? n_out : cap_out; k2 = 0; continue; }
                    sx = x; min_intv = 1; begin = true;
                } else if (pass == 2) {                                              // bwamem.c:156-165
                    while (k2 < old_n) {
                        Intv p = load_intv(&out[k2]);
                        int s0 = (int)(p.info >> 32), e0 = (int)(u32)p.info;
                        if (e0 - s0 < split_len || p.x2 > (u64)split_width) { ++k2; continue; }
                        sx = (s0 + e0) >> 1; min_intv = (int)p.x2 + 1;
                        break;
                    }
                    if (k2 >= old_n) { pass = 3; x = 0; continue; }
                    if (q[sx] > 3) { ++k2; continue; }                               // cannot happen inside an exact match
                    begin = true;
                } else {                                                             // bwamem.c:167-182
                    if (max_mem_intv == 0) { st = ST_DONE; continue; }
                    while (x < len && q[x] > 3) ++x;
                    if (x >= len) { st = ST_DONE; continue; }
                    dev_set_intv(ix, q[x], ik0, ik1, ik2);
                    i = x + 1; st = ST_P3;
                    continue;
                }
                if (begin) {                                                         // bwt.c:301-304
                    dev_set_intv(ix, q[sx], ik0, ik1, ik2);
                    ik_end = sx + 1; i = sx + 1; nfw = 0; last_start = 0x7fffffff;
                    st = ST_FWD;
                }
            } else if (st == ST_FWD) {
                bool stop = false;
                if (i >= len) stop = true;                                           // bwt.c:321
                else if (ik2 < max_intv) stop = true;                                // bwt.c:305-307
                else if (q[i] < 4) { r0 = ik0; r1 = ik1; r2 = ik2; rback = 0; need = true; }
                else stop = true;                                                    // bwt.c:316-318
                if (stop) {
                    if (nfw < list_cap) { store_intv(&LA[list_cap - 1 - nfw], ik0, ik1, ik2, (u64)ik_end); ++nfw; }
                    else *ovf = 1;
                    goto fwd_done;
                }
            } else if (st == ST_BWD) {
                if (j >= nprev) {                                                    // bwt.c:343-344
                    bool fin = false;
                    if (ncurr == 0) fin = true;
                    else {
                        Intv *t = prev; prev = curr; curr = (t >= LA && t < LA + list_cap) ? LA : LB;
                        nprev = ncurr; ncurr = 0; --i; j = 0;
                        if (i < -1) fin = true;
                    }
                    if (fin) {
                        if (pass == 1) x = ret; else if (pass == 2) ++k2;
                        st = ST_NEXT;
                    }
                    continue;
                }
                int c = i < 0 ? -1 : (q[i] < 4 ? q[i] : -1);                         // bwt.c:327
                pj = load_intv(&prev[j]);
                if (c >= 0 && ik2 >= max_intv) { r0 = pj.x0; r1 = pj.x1; r2 = pj.x2; rback = 1; need = true; }
                else {                                                               // keep-hit branch, bwt.c:331-337
                    if (ncurr == 0 && (last_start == 0x7fffffff || i + 1 < last_start)) {
                        ik0 = pj.x0; ik1 = pj.x1; ik2 = pj.x2;
                        last_start = i + 1;
                        if ((int)(u32)pj.info - (i + 1) >= min_len) {
                            if (n_out < cap_out) store_intv(&out[n_out], pj.x0, pj.x1, pj.x2, pj.info | (u64)(i + 1) << 32);
                            else *ovf = 1;
                            ++n_out;
                        }
                    }
                    ++j;
                }
            } else if (st == ST_P3) {                                                // bwt.c:366-378
                if (i >= len) { x = len; st = ST_DONE; }
                else if (q[i] < 4) { r0 = ik0; r1 = ik1; r2 = ik2; rback = 0; need = true; }
                else { x = i + 1; st = ST_NEXT; }
            } else {
                if (ret_single) *ret_single = ret;
                return n_out;
            }
            continue;
        fwd_done:                                                                    // bwt.c:322-324
            prev = LA + (list_cap - nfw); nprev = nfw; curr = LB; ncurr = 0;
            ret = (int)(u32)load_intv(&prev[0]).info;
            i = sx - 1; j = 0; st = ST_BWD;
        }

        // ---- the single gather site
        u64 o0[4], o1[4], o2[4]; int two = 0;
        dev_extend(ix, r0, r1, r2, rback, o0, o1, o2, &two);
        ++n_ext; n_two += two;

        // ---- consume
        if (st == ST_FWD) {                                                          // bwt.c:308-315
            int c = 3 - q[i];
            bool brk = false;
            if (o2[c] != ik2) {
                if (nfw < list_cap) { store_intv(&LA[list_cap - 1 - nfw], ik0, ik1, ik2, (u64)ik_end); ++nfw; }
                else *ovf = 1;
                if (o2[c] < (u64)min_intv) brk = true;
            }
            if (brk) {
                prev = LA + (list_cap - nfw); nprev = nfw; curr = LB; ncurr = 0;
                ret = (int)(u32)load_intv(&prev[0]).info;
                i = sx - 1; j = 0; st = ST_BWD;
            } else { ik0 = o0[c]; ik1 = o1[c]; ik2 = o2[c]; ik_end = i + 1; ++i; }
        } else if (st == ST_BWD) {                                                   // bwt.c:331-341
            int c = q[i];
            if (o2[c] < (u64)min_intv) {
                if (ncurr == 0 && (last_start == 0x7fffffff || i + 1 < last_start)) {
                    ik0 = pj.x0; ik1 = pj.x1; ik2 = pj.x2;
                    last_start = i + 1;
                    if ((int)(u32)pj.info - (i + 1) >= min_len) {
                        if (n_out < cap_out) store_intv(&out[n_out], pj.x0, pj.x1, pj.x2, pj.info | (u64)(i + 1) << 32);
                        else *ovf = 1;
                        ++n_out;
                    }
                }
            } else if (ncurr == 0 || o2[c] != load_intv(&curr[ncurr - 1]).x2) {
                store_intv(&curr[ncurr], o0[c], o1[c], o2[c], pj.info);
                ++ncurr;
            }
            ++j;
        } else {                                                                     // ST_P3, bwt.c:369-375
            int c = 3 - q[i];
            if (o2[c] < max_mem_intv && i - x >= min_seed_len) {
                if (o2[c] > 0) {                                                     // bwamem.c:174
                    if (n_out < cap_out) store_intv(&out[n_out], o0[c], o1[c], o2[c], (u64)x << 32 | (u64)(i + 1));
                    else *ovf = 1;
                    ++n_out;
                }
                x = i + 1; st = ST_NEXT;
            } else { ik0 = o0[c]; ik1 = o1[c]; ik2 = o2[c]; ++i; }
        }
    }
}

__device__ __forceinline__ void sort_by_info(Intv *a, int n)
{
    for (int i = 1; i < n; ++i) {
        Intv v = load_intv(&a[i]);
        int j = i;
        while (j > 0) {
            Intv u = load_intv(&a[j - 1]);
            if (u.info <= v.info) break;
            store_intv(&a[j], u.x0, u.x1, u.x2, u.info);
            --j;
        }
        if (j != i) store_intv(&a[j], v.x0, v.x1, v.x2, v.info);
    }
}

__global__ void __launch_bounds__(128) smem_collect_kernel(SmemParams P)
{
    const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    Intv *LA = P.lists + gtid * 2 * (size_t)P.list_cap, *LB = LA + P.list_cap;
    u64 n_ext = 0, n_two = 0;
    int ovf = 0;
    const unsigned n_work = P.use_retry ? *P.retry_n : (unsigned)P.n_reads;
    for (;;) {
        unsigned r = atomicAdd(P.work_counter, 1u);
        if (r >= n_work) break;
        if (P.use_retry) r = (unsigned)P.retry_ids[r];
        const i64 o = P.off[r];
        const int len = (int)(P.off[r + 1] - o);
        Intv *out = P.out + (size_t)r * P.cap_out;
        int n = 0;
        if (len >= P.min_seed_len) {                                                 // bwamem.c:267
            n = smem_machine(P.ix, P.seq + o, len, 1, 0, 1, 0, P.min_seed_len, P.split_len, P.split_width,
                             P.max_mem_intv, LA, LB, P.list_cap, out, P.cap_out, nullptr, &ovf, n_ext, n_two);
            if (n > P.cap_out) { atomicMax(P.overflow, n); n = P.cap_out; }
            sort_by_info(out, n);                                                    // bwamem.c:184
        }
        P.out_n[r] = n;
    }
    if (ovf) atomicMax(P.overflow + 1, 1);
    if (n_ext) { atomicAdd(&P.stats[CNT_EXTEND], (unsigned long long)n_ext); atomicAdd(&P.stats[CNT_TWO_BLOCK], (unsigned long long)n_two); }
}

// one bwt_smem1a call per thread (the scalar drop-in and b200_smem1a_batch)
__global__ void __launch_bounds__(128) smem_single_kernel(SmemParams P)
{
    const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    Intv *LA = P.lists + gtid * 2 * (size_t)P.list_cap, *LB = LA + P.list_cap;
    u64 n_ext = 0, n_two = 0;
    int ovf = 0;
    for (;;) {
        unsigned r = atomicAdd(P.work_counter, 1u);
        if (r >= (unsigned)P.n_reads) break;
        const i64 o = P.off[r];
        const int len = (int)(P.off[r + 1] - o);
        Intv *out = P.out + (size_t)r * P.cap_out;
        int ret = 0;
        int n = smem_machine(P.ix, P.seq + o, len, 0, P.px[r], P.pmin_intv[r], P.max_intv, 0, 0, 0, 0,
                             LA, LB, P.list_cap, out, P.cap_out, &ret, &ovf, n_ext, n_two);
        if (n > P.cap_out) { atomicMax(P.overflow, n); n = P.cap_out; }
        // bwt.c:346: reversed so that the list is sorted by start coordinate
        for (int a = 0, b = n - 1; a < b; ++a, --b) {
            Intv u = load_intv(&out[a]), v = load_intv(&out[b]);
            store_intv(&out[a], v.x0, v.x1, v.x2, v.info);
            store_intv(&out[b], u.x0, u.x1, u.x2, u.info);
        }
        P.out_n[r] = n;
        P.out_ret[r] = ret;
    }
    if (ovf) atomicMax(P.overflow + 1, 1);
}

// ASCII -> nt4 in place (nst_nt4_table, bntseq.c:40-57; mem_align1_core, bwamem.c:1067-1068).
__global__ void pack_reads_kernel(u8 *seq, i64 n)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    const i64 stride = (i64)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        u8 c = seq[i];
        if (c >= 4) {
            u8 u = c & 0xdf;   // fold case
            seq[i] = u == 'A' ? 0 : u == 'C' ? 1 : u == 'G' ? 2 : u == 'T' ? 3 : (c == '-' ? 5 : 4);
        }
    }
}

int launch_pack_reads(u8 *seq, i64 n_bytes, cudaStream_t st)
{
    if (n_bytes <= 0) return 0;
    int grid = (int)((n_bytes + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    B200_LAUNCH(pack_reads_kernel, grid, 256, 0, st, seq, n_bytes);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_smem_collect(const SmemParams &p, int grid, int block, cudaStream_t st)
{
    B200_LAUNCH(smem_collect_kernel, grid, block, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_smem_single(const SmemParams &p, int grid, int block, cudaStream_t st)
{
    B200_LAUNCH(smem_single_kernel, grid, block, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}
