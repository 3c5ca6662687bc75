// Stage 1, fast path: mem_collect_intv (bwamem.c:137-185) = three seeding passes built from bwt_smem1a
// (bwt.c:289-351) and bwt_seed_strategy1 (bwt.c:358-379), one read per thread over the occ32 index.
//
// Why this shape.  Per 150-bp read the reference issues ~700 DEPENDENT bwt_extend calls; each is one or two
// random index sectors followed by a little integer work.  The batch holds millions of independent reads, so
// the machine is filled across reads (one read per lane, persistent lanes pulling reads from an atomic
// counter) and the per-lane code is a flattened state machine with exactly ONE extension site per iteration:
// whatever phase the 32 reads of a warp are in (forward sweep, lock-step backward sweep, re-seeding,
// LAST-like sweep) all lanes meet at the gather, issue their sector loads together and do the same popcount
// arithmetic.  Everything that is not the gather is kept small:
//   * only the child ok[c] that the algorithm consumes is computed (occ32.cuh);
//   * the forward change-point list lives in SHARED memory, [entry][thread] interleaved 16/32-byte records
//     (conflict-free 128-bit accesses), and the backward sweep compacts it IN PLACE (the survivor list of a
//     sweep is written over entries already consumed), so there is no prev/curr pair and no global scratch;
//   * positions are 32-bit when seq_len < 2^32 (template PT), 64-bit otherwise;
//   * reads whose list would exceed the shared-memory capacity (deep repeats) are queued for the general
//     kernel (smem.cu) -- results are identical, only the scratch differs.
// The emitted interval set equals the reference's; records are determined by (start,end), so after the sort
// by `info` (bwamem.c:184) the order is unique.
#include "b200_internal.h"
#include "occ32.cuh"

#define S2_THREADS 128

enum { S2_TASK = 0, S2_FWD = 1, S2_BWD = 2, S2_P3 = 3, S2_FIN = 4 };

template <class PT> struct ListOps;
template <> struct ListOps<u32> {       // 16-byte entries
    static constexpr int kBytes = 16;
    static __device__ __forceinline__ void st(uint4 *L, int e, u32 x0, u32 x1, u32 x2, u32 end)
    { L[e * S2_THREADS] = make_uint4(x0, x1, x2, end); }
    static __device__ __forceinline__ void ld(const uint4 *L, int e, u32 &x0, u32 &x1, u32 &x2, u32 &end)
    { uint4 v = L[e * S2_THREADS]; x0 = v.x; x1 = v.y; x2 = v.z; end = v.w; }
};
template <> struct ListOps<u64> {       // 32-byte entries
    static constexpr int kBytes = 32;
    static __device__ __forceinline__ void st(uint4 *L, int e, u64 x0, u64 x1, u64 x2, u32 end)
    {
        L[(2 * e) * S2_THREADS] = make_uint4((u32)x0, (u32)(x0 >> 32), (u32)x1, (u32)(x1 >> 32));
        L[(2 * e + 1) * S2_THREADS] = make_uint4((u32)x2, (u32)(x2 >> 32), end, 0);
    }
    static __device__ __forceinline__ void ld(const uint4 *L, int e, u64 &x0, u64 &x1, u64 &x2, u32 &end)
    {
        uint4 a = L[(2 * e) * S2_THREADS], b = L[(2 * e + 1) * S2_THREADS];
        x0 = (u64)a.y << 32 | a.x; x1 = (u64)a.w << 32 | a.z; x2 = (u64)b.y << 32 | b.x; end = b.z;
    }
};

__device__ __forceinline__ void s2_emit(Intv *out, int &n_out, int cap_out, u64 x0, u64 x1, u64 x2, int start, int end)
{
    if (n_out < cap_out) {
        uint4 *q = (uint4 *)&out[n_out];
        q[0] = make_uint4((u32)x0, (u32)(x0 >> 32), (u32)x1, (u32)(x1 >> 32));
        q[1] = make_uint4((u32)x2, (u32)(x2 >> 32), (u32)end, (u32)start);
    }
    ++n_out;
}

template <class PT>
__global__ void __launch_bounds__(S2_THREADS) smem_fast_kernel(SmemParams P)
{
    B200_DYN_SMEM(s2_raw);
    uint4 *L = (uint4 *)s2_raw + threadIdx.x;              // this thread's column of the interleaved list
    typedef ListOps<PT> LO;
    const DevIndex &ix = P.ix;
    const int lcap = P.list_cap;
    const int min_seed_len = P.min_seed_len;
    const PT max_mem_intv = (PT)P.max_mem_intv;
    u64 n_ext = 0, n_two = 0;

    // per-read state
    const u8 *q = nullptr; Intv *out = nullptr;
    int len = 0, r = -1, n_out = 0;
    int st = S2_FIN, pass = 0, x = 0, k2 = 0, old_n = 0;
    int i = 0, sx = 0, nfw = 0, base = 0, nprev = 0, ncurr = 0, j = 0, last_start = 0, ret = 0;
    PT min_intv = 1, ik0 = 0, ik1 = 0, ik2 = 0, last_x2 = 0;
    PT p0 = 0, p1 = 0, p2 = 0; u32 pend = 0;               // the backward entry being extended
    bool spilled = false;

    for (;;) {
        // ------------------------------------------------------------ advance to the next extension request
        bool have = false; int c = 0, is_back = 0;
        PT r0 = 0, r1 = 0, r2 = 0;
        while (!have) {
            if (st == S2_FIN) {
                if (r >= 0) {                                              // finish the previous read
                    if (spilled) {
                        unsigned slot = atomicAdd(P.retry_n, 1u);
                        P.retry_ids[slot] = r;
                    } else {
                        int n = n_out;
                        if (n > P.cap_out) { atomicMax(P.overflow, n); n = P.cap_out; }
                        for (int a = 1; a < n; ++a) {                      // sort by info (bwamem.c:184)
                            uint4 v0 = ((uint4 *)&out[a])[0], v1 = ((uint4 *)&out[a])[1];
                            const u64 key = (u64)v1.w << 32 | v1.z;
                            int b = a;
                            while (b > 0) {
                                uint4 u1 = ((uint4 *)&out[b - 1])[1];
                                if (((u64)u1.w << 32 | u1.z) <= key) break;
                                ((uint4 *)&out[b])[0] = ((uint4 *)&out[b - 1])[0]; ((uint4 *)&out[b])[1] = u1;
                                --b;
                            }
                            if (b != a) { ((uint4 *)&out[b])[0] = v0; ((uint4 *)&out[b])[1] = v1; }
                        }
                        P.out_n[r] = n;
                    }
                }
                const unsigned nr = atomicAdd(P.work_counter, 1u);
                if (nr >= (unsigned)P.n_reads) { r = -1; break; }
                r = (int)nr;
                const i64 o = P.off[r];
                len = (int)(P.off[r + 1] - o);
                q = P.seq + o;
                out = P.out + (size_t)r * P.cap_out;
                n_out = 0; spilled = false;
                if (len < min_seed_len) continue;                           // bwamem.c:267 (st stays FIN -> out_n = 0)
                pass = 1; x = 0; st = S2_TASK;
            } else if (st == S2_TASK) {
                bool begin = false;
                if (pass == 1) {                                            // bwamem.c:144-154
                    while (x < len && q[x] > 3) ++x;
                    if (x >= len) { pass = 2; old_n = n_out < P.cap_out ? n_out : P.cap_out; k2 = 0; continue; }
                    sx = x; min_intv = 1; begin = true;
                } else if (pass == 2) {                                     // bwamem.c:156-165
                    while (k2 < old_n) {
                        const uint4 v1 = ((const uint4 *)&out[k2])[1];
                        const u64 sz = (u64)v1.y << 32 | v1.x;
                        const int s0 = (int)v1.w, e0 = (int)v1.z;
                        if (e0 - s0 < P.split_len || sz > (u64)P.split_width) { ++k2; continue; }
                        sx = (s0 + e0) >> 1; min_intv = (PT)sz + 1;
                        break;
                    }
                    if (k2 >= old_n) { pass = 3; x = 0; continue; }
                    begin = true;                                           // q[sx] < 4: inside an exact match
                } else {                                                    // bwamem.c:167-182
                    if (max_mem_intv == 0) { st = S2_FIN; continue; }
                    while (x < len && q[x] > 3) ++x;
                    if (x >= len) { st = S2_FIN; continue; }
                    const int b0 = q[x];
                    ik0 = (PT)ix.L2[b0] + 1; ik2 = (PT)(ix.L2[b0 + 1] - ix.L2[b0]); ik1 = (PT)ix.L2[3 - b0] + 1;
                    i = x + 1; st = S2_P3;
                    continue;
                }
                if (begin) {                                                // bwt.c:301-304
                    const int b0 = q[sx];
                    ik0 = (PT)ix.L2[b0] + 1; ik2 = (PT)(ix.L2[b0 + 1] - ix.L2[b0]); ik1 = (PT)ix.L2[3 - b0] + 1;
                    i = sx + 1; nfw = 0; st = S2_FWD;
                }
            } else if (st == S2_FWD) {
                if (i < len && q[i] < 4) { r0 = ik0; r1 = ik1; r2 = ik2; is_back = 0; c = 3 - q[i]; have = true; }
                else {                                                      // bwt.c:316-318, 321
                    if (nfw < lcap) LO::st(L, nfw, ik0, ik1, ik2, (u32)i); else spilled = true;
                    ++nfw;
                    ret = i; base = nfw - 1; nprev = nfw; ncurr = 0; j = 0; i = sx - 1; last_start = 0x7fffffff;
                    st = spilled ? S2_FIN : S2_BWD;
                }
            } else if (st == S2_BWD) {                                      // bwt.c:325-345, one entry per round
                const int cb = i < 0 ? -1 : (q[i] < 4 ? (int)q[i] : -1);
                LO::ld(L, base - j, p0, p1, p2, pend);
                if (cb >= 0) { r0 = p0; r1 = p1; r2 = p2; is_back = 1; c = cb; have = true; }
                else {
                    // no extension possible: only the first (longest) entry can be emitted, the sweep ends
                    if ((int)pend - (i + 1) >= min_seed_len) s2_emit(out, n_out, P.cap_out, p0, p1, p2, i + 1, (int)pend);
                    if (pass == 1) x = ret; else ++k2;
                    st = S2_TASK;
                }
            } else {                                                        // S2_P3, bwt.c:366-378
                if (i >= len) { st = S2_FIN; }
                else if (q[i] < 4) { r0 = ik0; r1 = ik1; r2 = ik2; is_back = 0; c = 3 - q[i]; have = true; }
                else { x = i + 1; st = S2_TASK; }
            }
        }
        if (!have) break;                                                   // no reads left for this lane

        // ------------------------------------------------------------ the single gather site
        PT n0, n1, n2; int two;
        occ32_child<PT>(ix, r0, r1, r2, is_back, c, n0, n1, n2, two);
        ++n_ext; n_two += two;

        // ------------------------------------------------------------ consume
        if (st == S2_FWD) {                                                 // bwt.c:308-315
            bool stop = false;
            if (n2 != ik2) {
                if (nfw < lcap) LO::st(L, nfw, ik0, ik1, ik2, (u32)i); else spilled = true;
                ++nfw;
                stop = n2 < min_intv;
            }
            if (stop) {
                ret = i; base = nfw - 1; nprev = nfw; ncurr = 0; j = 0; i = sx - 1; last_start = 0x7fffffff;
                st = spilled ? S2_FIN : S2_BWD;
            } else { ik0 = n0; ik1 = n1; ik2 = n2; ++i; }
        } else if (st == S2_BWD) {                                          // bwt.c:331-341
            if (n2 < min_intv) {
                if (ncurr == 0 && i + 1 < last_start) {
                    last_start = i + 1;
                    if ((int)pend - (i + 1) >= min_seed_len) s2_emit(out, n_out, P.cap_out, p0, p1, p2, i + 1, (int)pend);
                }
            } else if (ncurr == 0 || n2 != last_x2) {
                LO::st(L, base - ncurr, n0, n1, n2, pend);                  // ncurr <= j: slot already consumed
                ++ncurr; last_x2 = n2;
            }
            if (++j == nprev) {                                             // bwt.c:343-344
                if (ncurr == 0) { if (pass == 1) x = ret; else ++k2; st = S2_TASK; }
                else { nprev = ncurr; ncurr = 0; j = 0; --i; }
            }
        } else {                                                            // S2_P3, bwt.c:369-375
            if (n2 < max_mem_intv && i - x >= min_seed_len) {
                if (n2 > 0) s2_emit(out, n_out, P.cap_out, n0, n1, n2, x, i + 1);   // bwamem.c:174
                x = i + 1; st = S2_TASK;
            } else { ik0 = n0; ik1 = n1; ik2 = n2; ++i; }
        }
    }
    if (n_ext) { atomicAdd(&P.stats[CNT_EXTEND], (unsigned long long)n_ext); atomicAdd(&P.stats[CNT_TWO_BLOCK], (unsigned long long)n_two); }
}

int launch_smem_fast(const SmemParams &p, int grid, int wide, cudaStream_t st)
{
    const size_t smem = (size_t)p.list_cap * S2_THREADS * (wide ? 32 : 16);
#ifndef B200_EMU
    static bool attr_done[2] = { false, false };
    if (!attr_done[wide]) {
        if (wide) B200_CHECK(cudaFuncSetAttribute(smem_fast_kernel<u64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        else B200_CHECK(cudaFuncSetAttribute(smem_fast_kernel<u32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_done[wide] = true;
    }
#endif
    if (wide) B200_LAUNCH(smem_fast_kernel<u64>, grid, S2_THREADS, smem, st, p);
    else B200_LAUNCH(smem_fast_kernel<u32>, grid, S2_THREADS, smem, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}
