// mem_flt_chained_seeds / mem_seed_sw (bwamem.c:605-622, 578-603), seed-parallel: ONE Smith-Waterman PER LANE.
//
// mem_seed_sw consumes nothing but the SCORE of ksw_align2 (bwamem.c:600-602; 16-bit path, ksw_i16 ksw.c:232-334).
// That score is the plain local-alignment optimum, independent of the striped evaluation order: the vector code
// derives E(i+1,j) from H(i,j) before the lazy-F correction, but a path "insertion, then deletion" that this misses
// has the same score as "deletion, then insertion", which it evaluates (the lazy-F loop fixes every H of a row), so
// every H -- hence the maximum -- equals the textbook recurrence; and with windows under 200 x 200 bases and int8
// scores nothing reaches the 16-bit saturation bound.  (te / qe / score2 / tb / qb of the full ksw_align2 DO depend on
// the striped layout; mate rescue therefore keeps the lane-group restatement in ksw_align.cuh.  seedsw.cu, the
// restatement applied to this stage, stays as the checker of this file in tests/ and as the path for scoring
// parameters outside the bounds asserted by the launcher.)
//
// Three kernels: (1) one thread per long read lists the windows to score (same clamps as the reference, job = seed
// index + window); (2) one job per lane: the H/E row lives in shared memory as one 64-bit word per column PAIR
// ([pair][lane]), two columns per 16x2 DPX instruction like ext3.cu, all lanes of a warp advance one target row per
// step; (3) one thread per long read applies the score threshold and compacts each chain's seed array in order.
#include "ctx.h"

#define SW2_THREADS 64
#define SW2_SHORT_EXT 50        // MEM_SHORT_EXT, bwamem.c:570
#define SW2_SHORT_LEN 200       // MEM_SHORT_LEN, bwamem.c:571

struct SswJob { i64 rb; i32 seed, read, qb, qlen, tlen, pad; };

__device__ __forceinline__ u32 sw2_prmt(u32 a, u32 b, u32 s)
{
#ifdef B200_EMU
    return emu_prmt(a, b, s);
#else
    u32 r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(s));
    return r;
#endif
}

// ---- (1) windows, in the reference's order of tests (mem_seed_sw, bwamem.c:578-598)
__global__ void __launch_bounds__(128) ssw_jobs_kernel(SeedSwParams P, SswJob *jobs, unsigned *n_jobs)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= P.n_long) return;
    const int r = P.long_reads[u];
    const int l_query = (int)(P.off[r + 1] - P.off[r]);
    const i64 l_pac = P.ix.l_pac;
    const ChainRec *chains = P.chains + P.chn_off[r];
    const int n_chn = P.chn_n[r];
    for (int ci = 0; ci < n_chn; ++ci) {
        const ChainRec c = chains[ci];
        const Seed *seeds = P.cseeds + c.seed_off;
        for (int j = 0; j < c.n; ++j) {
            const Seed s = seeds[j];
            if (s.len >= SW2_SHORT_LEN) continue;
            int qb = s.qbeg, qe = s.qbeg + s.len;
            i64 rb = s.rbeg, re = s.rbeg + s.len;
            const i64 mid = (rb + re) >> 1;
            qb -= SW2_SHORT_EXT; qb = qb > 0 ? qb : 0;
            qe += SW2_SHORT_EXT; qe = qe < l_query ? qe : l_query;
            rb -= SW2_SHORT_EXT; rb = rb > 0 ? rb : 0;
            re += SW2_SHORT_EXT; re = re < l_pac << 1 ? re : l_pac << 1;
            if (rb < l_pac && l_pac < re) { if (mid < l_pac) re = l_pac; else rb = l_pac; }
            if (qe - qb >= SW2_SHORT_LEN || re - rb >= SW2_SHORT_LEN) continue;
            int rev;                                                                     // bns_fetch_seq: clamp to the contig of mid
            const int rid = dev_pos2rid(P.ix, dev_depos(P.ix, mid, &rev));
            i64 far_beg = P.ix.ann_off[rid], far_end = far_beg + P.ix.ann_len[rid];
            if (rev) { const i64 t = far_beg; far_beg = (l_pac << 1) - far_end; far_end = (l_pac << 1) - t; }
            rb = rb > far_beg ? rb : far_beg;
            re = re < far_end ? re : far_end;
            SswJob jb;
            jb.rb = rb; jb.seed = (i32)(c.seed_off + j); jb.read = r; jb.qb = qb; jb.qlen = qe - qb; jb.tlen = (int)(re - rb); jb.pad = 0;
            jobs[atomicAdd(n_jobs, 1u)] = jb;
        }
    }
}

// ---- (2) one local alignment per lane
__global__ void __launch_bounds__(SW2_THREADS) ssw_dp_kernel(SeedSwParams P, const SswJob *jobs, const unsigned *n_jobs_p, i32 *score_out, int np_cap)
{
    B200_DYN_SMEM(sw2_raw);
    __shared__ u32 s_rows[5 * 2];                                                        // matrix row of each target base, 8 bytes (5 used)
    uint2 *HE = (uint2 *)sw2_raw + threadIdx.x;                                          // pair p of this lane: HE[p * SW2_THREADS] = (H2, E2)
    u16 *QS = (u16 *)(sw2_raw + (size_t)np_cap * SW2_THREADS * 8) + threadIdx.x;         // prmt selector of the pair's two substitution scores
    if (threadIdx.x < 10) {
        u32 v = 0;
        for (int k = 0; k < 4; ++k) {
            const int col = (threadIdx.x & 1) * 4 + k;
            if (col < 5) v |= (u32)(u8)P.opt.mat[(threadIdx.x >> 1) * 5 + col] << (8 * k);
        }
        s_rows[threadIdx.x] = v;
    }
    __syncthreads();
    const mem_opt_t &opt = P.opt;
    const int oe_del = opt.o_del + opt.e_del, oe_ins = opt.o_ins + opt.e_ins;
    const u32 ne_del2 = ((u32)(-opt.e_del) & 0xffffu) * 0x10001u, noe_del2 = ((u32)(-oe_del) & 0xffffu) * 0x10001u;
    const u32 noe_ins2 = ((u32)(-oe_ins) & 0xffffu) * 0x10001u, ne_ins2 = ((u32)(-opt.e_ins) & 0xffffu) * 0x10001u;
    const u32 ne_ins_hi = ((u32)(-opt.e_ins) & 0xffffu) << 16;
    const unsigned n_jobs = *n_jobs_p;
    const i64 l_pac = P.ix.l_pac;
    const u32 *pacw32 = (const u32 *)P.ix.pac;
    u64 n_cells = 0, n_sw = 0;
    bool active = false, exhausted = false;
    int np = 0, tlen = 0, i = 0, seed = 0, fstep = 0, comp = 0;
    i64 fpos = 0;
    u32 pacw = 0, pacw_tag = 0xffffffffu, gmax2 = 0;
    for (;;) {
        if (!active && !exhausted) {                                                     // next job: clear the row, stage the query selectors
            const unsigned slot = atomicAdd(P.work_counter, 1u);
            if (slot >= n_jobs) exhausted = true;
            else {
                const SswJob jb = jobs[slot];
                const u8 *q = P.seq + P.off[jb.read] + jb.qb;
                np = (jb.qlen + 1) >> 1; tlen = jb.tlen; seed = jb.seed; i = 0; gmax2 = 0;
                for (int p = 0; p < np; ++p) {
                    const u32 c0 = q[2 * p], c1 = 2 * p + 1 < jb.qlen ? q[2 * p + 1] : 5u;   // code 5: a score-0 column past the query end
                    QS[p * SW2_THREADS] = (u16)((c0 * 0x11u + 0x80u) | (c1 * 0x11u + 0x80u) << 8);
                    HE[p * SW2_THREADS] = make_uint2(0u, 0u);
                }
                if (jb.rb >= l_pac) { fpos = (l_pac << 1) - 1 - jb.rb; fstep = -1; comp = 3; }
                else { fpos = jb.rb; fstep = 1; comp = 0; }
                pacw_tag = 0xffffffffu;
                n_cells += (u64)jb.qlen * (u64)jb.tlen; ++n_sw;
                active = tlen > 0 && np > 0;
                if (!active) score_out[seed] = 0;
            }
        }
        __syncwarp();
        if (active) {                                                                    // one target row
            const u32 tag = (u32)(fpos >> 4);
            if (tag != pacw_tag) { pacw = pacw32[fpos >> 4]; pacw_tag = tag; }
            const int f = (int)fpos;
            const int tb = (int)((pacw >> ((((f >> 2) & 3) << 3) + ((~f & 3) << 1))) & 3) ^ comp;
            const u32 srow_lo = s_rows[2 * tb], srow_hi = s_rows[2 * tb + 1];
            u32 hc = 0, fc = 0;                                                          // H(i-1, j-1) in the high half; F entering column j in the high half
            for (int p = 0; p < np; ++p) {
                const uint2 he = HE[p * SW2_THREADS];
                const u32 sc2 = sw2_prmt(srow_lo, srow_hi, (u32)QS[p * SW2_THREADS]);
                const u32 Hd2 = sw2_prmt(hc, he.x, 0x5432u);                             // (H(i-1,j-1), H(i-1,j))
                hc = he.x;
                const u32 M2 = __vadd2(Hd2, sc2);
                const u32 G2 = __vimax3_s16x2_relu(M2, he.y, M2);                        // max(0, diagonal, E)
                const u32 T2 = __viaddmax_s16x2_relu(G2, noe_ins2, 0u);                  // gap opened from this column
                const u32 A = sw2_prmt(fc, 0u, 0x3232u);
                const u32 F2 = __viaddmax_s16x2(A, ne_ins_hi, T2 << 16);                 // (F_j, F_j+1)
                fc = __viaddmax_s16x2(F2, ne_ins2, T2);                                  // high half: F_j+2
                const u32 H2 = __vmaxs2(G2, F2);
                const u32 E2 = __viaddmax_s16x2_relu(H2, noe_del2, __vadd2(he.y, ne_del2));
                gmax2 = __vmaxs2(gmax2, H2);
                HE[p * SW2_THREADS] = make_uint2(H2, E2);
            }
            fpos += fstep;
            if (++i == tlen) {
                const int lo = (int)(gmax2 & 0xffffu), hi = (int)(gmax2 >> 16);
                score_out[seed] = lo > hi ? lo : hi;
                active = false;
            }
        }
        if (__all_sync(B200_FULL_MASK, exhausted && !active)) break;
    }
    if (n_sw) { atomicAdd((unsigned long long *)&P.stats[0], (unsigned long long)n_sw); atomicAdd((unsigned long long *)&P.stats[1], (unsigned long long)n_cells); }
}

// ---- (3) bwamem.c:612-620: keep a seed when it was not re-scored or scores at least min_HSP_score
__global__ void __launch_bounds__(128) ssw_filter_kernel(SeedSwParams P, const i32 *score_out)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= P.n_long) return;
    const int r = P.long_reads[u];
    const int min_hsp = P.min_hsp[r];
    ChainRec *chains = P.chains + P.chn_off[r];
    const int n_chn = P.chn_n[r];
    for (int ci = 0; ci < n_chn; ++ci) {
        ChainRec *c = &chains[ci];
        Seed *seeds = P.cseeds + c->seed_off;
        const int n = c->n;
        int k = 0;
        for (int j = 0; j < n; ++j) {
            const int score = score_out[c->seed_off + j];
            if (score < 0 || score >= min_hsp) {
                Seed t = seeds[j];
                t.score = score < 0 ? t.len * P.opt.a : score;
                seeds[k++] = t;
            }
        }
        c->n = k;
    }
}

// scoring parameters for which the 16-bit halves cannot wrap (|values| stay far below 2^15 for windows < 200 bases)
bool seed_sw2_supported(const mem_opt_t *opt)
{
    if (opt->o_del < 0 || opt->e_del < 0 || opt->o_ins < 0 || opt->e_ins < 0) return false;
    if (opt->o_del + opt->e_del > 8000 || opt->o_ins + opt->e_ins > 8000) return false;
    return true;                                             // mat[] is int8: 200 * 127 < 2^15
}

int launch_seed_sw2(const SeedSwParams &P, SswJob *jobs, i64 n_seeds_total, i32 *score_out, unsigned *n_jobs, int n_sms, cudaStream_t st)
{
    const int np_cap = SW2_SHORT_LEN / 2;
    const size_t smem = (size_t)np_cap * SW2_THREADS * (8 + 2);
    B200_CHECK(cudaMemsetAsync(score_out, 0xff, (size_t)(n_seeds_total + 1) * 4, st));
    B200_CHECK(cudaMemsetAsync(n_jobs, 0, 4, st));
    B200_LAUNCH(ssw_jobs_kernel, (P.n_long + 127) / 128, 128, 0, st, P, jobs, n_jobs);
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute(ssw_dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / (smem + 1024)));
    const i64 lanes_wanted = (n_seeds_total + SW2_THREADS - 1) / SW2_THREADS;
    const int grid = (int)std::max<i64>(1, std::min<i64>(lanes_wanted, (i64)n_sms * per_sm));
    B200_LAUNCH(ssw_dp_kernel, grid, SW2_THREADS, smem, st, P, jobs, n_jobs, score_out, np_cap);
    B200_LAUNCH(ssw_filter_kernel, (P.n_long + 127) / 128, 128, 0, st, P, score_out);
    B200_CHECK(cudaGetLastError());
    return 0;
}
