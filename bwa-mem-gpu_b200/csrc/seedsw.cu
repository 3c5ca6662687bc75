// mem_flt_chained_seeds / mem_seed_sw (bwamem.c:605-622, 578-603): for reads long enough that short random seed hits
// become likely, every seed shorter than 200 bp is re-scored by a local Smith-Waterman of the seed +-50 bp against
// the reference +-50 bp (ksw_align2, 16-bit scores) and dropped when the score stays below a*5.5*ln(l_query).
//
// One read per 16-lane group (its chains and seeds in order, the chain's seed array compacted in place by lane 0);
// only the score of ksw_align2 is used, so the KSW_XSTART reverse pass of the reference is not run.
#include "ctx.h"
#include "ksw_align.cuh"

#define SSW_SHORT_EXT 50        // MEM_SHORT_EXT, bwamem.c:570
#define SSW_SHORT_LEN 200       // MEM_SHORT_LEN, bwamem.c:571

struct SswRefWin { const DevIndex *ix; i64 rb; __device__ __forceinline__ int operator()(int i) const { return dev_ref_base(*ix, rb + i); } };

__global__ void __launch_bounds__(128) seed_sw_kernel(SeedSwParams P)
{
    B200_DYN_SMEM(ssw_raw);
    const int glane = threadIdx.x & 15, gib = threadIdx.x >> 4;
    const unsigned gmask = 0xffffu << (threadIdx.x & 16);
    KswaScratch S;
    S.sm = (u16 *)ssw_raw + (size_t)gib * 5 * P.slen_cap * 16;
    S.slen_cap = P.slen_cap; S.b = nullptr; S.b_cap = 0;
    const mem_opt_t &opt = P.opt;
    const i64 l_pac = P.ix.l_pac;
    int ovf = 0;
    u64 n_sw = 0, n_cells = 0;
    for (;;) {
        unsigned u = 0;
        if (glane == 0) u = atomicAdd(P.work_counter, 1u);
        u = __shfl_sync(gmask, u, 0, 16);
        if (u >= (unsigned)P.n_long) break;
        const int r = P.long_reads[u];
        const int min_hsp = P.min_hsp[r];
        const u8 *query = P.seq + P.off[r];
        const int l_query = (int)(P.off[r + 1] - P.off[r]);
        ChainRec *chains = P.chains + P.chn_off[r];
        const int n_chn = P.chn_n[r];
        for (int ci = 0; ci < n_chn; ++ci) {
            ChainRec *c = &chains[ci];
            Seed *seeds = P.cseeds + c->seed_off;
            const int n = c->n;
            int k = 0;
            for (int j = 0; j < n; ++j) {
                const Seed s = seeds[j];
                int score = -1;                                                          // mem_seed_sw
                if (s.len < SSW_SHORT_LEN) {
                    int qb = s.qbeg, qe = s.qbeg + s.len;
                    i64 rb = s.rbeg, re = s.rbeg + s.len;
                    const i64 mid = (rb + re) >> 1;
                    qb -= SSW_SHORT_EXT; qb = qb > 0 ? qb : 0;
                    qe += SSW_SHORT_EXT; qe = qe < l_query ? qe : l_query;
                    rb -= SSW_SHORT_EXT; rb = rb > 0 ? rb : 0;
                    re += SSW_SHORT_EXT; re = re < l_pac << 1 ? re : l_pac << 1;
                    if (rb < l_pac && l_pac < re) { if (mid < l_pac) re = l_pac; else rb = l_pac; }
                    if (!(qe - qb >= SSW_SHORT_LEN || re - rb >= SSW_SHORT_LEN)) {
                        int rev;                                                         // bns_fetch_seq: clamp to the contig of mid
                        const int rid = dev_pos2rid(P.ix, dev_depos(P.ix, mid, &rev));
                        i64 far_beg = P.ix.ann_off[rid], far_end = far_beg + P.ix.ann_len[rid];
                        if (rev) { const i64 t = far_beg; far_beg = (l_pac << 1) - far_end; far_end = (l_pac << 1) - t; }
                        rb = rb > far_beg ? rb : far_beg;
                        re = re < far_end ? re : far_end;
                        const KswAlnRes x = kswa_pass(2, qe - qb, KswaFwd{query + qb}, (int)(re - rb), SswRefWin{&P.ix, rb}, opt.mat,
                                                      opt.o_del, opt.e_del, opt.o_ins, opt.e_ins, KSWA_XSTART, S, gmask, glane, &ovf);
                        score = x.score;
                        ++n_sw; n_cells += (u64)(qe - qb) * (u64)(re - rb);
                    }
                }
                if (score < 0 || score >= min_hsp) {                                     // bwamem.c:614-618
                    if (glane == 0) { Seed t = s; t.score = score < 0 ? s.len * opt.a : score; seeds[k] = t; }
                    ++k;
                }
                __syncwarp(gmask);
            }
            if (glane == 0) c->n = k;
        }
    }
    if (glane == 0 && n_sw) { atomicAdd((unsigned long long *)&P.stats[0], (unsigned long long)n_sw); atomicAdd((unsigned long long *)&P.stats[1], (unsigned long long)n_cells); }
    if (ovf) atomicMax(P.overflow, 1);
}

int launch_seed_sw(const SeedSwParams &P, int n_sms, cudaStream_t st)
{
    const size_t smem = (size_t)8 * 5 * P.slen_cap * 16 * 2;
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(6, (200 * 1024) / (smem + 1024)));
    const int grid = std::max(1, std::min((P.n_long + 7) / 8, n_sms * per_sm));
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute(seed_sw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
    B200_LAUNCH(seed_sw_kernel, grid, 128, smem, st, P);
    B200_CHECK(cudaGetLastError());
    return 0;
}
