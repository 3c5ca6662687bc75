// Stage 3: seed extension (mem_chain2aln, bwamem.c:639-793; cal_max_gap :628-635; bns_fetch_seq clamp,
// bntseq.c:426-451) followed by mem_sort_dedup_patch (bwamem.c:444-496) and the is_alt marking of
// mem_align1_core (bwamem.c:1091-1095).
//
// One warp owns one read and walks its chains and seeds in the reference's order: whether a seed is
// extended depends on every region already produced for the read (bwamem.c:678-713), so that control
// flow is replayed as-is (warp-uniform scalar code), while each extension itself runs on all 32 lanes
// (ksw_warp.cuh).  Reference bases are decoded straight from the 2-bit pac (no window is materialised:
// a DP row needs one target base), 32 rows per fetch.  Warps are persistent and pull reads from an
// atomic work counter, so long and short reads balance dynamically.
#include "b200_internal.h"
#include "ksort.cuh"
#include "ksw_warp.cuh"
#include "extend_common.cuh"

#define EXT_WARPS 4
#define EXT_SM_EH 320          // H/E rows kept in shared memory when qlen + 1 <= this

struct QueryRev { const u8 *q; int qbeg; __device__ __forceinline__ int operator()(int j) const { return q[qbeg - 1 - j]; } };
struct QueryFwd { const u8 *q; int qe;   __device__ __forceinline__ int operator()(int j) const { return q[qe + j]; } };
struct RefRev { const DevIndex *ix; i64 p0; __device__ __forceinline__ int operator()(int i) const { return dev_ref_base(*ix, p0 - i); } };
struct RefFwd { const DevIndex *ix; i64 p0; __device__ __forceinline__ int operator()(int i) const { return dev_ref_base(*ix, p0 + i); } };

__global__ void __launch_bounds__(EXT_WARPS * 32) extend_kernel(ExtParams P)
{
    __shared__ int8_t s_mat[32];
    __shared__ i32 s_eh[EXT_WARPS][2 * EXT_SM_EH];
    const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const size_t gw = (size_t)blockIdx.x * EXT_WARPS + wib;
    if (threadIdx.x < 25) s_mat[threadIdx.x] = P.opt.mat[threadIdx.x];
    __syncthreads();
    const mem_opt_t &opt = P.opt;
    const DevIndex &ix = P.ix;
    const i64 l_pac = ix.l_pac;
    mem_alnreg_t *av = P.regs + gw * (size_t)P.cap_regs;
    i32 *gH = P.eh + gw * 2 * (size_t)P.eh_cap, *gE = gH + P.eh_cap;
    u64 *srt = P.srt + gw * (size_t)P.srt_cap;
    u64 n_ksw = 0, n_cells = 0;

    for (;;) {
        unsigned r = 0;
        if (lane == 0) r = atomicAdd(P.work_counter, 1u);
        r = __shfl_sync(B200_FULL_MASK, r, 0);
        if (r >= (unsigned)P.n_reads) break;
        const u8 *query = P.seq + P.off[r];
        const int l_query = (int)(P.off[r + 1] - P.off[r]);
        const int nchn = P.chn_n[r];
        const ChainRec *chains = P.chains + P.chn_off[r];
        int nav = 0, bad = 0;
        i32 *H = (l_query + 1 <= EXT_SM_EH) ? &s_eh[wib][0] : gH;
        i32 *E = (l_query + 1 <= EXT_SM_EH) ? &s_eh[wib][EXT_SM_EH] : gE;
        if (l_query + 1 > P.eh_cap && l_query + 1 > EXT_SM_EH) bad = 1;

        for (int ci = 0; ci < nchn && !bad; ++ci) {
            const ChainRec c = chains[ci];
            if (c.n == 0) continue;
            if (c.n > P.srt_cap) { bad = 1; break; }
            const Seed *seeds = P.cseeds + c.seed_off;
            // max possible span (bwamem.c:649-664)
            i64 rmax0 = l_pac << 1, rmax1 = 0;
            for (int i = 0; i < c.n; ++i) {
                const Seed t = seeds[i];
                i64 b = t.rbeg - (t.qbeg + cal_max_gap(opt, t.qbeg));
                i64 e = t.rbeg + t.len + ((l_query - t.qbeg - t.len) + cal_max_gap(opt, l_query - t.qbeg - t.len));
                rmax0 = rmax0 < b ? rmax0 : b;
                rmax1 = rmax1 > e ? rmax1 : e;
            }
            rmax0 = rmax0 > 0 ? rmax0 : 0;
            rmax1 = rmax1 < l_pac << 1 ? rmax1 : l_pac << 1;
            if (rmax0 < l_pac && l_pac < rmax1) {
                if (seeds[0].rbeg < l_pac) rmax1 = l_pac; else rmax0 = l_pac;
            }
            {   // bns_fetch_seq clamp to the contig of seeds[0].rbeg (bntseq.c:432-443)
                int is_rev;
                int rid = dev_pos2rid(ix, dev_depos(ix, seeds[0].rbeg, &is_rev));
                i64 far_beg = ix.ann_off[rid], far_end = far_beg + ix.ann_len[rid];
                if (is_rev) { i64 tmp = far_beg; far_beg = (l_pac << 1) - far_end; far_end = (l_pac << 1) - tmp; }
                rmax0 = rmax0 > far_beg ? rmax0 : far_beg;
                rmax1 = rmax1 < far_end ? rmax1 : far_end;
            }
            // seeds by score (bwamem.c:669-672); keys are unique so any exact sort gives the same order
            __syncwarp();
            for (int i = lane; i < c.n; i += 32) srt[i] = (u64)(u32)seeds[i].score << 32 | (u32)i;
            __syncwarp();
            if (lane == 0) ks_introsort(srt, c.n, U64Less());
            __syncwarp();

            for (int k = c.n - 1; k >= 0; --k) {
                const Seed s = seeds[(u32)srt[k]];
                int i;
                for (i = 0; i < nav; ++i) {                                   // bwamem.c:678-694
                    const mem_alnreg_t *p = &av[i];
                    i64 rd; int qd, w, max_gap;
                    if (s.rbeg < p->rb || s.rbeg + s.len > p->re || s.qbeg < p->qb || s.qbeg + s.len > p->qe) continue;
                    if (s.len - p->seedlen0 > .1 * l_query) continue;
                    qd = s.qbeg - p->qb; rd = s.rbeg - p->rb;
                    max_gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
                    w = max_gap < p->w ? max_gap : p->w;
                    if (qd - rd < w && rd - qd < w) break;
                    qd = p->qe - (s.qbeg + s.len); rd = p->re - (s.rbeg + s.len);
                    max_gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
                    w = max_gap < p->w ? max_gap : p->w;
                    if (qd - rd < w && rd - qd < w) break;
                }
                if (i < nav) {                                                 // bwamem.c:695-713
                    for (i = k + 1; i < c.n; ++i) {
                        if (srt[i] == 0) continue;
                        const Seed t = seeds[(u32)srt[i]];
                        if (t.len < s.len * .95) continue;
                        if (s.qbeg <= t.qbeg && s.qbeg + s.len - t.qbeg >= s.len >> 2 && t.qbeg - s.qbeg != t.rbeg - s.rbeg) break;
                        if (t.qbeg <= s.qbeg && t.qbeg + t.len - s.qbeg >= s.len >> 2 && s.qbeg - t.qbeg != s.rbeg - t.rbeg) break;
                    }
                    if (i == c.n) {
                        __syncwarp();
                        if (lane == 0) srt[k] = 0;
                        __syncwarp();
                        continue;
                    }
                }
                if (nav >= P.cap_regs) { bad = 1; break; }

                mem_alnreg_t a;
                memset(&a, 0, sizeof(a));
                int aw0 = opt.w, aw1 = opt.w;
                a.w = opt.w; a.score = a.truesc = -1; a.rid = c.rid;
                if (s.qbeg) {                                                  // left extension, bwamem.c:722-751
                    KswResult kr; kr.score = -1; kr.qle = kr.tle = kr.gtle = kr.gscore = kr.max_off = 0;
                    const i64 tmp = s.rbeg - rmax0;
                    for (int t = 0; t < 2; ++t) {                              // MAX_BAND_TRY
                        int prev = a.score;
                        aw0 = opt.w << t;
                        kr = ksw_extend_warp(s.qbeg, QueryRev{query, s.qbeg}, (int)tmp, RefRev{&ix, s.rbeg - 1}, s_mat,
                                             opt.o_del, opt.e_del, opt.o_ins, opt.e_ins, aw0, opt.pen_clip5, opt.zdrop,
                                             s.len * opt.a, H, E, &n_cells);
                        ++n_ksw;
                        a.score = kr.score;
                        if (a.score == prev || kr.max_off < (aw0 >> 1) + (aw0 >> 2)) break;
                    }
                    if (kr.gscore <= 0 || kr.gscore <= a.score - opt.pen_clip5) {
                        a.qb = s.qbeg - kr.qle; a.rb = s.rbeg - kr.tle; a.truesc = a.score;
                    } else {
                        a.qb = 0; a.rb = s.rbeg - kr.gtle; a.truesc = kr.gscore;
                    }
                } else { a.score = a.truesc = s.len * opt.a; a.qb = 0; a.rb = s.rbeg; }

                if (s.qbeg + s.len != l_query) {                               // right extension, bwamem.c:753-778
                    KswResult kr; kr.score = -1; kr.qle = kr.tle = kr.gtle = kr.gscore = kr.max_off = 0;
                    const int sc0 = a.score, qe = s.qbeg + s.len;
                    const i64 re = s.rbeg + s.len - rmax0;
                    for (int t = 0; t < 2; ++t) {
                        int prev = a.score;
                        aw1 = opt.w << t;
                        kr = ksw_extend_warp(l_query - qe, QueryFwd{query, qe}, (int)(rmax1 - rmax0 - re), RefFwd{&ix, rmax0 + re}, s_mat,
                                             opt.o_del, opt.e_del, opt.o_ins, opt.e_ins, aw1, opt.pen_clip3, opt.zdrop,
                                             sc0, H, E, &n_cells);
                        ++n_ksw;
                        a.score = kr.score;
                        if (a.score == prev || kr.max_off < (aw1 >> 1) + (aw1 >> 2)) break;
                    }
                    if (kr.gscore <= 0 || kr.gscore <= a.score - opt.pen_clip3) {
                        a.qe = qe + kr.qle; a.re = rmax0 + re + kr.tle; a.truesc += a.score - sc0;
                    } else {
                        a.qe = l_query; a.re = rmax0 + re + kr.gtle; a.truesc += kr.gscore - sc0;
                    }
                } else { a.qe = l_query; a.re = s.rbeg + s.len; }

                a.seedcov = 0;                                                 // bwamem.c:781-790
                for (i = 0; i < c.n; ++i) {
                    const Seed t = seeds[i];
                    if (t.qbeg >= a.qb && t.qbeg + t.len <= a.qe && t.rbeg >= a.rb && t.rbeg + t.len <= a.re) a.seedcov += t.len;
                }
                a.w = aw0 > aw1 ? aw0 : aw1;
                a.seedlen0 = s.len;
                a.frac_rep = c.frac_rep;
                __syncwarp();
                if (lane == 0) av[nav] = a;
                ++nav;
                __syncwarp();
            }
        }

        // mem_sort_dedup_patch + is_alt (bwamem.c:1083-1095)
        i64 obase = 0;
        if (!bad) {
            __syncwarp();
            if (lane == 0) {
                nav = sort_dedup_patch(ix, opt, query, nav, av, H, E);
                for (int i = 0; i < nav; ++i)
                    if (av[i].rid >= 0 && ix.ann_alt[av[i].rid]) av[i].is_alt = 1;
                obase = (i64)atomicAdd(P.out_used, (unsigned long long)nav);
            }
            nav = __shfl_sync(B200_FULL_MASK, nav, 0);
            obase = __shfl_sync(B200_FULL_MASK, obase, 0);
            __syncwarp();
            if (obase + nav <= P.out_cap) {
                const u32 *src = (const u32 *)av;
                u32 *dst = (u32 *)(P.out + obase);
                const int nw = nav * (int)(sizeof(mem_alnreg_t) / 4);
                for (int i = lane; i < nw; i += 32) dst[i] = src[i];
            } else if (lane == 0) atomicMax(P.overflow, 2);
            if (lane == 0) { P.reg_off[r] = obase; P.reg_n[r] = nav; }
        } else {
            if (lane == 0) { P.reg_off[r] = 0; P.reg_n[r] = 0; atomicMax(P.overflow, 1); }
        }
        __syncwarp();
    }
    if (lane == 0 && n_ksw) {
        atomicAdd(&P.stats[CNT_KSW], (unsigned long long)n_ksw);
        atomicAdd(&P.stats[CNT_CELLS], (unsigned long long)n_cells);
    }
}

int launch_extend(const ExtParams &p, int grid, int block, cudaStream_t st)
{
    (void)block;
    B200_LAUNCH(extend_kernel, grid, EXT_WARPS * 32, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------- standalone ksw_extend2 batch (one warp per problem)
struct BufAt { const u8 *p; __device__ __forceinline__ int operator()(int i) const { return p[i]; } };

__global__ void __launch_bounds__(EXT_WARPS * 32) ksw_batch_kernel(KswBatchParams P)
{
    __shared__ int8_t s_mat[32];
    const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const size_t gw = (size_t)blockIdx.x * EXT_WARPS + wib;
    if (threadIdx.x < 25) s_mat[threadIdx.x] = P.mat[threadIdx.x];
    __syncthreads();
    i32 *H = P.eh + gw * 2 * (size_t)P.eh_cap, *E = H + P.eh_cap;
    for (;;) {
        unsigned r = 0;
        if (lane == 0) r = atomicAdd(P.work_counter, 1u);
        r = __shfl_sync(B200_FULL_MASK, r, 0);
        if ((i64)r >= P.n) break;
        KswResult kr = ksw_extend_warp(P.qlen[r], BufAt{P.q + P.qoff[r]}, P.tlen[r], BufAt{P.t + P.toff[r]}, s_mat,
                                       P.o_del, P.e_del, P.o_ins, P.e_ins, P.w[r], P.end_bonus, P.zdrop, P.h0[r], H, E, nullptr);
        if (lane == 0) {
            i32 *o = P.out + 6 * (size_t)r;
            o[0] = kr.score; o[1] = kr.qle; o[2] = kr.tle; o[3] = kr.gtle; o[4] = kr.gscore; o[5] = kr.max_off;
        }
        __syncwarp();
    }
}

int launch_ksw_batch(const KswBatchParams &p, int grid, int block, cudaStream_t st)
{
    (void)block;
    B200_LAUNCH(ksw_batch_kernel, grid, EXT_WARPS * 32, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}

// regions of read r: src[src_off[r] .. +n[r])  ->  dst[dst_off[r] ..): read-ordered compaction
__global__ void gather_regs_kernel(const mem_alnreg_t *src, const i64 *src_off, const i32 *n, const i64 *dst_off,
                                   int n_reads, mem_alnreg_t *dst)
{
    const int per = (int)(sizeof(mem_alnreg_t) / 4);
    int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (r >= n_reads) return;
    const u32 *s = (const u32 *)(src + src_off[r]);
    u32 *d = (u32 *)(dst + dst_off[r]);
    const int nw = n[r] * per;
    for (int i = lane; i < nw; i += 32) d[i] = s[i];
}

int launch_gather_regs(const mem_alnreg_t *src, const i64 *src_off, const i32 *n, const i64 *dst_off,
                       int n_reads, mem_alnreg_t *dst, cudaStream_t st)
{
    if (n_reads == 0) return 0;
    B200_LAUNCH(gather_regs_kernel, (n_reads * 32 + 255) / 256, 256, 0, st, src, src_off, n, dst_off, n_reads, dst);
    B200_CHECK(cudaGetLastError());
    return 0;
}
