// Mate rescue: the block of mem_sam_pe that runs before pairing (bwamem_pair.c:291-302) around mem_matesw
// (bwamem_pair.c:137-206), for every pair of the resident batch.
//
// Work unit = one read c whose region list `ma` may grow; the anchors are the regions of its mate (read c^1) as they
// were BEFORE any rescue (the reference snapshots them into b[0], b[1] first), so the two units of a pair are
// independent.  Inside a unit everything is sequential by definition: an inserted hit changes which orientations the
// next anchor still has to try.  One unit is run by a group of 16 lanes: the control flow is uniform across the
// group, the list is edited by lane 0, the Smith-Waterman (ksw_align2, ksw_align.cuh) uses all lanes.  The target
// window is never materialised: bases are decoded from the packed reference on the fly (bns_fetch_seq only clamps).
#include "ctx.h"
#include "extend_common.cuh"
#include "ksw_align.cuh"

struct RescueParams {
    DevIndex ix; mem_opt_t opt; mem_pestat_t pes[4];
    const u8 *seq; const i64 *off; int n_reads;             // nt4 reads of the batch
    const mem_alnreg_t *in; const i64 *in_off; const i32 *in_n;
    mem_alnreg_t *out; const i64 *out_off; i32 *out_n;      // padded lists, capacity from rescue_cap_kernel
    int slen_cap; u64 *b; int b_cap;
    unsigned *work_counter; u64 *stats; int *overflow;
};

// mem_infer_dir (bwamem_pair.c:49-56)
__device__ __forceinline__ int infer_dir(i64 l_pac, i64 b1, i64 b2, i64 *dist)
{
    const int r1 = b1 >= l_pac, r2 = b2 >= l_pac;
    const i64 p2 = r1 == r2 ? b2 : (l_pac << 1) - 1 - b2;
    *dist = p2 > b1 ? p2 - b1 : b1 - p2;
    return (r1 == r2 ? 0 : 1) ^ (p2 > b1 ? 0 : 3);
}

// number of anchors mem_sam_pe takes from a list (bwamem_pair.c:294-299)
__device__ __forceinline__ int n_anchors(const mem_alnreg_t *a, int n, const mem_opt_t &opt)
{
    int k = 0;
    for (int j = 0; j < n; ++j) if (a[j].score >= a[0].score - opt.pen_unpaired) ++k;
    return k < opt.max_matesw ? k : opt.max_matesw;
}

// list capacity a unit can need: every (anchor, usable orientation) adds at most one region
__global__ void rescue_cap_kernel(RescueParams P, i32 *cap)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P.n_reads) return;
    int n_or = 0;
    for (int r = 0; r < 4; ++r) n_or += P.pes[r].failed ? 0 : 1;
    const int m = c ^ 1;
    int extra = 0;
    if (m < P.n_reads) extra = n_or * n_anchors(P.in + P.in_off[m], P.in_n[m], P.opt);
    cap[c] = P.in_n[c] + extra;
}

struct MateFwd { const u8 *p; __device__ __forceinline__ int operator()(int i) const { return p[i]; } };
struct MateRev {                                                     // reverse complement, bwamem_pair.c:159-160
    const u8 *p; int last;
    __device__ __forceinline__ int operator()(int i) const { const int c = p[last - i]; return c < 4 ? 3 - c : 4; }
};
struct RefWin { const DevIndex *ix; i64 rb; __device__ __forceinline__ int operator()(int i) const { return dev_ref_base(*ix, rb + i); } };

__global__ void __launch_bounds__(128) rescue_kernel(RescueParams P)
{
    B200_DYN_SMEM(rs_raw);
    const int glane = threadIdx.x & 15, gib = threadIdx.x >> 4;
    const unsigned gmask = 0xffffu << (threadIdx.x & 16);
    const size_t gg = (size_t)blockIdx.x * 8 + gib;
    KswaScratch S;
    S.sm = (u16 *)rs_raw + (size_t)gib * 5 * P.slen_cap * 16;
    S.slen_cap = P.slen_cap;
    S.b = P.b + gg * (size_t)P.b_cap; S.b_cap = P.b_cap;
    const mem_opt_t &opt = P.opt;
    const i64 l_pac = P.ix.l_pac;
    int ovf = 0;
    u64 n_sw = 0, n_cells = 0;
    for (;;) {
        unsigned cu = 0;
        if (glane == 0) cu = atomicAdd(P.work_counter, 1u);
        cu = __shfl_sync(gmask, cu, 0, 16);
        if (cu >= (unsigned)P.n_reads) break;
        const int c = (int)cu, m = c ^ 1;
        mem_alnreg_t *ma = P.out + P.out_off[c];
        const int ma_cap = (int)(P.out_off[c + 1] - P.out_off[c]);
        int n_ma = P.in_n[c];
        {   // ma starts as the read's own list
            const u64 *src = (const u64 *)(P.in + P.in_off[c]); u64 *dst = (u64 *)ma;
            for (int k = glane; k < n_ma * 11; k += 16) dst[k] = src[k];
        }
        __syncwarp(gmask);
        const int n_a = m < P.n_reads ? P.in_n[m] : 0;
        const mem_alnreg_t *A = P.in + (m < P.n_reads ? P.in_off[m] : 0);
        const u8 *ms = P.seq + P.off[c];
        const int l_ms = (int)(P.off[c + 1] - P.off[c]);
        const int thr = n_a ? A[0].score - opt.pen_unpaired : 0;
        for (int j = 0, taken = 0; j < n_a && taken < opt.max_matesw; ++j) {
            if (A[j].score < thr) continue;
            ++taken;
            const mem_alnreg_t a = A[j];
            // ---- mem_matesw(opt, bns, pac, pes, &a, l_ms, ms, ma)
            int skip = 0, n = 0;
            for (int r = 0; r < 4; ++r) if (P.pes[r].failed) skip |= 1 << r;
            for (int i = 0; i < n_ma; ++i) {
                i64 dist;
                const int r = infer_dir(l_pac, a.rb, ma[i].rb, &dist);
                if (dist >= P.pes[r].low && dist <= P.pes[r].high) skip |= 1 << r;
            }
            if (skip == 15) continue;
            for (int r = 0; r < 4; ++r) {
                if (skip >> r & 1) continue;
                const int is_rev = (r >> 1) != (r & 1), is_larger = !(r >> 1);
                i64 rb, re;
                if (!is_rev) {
                    rb = is_larger ? a.rb + P.pes[r].low : a.rb - P.pes[r].high;
                    re = (is_larger ? a.rb + P.pes[r].high : a.rb - P.pes[r].low) + l_ms;
                } else {
                    rb = (is_larger ? a.rb + P.pes[r].low : a.rb - P.pes[r].high) - l_ms;
                    re = is_larger ? a.rb + P.pes[r].high : a.rb - P.pes[r].low;
                }
                if (rb < 0) rb = 0;
                if (re > l_pac << 1) re = l_pac << 1;
                int rid = -1;
                if (rb < re) {                                       // bns_fetch_seq (bntseq.c:426-451): clamp to the contig of mid
                    int rev;
                    rid = dev_pos2rid(P.ix, dev_depos(P.ix, (rb + re) >> 1, &rev));
                    i64 far_beg = P.ix.ann_off[rid], far_end = far_beg + P.ix.ann_len[rid];
                    if (rev) { const i64 t = far_beg; far_beg = (l_pac << 1) - far_end; far_end = (l_pac << 1) - t; }
                    rb = rb > far_beg ? rb : far_beg;
                    re = re < far_end ? re : far_end;
                }
                if (rb < re && a.rid == rid && re - rb >= opt.min_seed_len) {
                    const int xtra = KSWA_XSUBO | KSWA_XSTART | (l_ms * opt.a < 250 ? KSWA_XBYTE : 0) | (opt.min_seed_len * opt.a);
                    const int tlen = (int)(re - rb);
                    KswAlnRes aln;
                    if (is_rev) aln = kswa_align2_f(l_ms, MateRev{ms, l_ms - 1}, tlen, RefWin{&P.ix, rb}, opt.mat, opt.o_del, opt.e_del,
                                                    opt.o_ins, opt.e_ins, xtra, S, gmask, glane, &ovf);
                    else aln = kswa_align2_f(l_ms, MateFwd{ms}, tlen, RefWin{&P.ix, rb}, opt.mat, opt.o_del, opt.e_del,
                                             opt.o_ins, opt.e_ins, xtra, S, gmask, glane, &ovf);
                    ++n_sw; n_cells += (u64)l_ms * tlen;
                    if (aln.score >= opt.min_seed_len && aln.qb >= 0) {
                        if (n_ma >= ma_cap) ovf = 1;
                        else if (glane == 0) {
                            mem_alnreg_t b;
                            memset(&b, 0, sizeof(b));
                            b.rid = a.rid; b.is_alt = a.is_alt;
                            b.qb = is_rev ? l_ms - (aln.qe + 1) : aln.qb;
                            b.qe = is_rev ? l_ms - aln.qb : aln.qe + 1;
                            b.rb = is_rev ? (l_pac << 1) - (rb + aln.te + 1) : rb + aln.tb;
                            b.re = is_rev ? (l_pac << 1) - (rb + aln.tb) : rb + aln.te + 1;
                            b.score = aln.score; b.csub = aln.score2; b.secondary = -1;
                            b.seedcov = (int)((b.re - b.rb < b.qe - b.qb ? b.re - b.rb : b.qe - b.qb) >> 1);
                            int i;
                            for (i = 0; i < n_ma; ++i) if (ma[i].score < b.score) break;      // keep ma sorted by score
                            for (int k = n_ma; k > i; --k) ma[k] = ma[k - 1];
                            ma[i] = b;
                        }
                        if (n_ma < ma_cap) ++n_ma;
                    }
                    ++n;
                }
                if (n) {
                    if (glane == 0) n_ma = sort_dedup_patch(P.ix, opt, nullptr, n_ma, ma, nullptr, nullptr);
                    n_ma = __shfl_sync(gmask, n_ma, 0, 16);
                }
                __syncwarp(gmask);
            }
        }
        if (glane == 0) P.out_n[c] = n_ma;
    }
    if (glane == 0 && n_sw) { atomicAdd((unsigned long long *)&P.stats[0], (unsigned long long)n_sw); atomicAdd((unsigned long long *)&P.stats[1], (unsigned long long)n_cells); }
    if (ovf) atomicMax(P.overflow, 1);
}

// Runs the rescue over the batch resident in the context (reads: d_seq/d_off, regions: d_final/d_final_n) and replaces
// the resident region lists by the rescued ones.  stats[0] = ksw_align2 calls, stats[1] = their DP cells (l_ms x tlen).
extern "C" int b200_batch_rescue(b200_ctx_t *ctx, const mem_opt_t *opt, const mem_pestat_t *pes, int64_t *stats, float *ms)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    const int n = ctx->n_reads;
    if (stats) stats[0] = stats[1] = 0;
    if (ms) *ms = 0;
    if (n <= 0) return 0;
    cudaStream_t st = ctx->st;
    int max_win = 0;
    for (int r = 0; r < 4; ++r) if (!pes[r].failed && pes[r].high - pes[r].low > max_win) max_win = pes[r].high - pes[r].low;
    const int slen_cap = (ctx->max_len + 7) / 8 + 1;
    const int b_cap = max_win + ctx->max_len + 2;
    const size_t smem = (size_t)8 * 5 * slen_cap * 16 * 2;
    if (smem > 200 * 1024) { b200_set_error("mate rescue: reads of %d bases do not fit the shared-memory profile", ctx->max_len); return -2; }
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / (smem + 1024)));
    const int grid = (int)std::min<i64>(((i64)n + 7) / 8, (i64)ctx->n_sms * per_sm);
    if (ctx->d_pa.ensure(((size_t)n + 1) * 8) || ctx->d_pb.ensure(((size_t)n + 1) * 4) || ctx->d_pc.ensure(((size_t)n + 1) * 8) ||
        ctx->d_pd.ensure(((size_t)n + 1) * 4) || ctx->d_scan_tmp.ensure(((size_t)n / 1024 + 2) * 8) ||
        ctx->d_eh.ensure((size_t)grid * 8 * b_cap * 8) || ctx->d_ctl.ensure(CTL_SLOTS * 8)) return -1;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ms) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, st); }
    RescueParams P;
    memset(&P, 0, sizeof(P));
    P.ix = ctx->ix; P.opt = *opt; memcpy(P.pes, pes, 4 * sizeof(mem_pestat_t));
    P.seq = ctx->d_seq.as<u8>(); P.off = ctx->d_off.as<i64>(); P.n_reads = n;
    P.in = ctx->d_final.as<mem_alnreg_t>(); P.in_off = ctx->d_pa.as<i64>(); P.in_n = ctx->d_final_n.as<i32>();
    P.out_off = ctx->d_pc.as<i64>(); P.out_n = ctx->d_pd.as<i32>();
    P.slen_cap = slen_cap; P.b = ctx->d_eh.as<u64>(); P.b_cap = b_cap;
    B200_CHECK(cudaMemsetAsync(ctx->d_ctl.p, 0, CTL_SLOTS * 8, st));
    P.work_counter = (unsigned *)(ctx->d_ctl.as<u64>() + CTL_WORK); P.stats = ctx->d_ctl.as<u64>() + CTL_STATS;
    P.overflow = (int *)(ctx->d_ctl.as<u64>() + CTL_OVF);
    if (launch_exclusive_scan(P.in_n, ctx->d_pa.as<i64>(), n, ctx->d_scan_tmp.as<i64>(), st)) return -1;
    B200_LAUNCH(rescue_cap_kernel, (n + 255) / 256, 256, 0, st, P, ctx->d_pb.as<i32>());
    if (launch_exclusive_scan(ctx->d_pb.as<i32>(), ctx->d_pc.as<i64>(), n, ctx->d_scan_tmp.as<i64>(), st)) return -1;
    i64 tot_cap = 0;
    B200_CHECK(cudaMemcpyAsync(&tot_cap, ctx->d_pc.as<i64>() + n, 8, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    if (ctx->d_stage.ensure((size_t)(tot_cap + 1) * sizeof(mem_alnreg_t))) return -1;
    P.out = ctx->d_stage.as<mem_alnreg_t>();
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute(rescue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
    B200_LAUNCH(rescue_kernel, grid, 128, smem, st, P);
    B200_CHECK(cudaGetLastError());
    // compact the padded lists back into the resident result, in read order
    if (launch_exclusive_scan(P.out_n, ctx->d_pa.as<i64>(), n, ctx->d_scan_tmp.as<i64>(), st)) return -1;
    i64 tot = 0; u64 ctl[CTL_SLOTS];
    B200_CHECK(cudaMemcpyAsync(&tot, ctx->d_pa.as<i64>() + n, 8, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaMemcpyAsync(ctl, ctx->d_ctl.p, CTL_SLOTS * 8, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    if (ctl[CTL_OVF]) { b200_set_error("mate rescue: scratch overflow"); return -2; }
    if (ctx->d_final.ensure((size_t)(tot + 1) * sizeof(mem_alnreg_t))) return -1;
    if (launch_gather_regs(P.out, P.out_off, P.out_n, ctx->d_pa.as<i64>(), n, ctx->d_final.as<mem_alnreg_t>(), st)) return -1;
    B200_CHECK(cudaMemcpyAsync(ctx->d_final_n.p, P.out_n, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    if (ms) { cudaEventRecord(e1, st); }
    B200_CHECK(cudaStreamSynchronize(st));
    if (ms) { cudaEventElapsedTime(ms, e0, e1); cudaEventDestroy(e0); cudaEventDestroy(e1); }
    ctx->total_regs = tot;
    if (stats) { stats[0] = (int64_t)ctl[CTL_STATS]; stats[1] = (int64_t)ctl[CTL_STATS + 1]; }
    return 0;
}
