// Batched ksw_align2 (ksw.c:343-365): one problem per 16-lane group, see ksw_align.cuh.
#include "ctx.h"
#include "ksw_align.cuh"

struct KswAlignParams {
    i64 n;
    const i32 *qlen; const i64 *qoff; const u8 *q;
    const i32 *tlen; const i64 *toff; const u8 *t;
    int8_t mat[25];
    int o_del, e_del, o_ins, e_ins; const i32 *xtra;
    i32 *out;                                   // 7 per problem: score, te, qe, score2, te2, tb, qb (kswr_t, ksw.h:14-19)
    int slen_cap; u64 *b; int b_cap;
    unsigned *work_counter; int *overflow;
};

__global__ void __launch_bounds__(128) ksw_align_batch_kernel(KswAlignParams P)
{
    B200_DYN_SMEM(ka_raw);
    __shared__ int8_t s_mat[32];
    if (threadIdx.x < 25) s_mat[threadIdx.x] = P.mat[threadIdx.x];
    __syncthreads();
    const int glane = threadIdx.x & 15, gib = threadIdx.x >> 4;
    const unsigned gmask = 0xffffu << (threadIdx.x & 16);
    const size_t gg = (size_t)blockIdx.x * 8 + gib;
    KswaScratch S;
    S.sm = (u16 *)ka_raw + (size_t)gib * 5 * P.slen_cap * 16;
    S.slen_cap = P.slen_cap;
    S.b = P.b + gg * (size_t)P.b_cap; S.b_cap = P.b_cap;
    int ovf = 0;
    for (;;) {
        unsigned r = 0;
        if (glane == 0) r = atomicAdd(P.work_counter, 1u);
        r = __shfl_sync(gmask, r, 0, 16);
        if ((i64)r >= P.n) break;
        const KswAlnRes a = kswa_align2(P.qlen[r], P.q + P.qoff[r], P.tlen[r], P.t + P.toff[r], s_mat, P.o_del, P.e_del,
                                        P.o_ins, P.e_ins, P.xtra[r], S, gmask, glane, &ovf);
        if (glane == 0) {
            i32 *o = P.out + 7 * (size_t)r;
            o[0] = a.score; o[1] = a.te; o[2] = a.qe; o[3] = a.score2; o[4] = a.te2; o[5] = a.tb; o[6] = a.qb;
        }
    }
    if (ovf) atomicMax(P.overflow, 1);
}

extern "C" int b200_ksw_align2_batch(b200_ctx_t *ctx, int64_t n, const int32_t *qlen, const int64_t *qoff, const uint8_t *q,
                                     const int32_t *tlen, const int64_t *toff, const uint8_t *t, const int8_t *mat,
                                     int o_del, int e_del, int o_ins, int e_ins, const int32_t *xtra, int32_t *out)
{
    if (!ctx) { b200_set_error("null context"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    if (n <= 0) return 0;
    i64 qbytes = 0, tbytes = 0; int maxq = 0, maxt = 0;
    for (i64 i = 0; i < n; ++i) {
        if (qoff[i] + qlen[i] > qbytes) qbytes = qoff[i] + qlen[i];
        if (toff[i] + tlen[i] > tbytes) tbytes = toff[i] + tlen[i];
        if (qlen[i] > maxq) maxq = qlen[i];
        if (tlen[i] > maxt) maxt = tlen[i];
    }
    const int slen_cap = (maxq + 7) / 8 + 1;
    const size_t smem = (size_t)8 * 5 * slen_cap * 16 * 2;
    if (smem > 200 * 1024) { b200_set_error("ksw_align2: query of %d bases does not fit the shared-memory profile", maxq); return -2; }
    const int grid = (int)std::min<i64>((n + 7) / 8, (i64)ctx->n_sms * std::max<size_t>(1, (200 * 1024) / (smem + 1)));
    const int b_cap = maxt + 1;
    if (ctx->d_pa.ensure(n * 4) || ctx->d_pb.ensure(n * 8) || ctx->d_pc.ensure(qbytes + 1) || ctx->d_pd.ensure(n * 4) ||
        ctx->d_pe.ensure(n * 8) || ctx->d_pf.ensure(tbytes + 1) || ctx->d_pg.ensure(n * 4) || ctx->d_ph.ensure(n * 28) ||
        ctx->d_eh.ensure((size_t)grid * 8 * b_cap * 8) || ctx->d_ctl.ensure(CTL_SLOTS * 8)) return -1;
    cudaStream_t st = ctx->st;
    B200_CHECK(cudaMemcpyAsync(ctx->d_pa.p, qlen, n * 4, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pb.p, qoff, n * 8, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pc.p, q, qbytes, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pd.p, tlen, n * 4, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pe.p, toff, n * 8, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pf.p, t, tbytes, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pg.p, xtra, n * 4, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemsetAsync(ctx->d_ctl.p, 0, CTL_SLOTS * 8, st));
    KswAlignParams P;
    P.n = n; P.qlen = ctx->d_pa.as<i32>(); P.qoff = ctx->d_pb.as<i64>(); P.q = ctx->d_pc.as<u8>();
    P.tlen = ctx->d_pd.as<i32>(); P.toff = ctx->d_pe.as<i64>(); P.t = ctx->d_pf.as<u8>();
    memcpy(P.mat, mat, 25);
    P.o_del = o_del; P.e_del = e_del; P.o_ins = o_ins; P.e_ins = e_ins; P.xtra = ctx->d_pg.as<i32>();
    P.out = ctx->d_ph.as<i32>(); P.slen_cap = slen_cap; P.b = ctx->d_eh.as<u64>(); P.b_cap = b_cap;
    P.work_counter = (unsigned *)(ctx->d_ctl.as<u64>() + CTL_WORK); P.overflow = (int *)(ctx->d_ctl.as<u64>() + CTL_OVF);
#ifndef B200_EMU
    B200_CHECK(cudaFuncSetAttribute(ksw_align_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
    B200_LAUNCH(ksw_align_batch_kernel, grid, 128, smem, st, P);
    B200_CHECK(cudaGetLastError());
    u64 ovf = 0;
    B200_CHECK(cudaMemcpyAsync(&ovf, ctx->d_ctl.as<u64>() + CTL_OVF, 8, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaMemcpyAsync(out, ctx->d_ph.p, n * 28, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    if (ovf) { b200_set_error("ksw_align2: scratch overflow"); return -2; }
    return 0;
}
