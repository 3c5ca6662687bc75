// Roofline denominators measured on the device the library runs on (SURVEY.md section 8d asks for both):
//   kind 0: dependent random 32-byte-sector gathers over a table (the access pattern of FM-index seeding):
//           every lane walks its own pointer-chasing chain, `mlp` independent chains per lane
//   kind 1: DPX issue rate: chains of __viaddmax_s32_relu / __vimax3_s32 (the cell arithmetic of ksw_extend2)
#include "ctx.h"

__global__ void __launch_bounds__(128) gather_bench_kernel(const uint4 *tab, u64 n_sectors, int iters, int mlp, u32 *sink)
{
    const u64 gtid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    u64 x0 = gtid * 0x9e3779b97f4a7c15ull + 12345, x1 = ~x0 * 0xbf58476d1ce4e5b9ull, x2 = x0 ^ 0x94d049bb133111ebull, x3 = x1 + 77;
    u32 acc = 0;
    for (int it = 0; it < iters; ++it) {
        const uint4 *p0 = tab + (__umul64hi(x0, n_sectors) << 1);
        uint4 a0 = __ldg(p0), b0 = __ldg(p0 + 1), a1 = a0, b1 = b0, a2 = a0, b2 = b0, a3 = a0, b3 = b0;
        if (mlp > 1) { const uint4 *p = tab + (__umul64hi(x1, n_sectors) << 1); a1 = __ldg(p); b1 = __ldg(p + 1); }
        if (mlp > 2) { const uint4 *p = tab + (__umul64hi(x2, n_sectors) << 1); a2 = __ldg(p); b2 = __ldg(p + 1); }
        if (mlp > 3) { const uint4 *p = tab + (__umul64hi(x3, n_sectors) << 1); a3 = __ldg(p); b3 = __ldg(p + 1); }
        x0 = x0 * 6364136223846793005ull + (a0.x ^ b0.w) + 1442695040888963407ull;
        x1 = x1 * 6364136223846793005ull + (a1.y ^ b1.z) + 1;
        x2 = x2 * 6364136223846793005ull + (a2.z ^ b2.y) + 3;
        x3 = x3 * 6364136223846793005ull + (a3.w ^ b3.x) + 5;
        acc += a0.y + b1.x + a2.w + b3.z;
    }
    if (acc == 0x12345678u) sink[0] = (u32)(x0 + x1 + x2 + x3);
}

__global__ void fill_table_kernel(uint4 *tab, u64 n_quads)
{
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; i < n_quads; i += stride) {
        u64 h = i * 0x9e3779b97f4a7c15ull; h ^= h >> 29;
        tab[i] = make_uint4((u32)h, (u32)(h >> 32), (u32)(h * 3), (u32)(h >> 17));
    }
}

__global__ void __launch_bounds__(256) dpx_bench_kernel(int iters, int seed, int *sink)
{
    int a0 = threadIdx.x + seed, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, f0 = 1, f1 = 2, f2 = 3, f3 = 4;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            a0 = __viaddmax_s32_relu(a0, -1, f0 - 7); f0 = __vimax3_s32(f0, a0, it);
            a1 = __viaddmax_s32_relu(a1, -1, f1 - 7); f1 = __vimax3_s32(f1, a1, it);
            a2 = __viaddmax_s32_relu(a2, -1, f2 - 7); f2 = __vimax3_s32(f2, a2, it);
            a3 = __viaddmax_s32_relu(a3, -1, f3 - 7); f3 = __vimax3_s32(f3, a3, it);
        }
    }
    if ((a0 ^ a1 ^ a2 ^ a3 ^ f0 ^ f1 ^ f2 ^ f3) == 0x7fffffff) sink[0] = a0;
}

// out[0] = milliseconds, out[1] = units (sectors gathered / DPX instructions executed)
extern "C" int b200_microbench(b200_ctx_t *ctx, int kind, int64_t table_bytes, int blocks_per_sm, int iters, int mlp, double *out)
{
    if (!ctx) { b200_set_error("null context"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    cudaEvent_t e0, e1;
    B200_CHECK(cudaEventCreate(&e0)); B200_CHECK(cudaEventCreate(&e1));
    float ms = 0;
    if (ctx->d_pa.ensure(64)) return -1;
    if (kind == 0) {
        const u64 n_sectors = (u64)table_bytes / 32;
        if (n_sectors == 0 || ctx->d_lists.ensure((size_t)n_sectors * 32)) return -1;
        B200_LAUNCH(fill_table_kernel, ctx->n_sms * 8, 256, 0, ctx->st, ctx->d_lists.as<uint4>(), n_sectors * 2);
        const int grid = ctx->n_sms * blocks_per_sm;
        B200_LAUNCH(gather_bench_kernel, grid, 128, 0, ctx->st, (const uint4 *)ctx->d_lists.p, n_sectors, iters / 8 + 1, mlp, ctx->d_pa.as<u32>());
        B200_CHECK(cudaEventRecord(e0, ctx->st));
        B200_LAUNCH(gather_bench_kernel, grid, 128, 0, ctx->st, (const uint4 *)ctx->d_lists.p, n_sectors, iters, mlp, ctx->d_pa.as<u32>());
        B200_CHECK(cudaEventRecord(e1, ctx->st));
        B200_CHECK(cudaEventSynchronize(e1));
        B200_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        out[0] = ms; out[1] = (double)grid * 128 * (double)iters * (mlp < 1 ? 1 : (mlp > 4 ? 4 : mlp));
    } else {
        const int grid = ctx->n_sms * blocks_per_sm;
        B200_LAUNCH(dpx_bench_kernel, grid, 256, 0, ctx->st, iters / 8 + 1, 1, (int *)ctx->d_pa.p);
        B200_CHECK(cudaEventRecord(e0, ctx->st));
        B200_LAUNCH(dpx_bench_kernel, grid, 256, 0, ctx->st, iters, 2, (int *)ctx->d_pa.p);
        B200_CHECK(cudaEventRecord(e1, ctx->st));
        B200_CHECK(cudaEventSynchronize(e1));
        B200_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        out[0] = ms; out[1] = (double)grid * 256 * (double)iters * 64;     // 8 x 4 x 2 DPX instructions per iteration
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    B200_CHECK(cudaGetLastError());
    return 0;
}
