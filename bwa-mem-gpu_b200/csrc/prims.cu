// Batched FM-index primitives (one problem per thread) and the device-side SA expansion.
#include "b200_internal.h"

__global__ void occ_batch_kernel(DevIndex ix, i64 n, const u64 *k, const u64 *l, u64 *ck, u64 *cl)
{
    i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u64 a[4], b[4];
    if (l) {                       // bwt_2occ4 (bwt.c:189)
        dev_2occ4(ix, k[t], l[t], a, b, nullptr);
        for (int c = 0; c < 4; ++c) { ck[4 * t + c] = a[c]; cl[4 * t + c] = b[c]; }
    } else {                       // bwt_occ4 (bwt.c:169)
        dev_occ4(ix, k[t], a);
        for (int c = 0; c < 4; ++c) ck[4 * t + c] = a[c];
    }
}

__global__ void extend_batch_kernel(DevIndex ix, i64 n, const u64 *ik, int is_back, u64 *ok)
{
    i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u64 o0[4], o1[4], o2[4];
    dev_extend(ix, ik[3 * t], ik[3 * t + 1], ik[3 * t + 2], is_back, o0, o1, o2, nullptr);
    for (int c = 0; c < 4; ++c) { ok[12 * t + 3 * c] = o0[c]; ok[12 * t + 3 * c + 1] = o1[c]; ok[12 * t + 3 * c + 2] = o2[c]; }
}

__global__ void sa_batch_kernel(DevIndex ix, i64 n, const u64 *k, u64 *out, int use_full)
{
    i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    out[t] = use_full ? dev_sa(ix, k[t], nullptr) : dev_sa_walk(ix, k[t], nullptr);
}

// Full suffix array from the sampled one: row k walks LF to the next sampled row (bwt_sa, bwt.c:86-96).
// One-off at index load; afterwards every SA lookup of the hot path is a single gather.
__global__ void build_full_sa_kernel(DevIndex ix, u32 *out32, u64 *out64)
{
    u64 k = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; k <= ix.seq_len; k += stride) {
        u64 v = dev_sa_walk(ix, k, nullptr);
        if (out32) out32[k] = (u32)v; else out64[k] = v;
    }
}

// occ32 block b (64 rows) from the reference-layout blocks (bwtindex.c:149-172): counts before the block
// (relative to its 2^32-row superblock) and the two bit planes of its 64 symbols (occ32.cuh).
__global__ void build_occ32_kernel(DevIndex ix, uint4 *out, u64 n_blocks)
{
    u64 b = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; b < n_blocks; b += stride) {
        const uint4 *p = ix.bwt + ((b >> 1) << 2);
        const uint4 c01 = p[0], c23 = p[1], w0 = p[2], w1 = p[3];
        u64 cnt[4] = { u64_of(c01.x, c01.y), u64_of(c01.z, c01.w), u64_of(c23.x, c23.y), u64_of(c23.z, c23.w) };
        const u32 wd[8] = { w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w };
        if (b & 1)
            for (int j = 0; j < 64; ++j) ++cnt[(wd[j >> 4] >> ((~j & 15) << 1)) & 3];
        const int h = (int)(b & 1) * 4;
        u64 lo = 0, hi = 0;
        for (int j = 0; j < 64; ++j) {
            const u32 s = (wd[h + (j >> 4)] >> ((~j & 15) << 1)) & 3;
            lo |= (u64)(s & 1) << j; hi |= (u64)(s >> 1) << j;
        }
        const int sb = (int)(b >> 26);
        out[2 * b] = make_uint4((u32)(cnt[0] - ix.sb_cnt[sb][0]), (u32)(cnt[1] - ix.sb_cnt[sb][1]),
                                (u32)(cnt[2] - ix.sb_cnt[sb][2]), (u32)(cnt[3] - ix.sb_cnt[sb][3]));
        out[2 * b + 1] = make_uint4((u32)lo, (u32)(lo >> 32), (u32)hi, (u32)(hi >> 32));
    }
}

int launch_build_occ32(const DevIndex &ix, uint4 *out, u64 n_blocks, cudaStream_t st)
{
    int grid = (int)((n_blocks + 255) / 256 < 148 * 32 ? (n_blocks + 255) / 256 : 148 * 32);
    B200_LAUNCH(build_occ32_kernel, grid, 256, 0, st, ix, out, n_blocks);
    B200_CHECK(cudaGetLastError());
    return 0;
}

int launch_occ_batch(const DevIndex &ix, i64 n, const u64 *k, const u64 *l, u64 *ck, u64 *cl, cudaStream_t st)
{
    if (n <= 0) return 0;
    B200_LAUNCH(occ_batch_kernel, (int)((n + 127) / 128), 128, 0, st, ix, n, k, l, ck, cl);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_extend_batch(const DevIndex &ix, i64 n, const u64 *ik, int is_back, u64 *ok, cudaStream_t st)
{
    if (n <= 0) return 0;
    B200_LAUNCH(extend_batch_kernel, (int)((n + 127) / 128), 128, 0, st, ix, n, ik, is_back, ok);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_sa_batch(const DevIndex &ix, i64 n, const u64 *k, u64 *out, int use_full, cudaStream_t st)
{
    if (n <= 0) return 0;
    B200_LAUNCH(sa_batch_kernel, (int)((n + 127) / 128), 128, 0, st, ix, n, k, out, use_full);
    B200_CHECK(cudaGetLastError());
    return 0;
}
// Full SA as 5 bytes per row: four rows = five 32-bit words, one thread per group of four (no two threads share a word).
template <class T>
__global__ void pack_sa40_kernel(const T *in, u32 *out, u64 n)
{
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 g = (u64)blockIdx.x * blockDim.x + threadIdx.x; g * 4 < n; g += stride) {
        u64 v[4];
        for (int t = 0; t < 4; ++t) { const u64 r = g * 4 + t; v[t] = r < n ? (u64)in[r] & 0xffffffffffull : 0; }
        u32 *o = out + g * 5;
        o[0] = (u32)v[0];
        o[1] = (u32)(v[0] >> 32) | (u32)(v[1] << 8);
        o[2] = (u32)(v[1] >> 24) | (u32)(v[2] << 16);
        o[3] = (u32)(v[2] >> 16) | (u32)(v[3] << 24);
        o[4] = (u32)(v[3] >> 8);
    }
}

size_t sa40_bytes(u64 n_rows) { return (size_t)((n_rows + 3) / 4) * 20 + 16; }      // whole groups + one spare word for the two-word read of the last row

int launch_pack_sa40(const void *in, int in_is64, u32 *out, u64 n_rows, cudaStream_t st)
{
    const u64 groups = (n_rows + 3) / 4;
    const int grid = (int)std::min<u64>((groups + 255) / 256, (u64)148 * 64);
    if (in_is64) B200_LAUNCH((pack_sa40_kernel<u64>), grid, 256, 0, st, (const u64 *)in, out, n_rows);
    else B200_LAUNCH((pack_sa40_kernel<u32>), grid, 256, 0, st, (const u32 *)in, out, n_rows);
    B200_CHECK(cudaGetLastError());
    return 0;
}

int launch_build_full_sa(const DevIndex &ix, u32 *out32, u64 *out64, cudaStream_t st)
{
    u64 n = ix.seq_len + 1;
    int grid = (int)((n + 255) / 256 < 148 * 32 ? (n + 255) / 256 : 148 * 32);
    B200_LAUNCH(build_full_sa_kernel, grid, 256, 0, st, ix, out32, out64);
    B200_CHECK(cudaGetLastError());
    return 0;
}
