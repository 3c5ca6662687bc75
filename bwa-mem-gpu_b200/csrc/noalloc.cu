// b200_ksw_extend2_noalloc: ksw_extend2 (ksw.c:380-479) over an array of problems WITHOUT any allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of
// ksw_batch_kernel (extend.cu): the same warp-per-problem row-at-a-time DP (ksw_warp.cuh), but the warp-private H / E
// rows and the work counter live in ONE file-scope __device__ array of fixed size instead of the context's grow-only
// cudaMalloc buffers.  Every GPU that loads the library holds the array until the process exits.
//
// The array's name is only a host-side stand-in in host code: its device address is taken once per device, with that
// device current, by cudaGetSymbolAddress, and that pointer is what the kernel receives.
//
// All calls share the array: calls that overlap on the GPU (two streams, two host threads) race on the rows and on the
// work counter.  The array keeps the previous call's contents; the only part that must start at zero is the work
// counter, cleared by a cudaMemsetAsync on the caller's stream in front of each launch (a memset node under capture).
// The H / E rows need no clearing: ksw_extend_warp writes columns 0..qlen of both rows before it reads them.
#include "b200_internal.h"
#include "ksw_warp.cuh"
#include <algorithm>
#include <climits>
#include <cstring>
#include <mutex>

#define NA_WARPS   4                                     // warps per block, as ksw_batch_kernel
#define NA_GRID    592                                   // resident blocks the scratch is sized for (148 SMs x 4)
#define NA_EH_CAP  (B200_NOALLOC_MAX_QLEN + 2)           // ints per row
#define NA_CTL_BYTES 256                                 // the work counter, in a 256-byte slot of its own
#define NA_SCRATCH_BYTES (NA_CTL_BYTES + (size_t)NA_GRID * NA_WARPS * 2 * NA_EH_CAP * 4)

static __device__ __align__(256) unsigned char g_na_scratch[NA_SCRATCH_BYTES];   // 19 398 912 bytes (18.5 MiB) per GPU

struct NaParams {
    i64 n;
    const i32 *qlen; const i64 *qoff; const u8 *q;
    const i32 *tlen; const i64 *toff; const u8 *t;
    int8_t mat[25];
    int o_del, e_del, o_ins, e_ins; const i32 *w; int end_bonus, zdrop; const i32 *h0;
    i32 *out;
    unsigned char *scratch;                              // device address of g_na_scratch (cudaGetSymbolAddress)
};

struct NaBufAt { const u8 *p; __device__ __forceinline__ int operator()(int i) const { return p[i]; } };

__global__ void __launch_bounds__(NA_WARPS * 32) ksw_batch_static_kernel(NaParams P)
{
    __shared__ int8_t s_mat[32];
    const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const size_t gw = (size_t)blockIdx.x * NA_WARPS + wib;
    if (threadIdx.x < 25) s_mat[threadIdx.x] = P.mat[threadIdx.x];
    __syncthreads();
    unsigned *work_counter = (unsigned *)P.scratch;
    i32 *H = (i32 *)(P.scratch + NA_CTL_BYTES) + gw * 2 * (size_t)NA_EH_CAP, *E = H + NA_EH_CAP;
    for (;;) {
        unsigned r = 0;
        if (lane == 0) r = atomicAdd(work_counter, 1u);
        r = __shfl_sync(B200_FULL_MASK, r, 0);
        if ((i64)r >= P.n) break;
        const int ql = P.qlen[r], h0 = P.h0[r];
        i32 *o = P.out + 6 * (size_t)r;
        if (ql < 0 || ql > B200_NOALLOC_MAX_QLEN || h0 <= 0) {              // beyond the scratch / ksw.c:385's assert: marked, not computed
            if (lane < 6) o[lane] = INT32_MIN;
        } else {
            KswResult kr = ksw_extend_warp(ql, NaBufAt{P.q + P.qoff[r]}, P.tlen[r], NaBufAt{P.t + P.toff[r]}, s_mat,
                                           P.o_del, P.e_del, P.o_ins, P.e_ins, P.w[r], P.end_bonus, P.zdrop, h0, H, E, nullptr);
            if (lane == 0) { o[0] = kr.score; o[1] = kr.qle; o[2] = kr.tle; o[3] = kr.gtle; o[4] = kr.gscore; o[5] = kr.max_off; }
        }
        __syncwarp();
    }
}

// device address of the scratch array, looked up once per device
static std::mutex g_na_mu;
static unsigned char *g_na_ptr[64];

static int na_scratch_for(int device, unsigned char **out)
{
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (!g_na_ptr[device]) {
        void *p = nullptr;
        B200_CHECK(cudaGetSymbolAddress(&p, g_na_scratch));                 // needs `device` current: the caller has set it
        if (!p) { b200_set_error("cudaGetSymbolAddress returned a null address for the scratch array"); return -1; }
        g_na_ptr[device] = (unsigned char *)p;
    }
    *out = g_na_ptr[device];
    return 0;
}

// Set-up, once per device (callers under graph capture do this before the capture starts): makes `device` current and
// takes the device address of the scratch array.  b200_ksw_extend2_noalloc does the same on its first call per device.
extern "C" int b200_noalloc_setup(int device)
{
    if (device < 0 || device >= 64) { b200_set_error("noalloc_setup: device %d out of range", device); return -1; }
    B200_CHECK(cudaSetDevice(device));
    unsigned char *scratch = nullptr;
    return na_scratch_for(device, &scratch);
}

extern "C" int b200_ksw_extend2_noalloc(int device, int64_t n, int max_qlen, const int32_t *d_qlen, const int64_t *d_qoff,
                                        const uint8_t *d_q, const int32_t *d_tlen, const int64_t *d_toff, const uint8_t *d_t,
                                        const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins, const int32_t *d_w,
                                        int end_bonus, int zdrop, const int32_t *d_h0, int32_t *d_out, void *stream)
{
    if (device < 0 || device >= 64) { b200_set_error("ksw_extend2_noalloc: device %d out of range", device); return -1; }
    if (n < 0 || n > B200_NOALLOC_MAX_PROBLEMS) {
        b200_set_error("ksw_extend2_noalloc: %lld problems, the limit is %lld per call", (long long)n, (long long)B200_NOALLOC_MAX_PROBLEMS);
        return -2;
    }
    if (max_qlen < 0 || max_qlen > B200_NOALLOC_MAX_QLEN) {
        b200_set_error("ksw_extend2_noalloc: query of %d bases, the static scratch holds rows of at most %d (use b200_ksw_extend2_batch)",
                       max_qlen, B200_NOALLOC_MAX_QLEN);
        return -2;
    }
    if (e_del <= 0 || e_ins <= 0) { b200_set_error("ksw_extend2_noalloc: gap extension penalties must be positive"); return -1; }
    if (n == 0) return 0;
    if (!d_qlen || !d_qoff || !d_q || !d_tlen || !d_toff || !d_t || !mat || !d_w || !d_h0 || !d_out) {
        b200_set_error("ksw_extend2_noalloc: null argument"); return -1;
    }
    B200_CHECK(cudaSetDevice(device));
    unsigned char *scratch = nullptr;
    if (na_scratch_for(device, &scratch)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    NaParams P;
    P.n = n; P.qlen = d_qlen; P.qoff = d_qoff; P.q = d_q; P.tlen = d_tlen; P.toff = d_toff; P.t = d_t;
    memcpy(P.mat, mat, 25);
    P.o_del = o_del; P.e_del = e_del; P.o_ins = o_ins; P.e_ins = e_ins; P.w = d_w; P.end_bonus = end_bonus; P.zdrop = zdrop;
    P.h0 = d_h0; P.out = d_out; P.scratch = scratch;
    B200_CHECK(cudaMemsetAsync(scratch, 0, NA_CTL_BYTES, st));              // the work counter starts at zero; the rows need not
    const int grid = (int)std::min<i64>(NA_GRID, (n + NA_WARPS - 1) / NA_WARPS);
    ksw_batch_static_kernel<<<grid, NA_WARPS * 32, 0, st>>>(P);
    B200_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int64_t b200_noalloc_scratch_bytes(void) { return (int64_t)NA_SCRATCH_BYTES; }
