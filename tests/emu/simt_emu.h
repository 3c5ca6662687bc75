// TEST HARNESS ONLY -- a minimal SIMT emulator so that the product's CUDA kernels can be exercised,
// unmodified, inside the GPU-less build container.  The product library (libbwamem_b200.so) is built
// by nvcc and never sees this header; the emulated build has a different name
// (tests/emu/_build/libbwamem_b200_emu.so), is loaded only by `-m "not gpu"` tests, and is never
// selected by the package at run time (there is no CPU fallback in the product).
//
// Model: a CUDA block = one OS thread running `blockDim.x` fibers (ucontext) round-robin.  Warp
// collectives (__shfl*_sync, __ballot_sync, ...) and __syncthreads() are rendezvous points at which
// a fiber yields until every participant has arrived.  Blocks are distributed over host threads.
// Only 1-D grids/blocks are supported (all our kernels are 1-D).
#pragma once
#ifndef B200_EMU
#error "simt_emu.h is only for the emulated test build"
#endif

#include <ucontext.h>
#include <sys/mman.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <atomic>
#include <functional>
#include <thread>
#include <vector>
#include <algorithm>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __noinline__ __attribute__((noinline))
#define __launch_bounds__(...)
#define __shared__ static thread_local
#define __constant__ static
#define __align__(n) __attribute__((aligned(n)))

namespace emu {

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};

struct Warp {
    uint64_t in[32], out[32];
    uint32_t arrived;
    uint32_t gen[32];
};

struct Fiber {
    ucontext_t ctx;
    char *stack;
    unsigned tid;
    bool done;
};

struct Block {
    ucontext_t sched;
    std::vector<Fiber> fibers;
    std::vector<Warp> warps;
    unsigned nthreads, bid, grid;
    unsigned bar_count, bar_gen;
    unsigned char *dyn_smem;
    std::function<void()> *body;
    unsigned long idle_spins;
};

extern thread_local Block *g_block;
extern thread_local Fiber *g_fiber;

static const size_t kStackBytes = 256 << 10;

inline void yield_() { swapcontext(&g_fiber->ctx, &g_block->sched); }

inline void spin_check(const char *what)
{
    if (++g_block->idle_spins > 50000000ul) {
        fprintf(stderr, "[simt_emu] deadlock suspected in %s (block %u thread %u)\n", what, g_block->bid, g_fiber->tid);
        abort();
    }
}

// all lanes named in `mask` deposit v; returns pointer to the 32 deposited values
inline const uint64_t *warp_gather(unsigned mask, uint64_t v)
{
    Fiber *f = g_fiber;
    Warp &w = g_block->warps[f->tid >> 5];
    unsigned lane = f->tid & 31;
    if (!(mask >> lane & 1)) { fprintf(stderr, "[simt_emu] lane %u not in mask %08x\n", lane, mask); abort(); }
    w.in[lane] = v;
    w.arrived |= 1u << lane;
    if ((w.arrived & mask) == mask) {
        for (unsigned l = 0; l < 32; ++l)
            if (mask >> l & 1) { w.out[l] = w.in[l]; ++w.gen[l]; }
        w.arrived &= ~mask;
        g_block->idle_spins = 0;
    } else {
        uint32_t g = w.gen[lane];
        while (w.gen[lane] == g) { spin_check("warp collective"); yield_(); }
    }
    return w.out;
}

inline void syncthreads_()
{
    Block *b = g_block;
    unsigned g = b->bar_gen;
    if (++b->bar_count == b->nthreads) { b->bar_count = 0; ++b->bar_gen; b->idle_spins = 0; }
    else while (b->bar_gen == g) { spin_check("__syncthreads"); yield_(); }
}

void launch(dim3 grid, dim3 block, size_t smem_bytes, std::function<void()> body);

struct Idx { unsigned x, y, z; };
inline Idx tid3() { return Idx{g_fiber->tid, 0, 0}; }
inline Idx bid3() { return Idx{g_block->bid, 0, 0}; }
inline Idx bdim3() { return Idx{g_block->nthreads, 1, 1}; }
inline Idx gdim3() { return Idx{g_block->grid, 1, 1}; }
inline unsigned char *dyn_smem() { return g_block->dyn_smem; }

template <class T> inline uint64_t to_bits(T v) { uint64_t b = 0; static_assert(sizeof(T) <= 8, ""); memcpy(&b, &v, sizeof(T)); return b; }
template <class T> inline T from_bits(uint64_t b) { T v; memcpy(&v, &b, sizeof(T)); return v; }

}  // namespace emu

using emu::dim3;
struct uint4 { unsigned x, y, z, w; };
struct uint2 { unsigned x, y; };
struct int4 { int x, y, z, w; };
struct int2 { int x, y; };
struct ulonglong2 { unsigned long long x, y; };
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
inline int2 make_int2(int x, int y) { return int2{x, y}; }
inline ulonglong2 make_ulonglong2(unsigned long long x, unsigned long long y) { return ulonglong2{x, y}; }
#define threadIdx (emu::tid3())
#define blockIdx (emu::bid3())
#define blockDim (emu::bdim3())
#define gridDim (emu::gdim3())
#define warpSize 32

// ------------------------------------------------------------------ synchronisation / collectives
inline void __syncthreads() { emu::syncthreads_(); }
inline void __syncwarp(unsigned mask = 0xffffffffu) { emu::warp_gather(mask, 0); }
inline void __threadfence() { std::atomic_thread_fence(std::memory_order_seq_cst); }
inline void __threadfence_block() {}

template <class T> inline T __shfl_sync(unsigned mask, T v, int src, int width = 32)
{
    unsigned lane = emu::g_fiber->tid & 31;
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    unsigned base = lane & ~(unsigned)(width - 1);
    return emu::from_bits<T>(o[base + ((unsigned)src & (unsigned)(width - 1))]);
}
template <class T> inline T __shfl_up_sync(unsigned mask, T v, unsigned delta, int width = 32)
{
    unsigned lane = emu::g_fiber->tid & 31;
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    unsigned base = lane & ~(unsigned)(width - 1);
    int src = (int)lane - (int)delta;
    return src < (int)base ? v : emu::from_bits<T>(o[src]);
}
template <class T> inline T __shfl_down_sync(unsigned mask, T v, unsigned delta, int width = 32)
{
    unsigned lane = emu::g_fiber->tid & 31;
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    unsigned base = lane & ~(unsigned)(width - 1);
    unsigned src = lane + delta;
    return src >= base + width ? v : emu::from_bits<T>(o[src]);
}
template <class T> inline T __shfl_xor_sync(unsigned mask, T v, int lm, int width = 32)
{
    unsigned lane = emu::g_fiber->tid & 31;
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    unsigned src = lane ^ (unsigned)lm;
    (void)width;
    return emu::from_bits<T>(o[src]);
}
inline unsigned __ballot_sync(unsigned mask, int pred)
{
    const uint64_t *o = emu::warp_gather(mask, pred ? 1 : 0);
    unsigned r = 0;
    for (unsigned l = 0; l < 32; ++l) if ((mask >> l & 1) && o[l]) r |= 1u << l;
    return r;
}
inline int __any_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) != 0; }
inline int __all_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) == mask; }
inline int __reduce_max_sync(unsigned mask, int v)
{
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    int r = v;
    for (unsigned l = 0; l < 32; ++l) if (mask >> l & 1) r = std::max(r, emu::from_bits<int>(o[l]));
    return r;
}
inline int __reduce_min_sync(unsigned mask, int v)
{
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    int r = v;
    for (unsigned l = 0; l < 32; ++l) if (mask >> l & 1) r = std::min(r, emu::from_bits<int>(o[l]));
    return r;
}
inline unsigned __reduce_max_sync(unsigned mask, unsigned v)
{
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    unsigned r = v;
    for (unsigned l = 0; l < 32; ++l) if (mask >> l & 1) r = std::max(r, emu::from_bits<unsigned>(o[l]));
    return r;
}
inline int __reduce_add_sync(unsigned mask, int v)
{
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    int r = 0;
    for (unsigned l = 0; l < 32; ++l) if (mask >> l & 1) r += emu::from_bits<int>(o[l]);
    return r;
}
inline unsigned __reduce_or_sync(unsigned mask, unsigned v)
{
    const uint64_t *o = emu::warp_gather(mask, emu::to_bits(v));
    unsigned r = 0;
    for (unsigned l = 0; l < 32; ++l) if (mask >> l & 1) r |= emu::from_bits<unsigned>(o[l]);
    return r;
}

// ------------------------------------------------------------------ integer intrinsics
inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b) { return (unsigned long long)(((unsigned __int128)a * b) >> 64); }
inline int __popc(unsigned x) { return __builtin_popcount(x); }
inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned sh) { return (unsigned)((((unsigned long long)hi << 32) | lo) >> (sh & 31)); }
inline int __clzll(long long x) { return x ? __builtin_clzll((unsigned long long)x) : 64; }
inline int __ffs(int x) { return __builtin_ffs(x); }
inline unsigned __brev(unsigned x)
{
    x = (x >> 1 & 0x55555555u) | (x & 0x55555555u) << 1;
    x = (x >> 2 & 0x33333333u) | (x & 0x33333333u) << 2;
    x = (x >> 4 & 0x0f0f0f0fu) | (x & 0x0f0f0f0fu) << 4;
    return __builtin_bswap32(x);
}
template <class T> inline T __ldg(const T *p) { return *p; }
template <class T> inline T __ldcs(const T *p) { return *p; }
template <class T> inline void __stcs(T *p, T v) { *p = v; }
// PRMT: byte select from {y:x}; selector nibble bit 3 = replicate the sign bit of the selected byte
inline unsigned __byte_perm(unsigned x, unsigned y, unsigned s)
{
    const uint64_t v = (uint64_t)y << 32 | x;
    unsigned r = 0;
    for (int k = 0; k < 4; ++k) {
        const unsigned sel = (s >> (4 * k)) & 0xf;
        unsigned b = (unsigned)(v >> (8 * (sel & 7))) & 0xff;
        if (sel & 8) b = (b & 0x80) ? 0xff : 0x00;
        r |= b << (8 * k);
    }
    return r;
}
// DPX (emulated exactly; on sm_100a these are single instructions)
inline int __vimax3_s32(int a, int b, int c) { return std::max(a, std::max(b, c)); }
inline int __vimin3_s32(int a, int b, int c) { return std::min(a, std::min(b, c)); }
inline int __vimax3_s32_relu(int a, int b, int c) { return std::max(0, std::max(a, std::max(b, c))); }
inline int __vimax_s32_relu(int a, int b) { return std::max(0, std::max(a, b)); }
inline int __viaddmax_s32(int a, int b, int c) { return std::max(a + b, c); }
inline int __viaddmax_s32_relu(int a, int b, int c) { return std::max(0, std::max(a + b, c)); }
// 16x2 DPX / video SIMD (per-half, wrap-around add like the hardware)
namespace emu {
inline int16_t lo16(unsigned x) { return (int16_t)(x & 0xffffu); }
inline int16_t hi16(unsigned x) { return (int16_t)(x >> 16); }
inline unsigned pk16(int lo, int hi) { return ((unsigned)lo & 0xffffu) | ((unsigned)hi << 16); }
inline int16_t wrap16(int v) { return (int16_t)(uint16_t)(unsigned)v; }
}
inline unsigned __vadd2(unsigned a, unsigned b) { return emu::pk16(emu::lo16(a) + emu::lo16(b), emu::hi16(a) + emu::hi16(b)); }
inline unsigned __viaddmax_s16x2(unsigned a, unsigned b, unsigned c)
{ return emu::pk16(std::max<int>(emu::wrap16(emu::lo16(a) + emu::lo16(b)), emu::lo16(c)), std::max<int>(emu::wrap16(emu::hi16(a) + emu::hi16(b)), emu::hi16(c))); }
inline unsigned __viaddmin_s16x2(unsigned a, unsigned b, unsigned c)
{ return emu::pk16(std::min<int>(emu::wrap16(emu::lo16(a) + emu::lo16(b)), emu::lo16(c)), std::min<int>(emu::wrap16(emu::hi16(a) + emu::hi16(b)), emu::hi16(c))); }
inline unsigned __viaddmax_s16x2_relu(unsigned a, unsigned b, unsigned c)
{ return emu::pk16(std::max<int>(0, std::max<int>(emu::wrap16(emu::lo16(a) + emu::lo16(b)), emu::lo16(c))), std::max<int>(0, std::max<int>(emu::wrap16(emu::hi16(a) + emu::hi16(b)), emu::hi16(c)))); }
inline unsigned __vimax3_s16x2(unsigned a, unsigned b, unsigned c)
{ return emu::pk16(std::max<int>(emu::lo16(a), std::max<int>(emu::lo16(b), emu::lo16(c))), std::max<int>(emu::hi16(a), std::max<int>(emu::hi16(b), emu::hi16(c)))); }
inline unsigned __vimax3_s16x2_relu(unsigned a, unsigned b, unsigned c)
{ return emu::pk16(std::max<int>(0, std::max<int>(emu::lo16(a), std::max<int>(emu::lo16(b), emu::lo16(c)))), std::max<int>(0, std::max<int>(emu::hi16(a), std::max<int>(emu::hi16(b), emu::hi16(c))))); }
inline unsigned __vmaxs2(unsigned a, unsigned b)
{ return emu::pk16(std::max<int>(emu::lo16(a), emu::lo16(b)), std::max<int>(emu::hi16(a), emu::hi16(b))); }
inline unsigned __vmaxu2(unsigned a, unsigned b)
{ return emu::pk16(std::max<unsigned>(a & 0xffffu, b & 0xffffu), std::max<unsigned>(a >> 16, b >> 16)); }
// prmt.b32 default mode: selector nibble bit 3 = replicate the sign bit of the selected byte
inline unsigned emu_prmt(unsigned a, unsigned b, unsigned s)
{
    const uint64_t v = (uint64_t)b << 32 | a;
    unsigned r = 0;
    for (int k = 0; k < 4; ++k) {
        const unsigned n = (s >> (4 * k)) & 0xf;
        unsigned byte = (unsigned)(v >> (8 * (n & 7))) & 0xff;
        if (n & 8) byte = (byte & 0x80) ? 0xff : 0x00;
        r |= byte << (8 * k);
    }
    return r;
}
inline int __viaddmin_s32(int a, int b, int c) { return std::min(a + b, c); }

// ------------------------------------------------------------------ atomics
inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
inline unsigned atomicAdd(unsigned *p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
inline unsigned long long atomicAdd(unsigned long long *p, unsigned long long v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
inline int atomicMax(int *p, int v)
{
    int o = __atomic_load_n(p, __ATOMIC_RELAXED);
    while (o < v && !__atomic_compare_exchange_n(p, &o, v, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
    return o;
}
inline unsigned atomicOr(unsigned *p, unsigned v) { return __atomic_fetch_or(p, v, __ATOMIC_RELAXED); }
inline int atomicOr(int *p, int v) { return __atomic_fetch_or(p, v, __ATOMIC_RELAXED); }
inline int atomicExch(int *p, int v) { return __atomic_exchange_n(p, v, __ATOMIC_RELAXED); }
inline int atomicCAS(int *p, int cmp, int v) { __atomic_compare_exchange_n(p, &cmp, v, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED); return cmp; }

// ------------------------------------------------------------------ runtime API subset
typedef int cudaError_t;
typedef void *cudaStream_t;
struct emuEvent { double t; };
typedef emuEvent *cudaEvent_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3, cudaMemcpyDefault = 4 };
enum { cudaHostAllocDefault = 0, cudaHostAllocPortable = 1, cudaStreamNonBlocking = 1, cudaEventDefault = 0, cudaEventDisableTiming = 2 };

inline double emu_now_ms() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
inline cudaError_t cudaMalloc(void **p, size_t n) { *p = n ? malloc(n) : nullptr; return (*p || !n) ? 0 : 2; }
template <class T> inline cudaError_t cudaMalloc(T **p, size_t n) { return cudaMalloc((void **)p, n); }
inline cudaError_t cudaFree(void *p) { free(p); return 0; }
inline cudaError_t cudaMallocHost(void **p, size_t n) { return cudaMalloc(p, n); }
template <class T> inline cudaError_t cudaMallocHost(T **p, size_t n) { return cudaMalloc((void **)p, n); }
inline cudaError_t cudaHostAlloc(void **p, size_t n, unsigned) { return cudaMalloc(p, n); }
inline cudaError_t cudaFreeHost(void *p) { free(p); return 0; }
inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { if (n) memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t = 0) { if (n) memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemset(void *d, int v, size_t n) { if (n) memset(d, v, n); return 0; }
inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t = 0) { if (n) memset(d, v, n); return 0; }
inline cudaError_t cudaStreamCreate(cudaStream_t *s) { *s = nullptr; return 0; }
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { *s = nullptr; return 0; }
inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return 0; }
inline cudaError_t cudaDeviceSynchronize() { return 0; }
inline cudaError_t cudaGetLastError() { return 0; }
inline cudaError_t cudaPeekAtLastError() { return 0; }
inline const char *cudaGetErrorString(cudaError_t e) { return e ? "emulated error" : "no error"; }
inline cudaError_t cudaSetDevice(int) { return 0; }
inline cudaError_t cudaGetDevice(int *d) { *d = 0; return 0; }
inline cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return 0; }
inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new emuEvent{0}; return 0; }
inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { *e = new emuEvent{0}; return 0; }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return 0; }
inline cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t = 0) { e->t = emu_now_ms(); return 0; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) { *ms = (float)(b->t - a->t); return 0; }
inline cudaError_t cudaMemGetInfo(size_t *f, size_t *t) { *f = *t = (size_t)16 << 30; return 0; }
