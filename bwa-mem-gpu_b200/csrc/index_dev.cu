// Device-side index construction: the arrays `bwa index` writes (bwa_idx_build, bwtindex.c:253-323) computed on
// the GPU from the 2-bit forward reference, so that a human-scale synthetic genome (3.1 Gbp, text of 6.2 G
// symbols) is indexed in seconds on the box that aligns against it instead of in hours on its host cores.
//
// The BWT and the suffix array are mathematically determined by the text (forward + reverse complement,
// bwt_pac2bwt bwtindex.c:63-126; sentinel smaller than every base), so any correct suffix sorter gives the
// byte-identical arrays; tests/test_index.py compares every array with the host SA-IS builder / the reference.
//
// Method (prefix-key sort; exact for any text, fast when 32-mers are mostly unique, as on random or real genomes):
//   1. text = fwd + revcomp as 2-bit codes, 16 per word, most significant first
//   2. histogram of the first 6 bases of every suffix -> contiguous key ranges ("groups") of at most `group_cap`
//      suffixes; per group: compact (key = first 32 bases as one u64, position) pairs, radix-sort them (CUB), append
//      to the full suffix array
//   3. suffixes whose 32-base keys collide are collected and ordered on the host by direct comparison (a few dozen
//      on a random genome; bounded by `tie_cap`, beyond which the call fails and the host builder has to be used)
//   4. BWT = text[SA-1] with the sentinel row removed (`primary`), Occ checkpoints every 128 rows interleaved into
//      64-byte blocks (bwt_bwtupdate_core, bwtindex.c:149-172), SA sampled every 32 rows with sa[0] = -1
//      (bwt_cal_sa, bwt.c:62-84); the full SA stays on the device for the hot path's single-gather SA lookups.
// Nothing here runs on the alignment hot path.
#include "ctx.h"
#include <string.h>
#include <algorithm>
#include <vector>
#ifndef B200_EMU
#include <cub/device/device_radix_sort.cuh>
#endif

namespace {

struct TieRec { u64 j; u64 pos; };      // j = (index inside the sorted group) << 1 | continues-previous-run

__device__ __forceinline__ u32 pac_base(const u8 *pac, u64 i) { return (pac[i >> 2] >> ((~i & 3) << 1)) & 3u; }
__device__ __forceinline__ u32 tx_base(const u32 *tx, u64 i) { return (tx[i >> 4] >> ((~i & 15) << 1)) & 3u; }

// first 32 bases of suffix i as one big-endian 2-bit word (zero padded past the end of the text)
__device__ __forceinline__ u64 key_at(const u32 *tx, u64 i)
{
    const u64 w = i >> 4;
    const int sh = (int)(i & 15) << 1;
    const u64 hi = (u64)tx[w] << 32 | tx[w + 1];
    return sh ? (hi << sh) | ((u64)tx[w + 2] >> (32 - sh)) : hi;
}

__global__ void text2_kernel(const u8 *pac, u64 l_pac, u32 *tx, u64 n_words)
{
    const u64 n = 2 * l_pac;
    u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; w < n_words; w += stride) {
        u32 v = 0;
        for (int j = 0; j < 16; ++j) {
            const u64 i = (w << 4) + j;
            u32 c = 0;
            if (i < l_pac) c = pac_base(pac, i);
            else if (i < n) c = 3u - pac_base(pac, n - 1 - i);
            v |= c << ((15 - j) << 1);
        }
        tx[w] = v;
    }
}

__global__ void hist_kernel(const u32 *tx, u64 n, unsigned long long *hist /* 4096 */)
{
    __shared__ unsigned sh[4096];
    for (int k = threadIdx.x; k < 4096; k += blockDim.x) sh[k] = 0;
    __syncthreads();
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; i < n; i += stride) atomicAdd(&sh[(unsigned)(key_at(tx, i) >> 52)], 1u);
    __syncthreads();
    for (int k = threadIdx.x; k < 4096; k += blockDim.x)
        if (sh[k]) atomicAdd(&hist[k], (unsigned long long)sh[k]);
}

// (key, position) pairs of every suffix whose 6-base bin lies in [bin_lo, bin_hi); order inside the group is arbitrary
template <class PT>
__global__ void compact_kernel(const u32 *tx, u64 n, unsigned bin_lo, unsigned bin_hi, u64 *keys, PT *vals, unsigned long long *counter)
{
    __shared__ unsigned wcnt[32];
    __shared__ unsigned long long bbase;
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const u64 n_up = (n + blockDim.x - 1) / blockDim.x * blockDim.x;
    for (u64 base = (u64)blockIdx.x * blockDim.x; base < n_up; base += (u64)gridDim.x * blockDim.x) {
        const u64 i = base + threadIdx.x;
        u64 key = 0;
        bool in = false;
        if (i < n) {
            key = key_at(tx, i);
            const unsigned bin = (unsigned)(key >> 52);
            in = bin >= bin_lo && bin < bin_hi;
        }
        const unsigned m = __ballot_sync(B200_FULL_MASK, in);
        if (lane == 0) wcnt[wid] = __popc(m);
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned tot = 0;
            for (unsigned k = 0; k < nw; ++k) { const unsigned c = wcnt[k]; wcnt[k] = tot; tot += c; }
            bbase = tot ? atomicAdd(counter, (unsigned long long)tot) : 0ull;
        }
        __syncthreads();
        if (in) {
            const u64 slot = bbase + wcnt[wid] + __popc(m & ((1u << lane) - 1u));
            keys[slot] = key;
            vals[slot] = (PT)i;
        }
        __syncthreads();
    }
}

template <class PT>
__global__ void tie_mark_kernel(const u64 *keys, const PT *vals, u64 m, TieRec *list, u64 cap, unsigned long long *counter)
{
    u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; j < m; j += stride) {
        const u64 k = keys[j];
        const bool prev = j > 0 && keys[j - 1] == k, next = j + 1 < m && keys[j + 1] == k;
        if (prev || next) {
            const u64 slot = atomicAdd(counter, 1ull);
            if (slot < cap) { list[slot].j = j << 1 | (prev ? 1u : 0u); list[slot].pos = (u64)vals[j]; }
        }
    }
}

template <class PT>
__global__ void scatter_kernel(PT *vals, const u64 *idx, const u64 *pos, u64 cnt)
{
    u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; t < cnt; t += stride) vals[idx[t]] = (PT)pos[t];
}

template <class PT>
__global__ void find_primary_kernel(const PT *sa_full, u64 n, unsigned long long *out)
{
    u64 r = 1 + (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; r <= n; r += stride)
        if (sa_full[r] == 0) *out = r;
}

// BWT with the sentinel row removed, 16 symbols per word (is_bwt, is.c:209-223): symbol o is text[SA[r] - 1], r = o + (o >= primary)
template <class PT>
__global__ void bwt_raw_kernel(const PT *sa_full, const u32 *tx, u64 n, u64 primary, u32 *raw, u64 n_words)
{
    u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; w < n_words; w += stride) {
        u32 v = 0;
        for (int j = 0; j < 16; ++j) {
            const u64 o = (w << 4) + j;
            if (o >= n) break;
            const u64 r = o + (o >= primary ? 1 : 0);
            const u64 pos = r == 0 ? n : (u64)sa_full[r];
            v |= tx_base(tx, pos - 1) << ((15 - j) << 1);
        }
        raw[w] = v;
    }
}

// symbols of one 128-row block (8 words), only those below n
__device__ __forceinline__ void block_counts(const u32 *raw, u64 b, u64 n, u32 c[4])
{
    c[0] = c[1] = c[2] = c[3] = 0;
    for (int k = 0; k < 8; ++k) {
        const u64 w = b * 8 + k, first = w << 4;
        if (first >= n) break;
        const u32 v = raw[w];
        const int valid = n - first >= 16 ? 16 : (int)(n - first);
        const u32 mask = valid == 16 ? 0x55555555u : (0x55555555u & ~((1u << ((16 - valid) << 1)) - 1u));
        const u32 lo = v & 0x55555555u, hi = (v >> 1) & 0x55555555u;
        const u32 n3 = __popc(lo & hi & mask), n1 = __popc(lo & ~hi & mask), n2 = __popc(hi & ~lo & mask);
        c[1] += n1; c[2] += n2; c[3] += n3; c[0] += (u32)valid - n1 - n2 - n3;
    }
}

enum { CHUNK_BLOCKS = 256 };           // 128-row blocks per chunk (one thread block of the two kernels below)

__global__ void chunk_count_kernel(const u32 *raw, u64 n, u64 n_blocks, unsigned long long *chunk_cnt)
{
    __shared__ unsigned s[4];
    if (threadIdx.x < 4) s[threadIdx.x] = 0;
    __syncthreads();
    const u64 b = (u64)blockIdx.x * CHUNK_BLOCKS + threadIdx.x;
    if (b < n_blocks) {
        u32 c[4];
        block_counts(raw, b, n, c);
        for (int k = 0; k < 4; ++k) if (c[k]) atomicAdd(&s[k], c[k]);
    }
    __syncthreads();
    if (threadIdx.x < 4) chunk_cnt[(u64)blockIdx.x * 4 + threadIdx.x] = s[threadIdx.x];
}

// reference layout (bwt_bwtupdate_core): per 128 rows 4 x u64 counts before the block, then its 8 words
__global__ void layout_kernel(const u32 *raw, u64 n, u64 n_blocks, u64 n_words, const unsigned long long *chunk_base, u32 *bwt)
{
    __shared__ unsigned s[4][CHUNK_BLOCKS];
    const u64 b = (u64)blockIdx.x * CHUNK_BLOCKS + threadIdx.x;
    u32 c[4] = {0, 0, 0, 0};
    if (b < n_blocks) block_counts(raw, b, n, c);
    for (int k = 0; k < 4; ++k) s[k][threadIdx.x] = c[k];
    __syncthreads();
    if (b >= n_blocks) return;
    u64 acc[4];
    for (int k = 0; k < 4; ++k) {
        u64 a = chunk_base[(u64)blockIdx.x * 4 + k];
        for (unsigned t = 0; t < threadIdx.x; ++t) a += s[k][t];
        acc[k] = a;
    }
    u32 *o = bwt + b * 16;
    for (int k = 0; k < 4; ++k) { o[2 * k] = (u32)acc[k]; o[2 * k + 1] = (u32)(acc[k] >> 32); }
    for (int k = 0; k < 8; ++k) {
        const u64 w = b * 8 + k;
        if (w < n_words) o[8 + k] = raw[w];
    }
}

template <class PT>
__global__ void sample_sa_kernel(const PT *sa_full, u64 n_sa, int sa_intv, u64 *sa)
{
    u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (; j < n_sa; j += stride) sa[j] = j == 0 ? (u64)-1 : (u64)sa_full[j * (u64)sa_intv];
}

inline int grid_for(u64 items, int block, int n_sms)
{
    const u64 g = (items + block - 1) / block, cap = (u64)n_sms * 16;
    return (int)(g < 1 ? 1 : (g < cap ? g : cap));
}

struct HostText {                       // forward + reverse complement over the host's 2-bit pac
    const uint8_t *pac; u64 l_pac, n;
    inline int at(u64 i) const
    {
        if (i < l_pac) return (pac[i >> 2] >> ((~i & 3) << 1)) & 3;
        const u64 k = n - 1 - i;
        return 3 - ((pac[k >> 2] >> ((~k & 3) << 1)) & 3);
    }
    bool less(u64 a, u64 b) const        // suffix a < suffix b; the end of the text is the smallest symbol
    {
        for (u64 d = 0;; ++d) {
            if (a + d >= n) return true;
            if (b + d >= n) return false;
            const int ca = at(a + d), cb = at(b + d);
            if (ca != cb) return ca < cb;
        }
    }
};

template <class PT>
int sort_pairs(b200_ctx *ctx, u64 *k0, u64 *k1, PT *v0, PT *v1, u64 m, DevBuf &tmp, u64 **k_out, PT **v_out)
{
#ifdef B200_EMU
    std::vector<std::pair<u64, PT>> a(m);
    for (u64 i = 0; i < m; ++i) a[i] = std::make_pair(k0[i], v0[i]);
    std::stable_sort(a.begin(), a.end(), [](const std::pair<u64, PT> &x, const std::pair<u64, PT> &y) { return x.first < y.first; });
    for (u64 i = 0; i < m; ++i) { k1[i] = a[i].first; v1[i] = a[i].second; }
    *k_out = k1; *v_out = v1;
    (void)ctx; (void)tmp;
#else
    cub::DoubleBuffer<u64> dk(k0, k1);
    cub::DoubleBuffer<PT> dv(v0, v1);
    size_t tb = 0;
    B200_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (int)m, 0, 64, ctx->st));
    if (tmp.ensure(tb + 16)) return -1;
    B200_CHECK(cub::DeviceRadixSort::SortPairs(tmp.p, tb, dk, dv, (int)m, 0, 64, ctx->st));
    *k_out = dk.Current(); *v_out = dv.Current();
#endif
    return 0;
}

template <class PT>
int build_device(b200_ctx *ctx, const uint8_t *pac, int64_t l_pac, u64 *L2_out, u64 *primary_out, u64 sb_cnt[8][4])
{
    cudaStream_t st = ctx->st;
    const u64 n = (u64)l_pac * 2;
    const int sms = ctx->n_sms;
    const u64 n_words = (n + 15) >> 4;
    DevBuf d_tx, d_raw, d_k0, d_k1, d_v0, d_v1, d_tmp, d_hist, d_ties, d_ti, d_tp, d_chunk;
    struct Cleanup { DevBuf *b[12]; ~Cleanup() { for (DevBuf *x : b) x->release(); } }
        cl{{&d_tx, &d_raw, &d_k0, &d_k1, &d_v0, &d_v1, &d_tmp, &d_hist, &d_ties, &d_ti, &d_tp, &d_chunk}};

    // 1. text
    if (d_tx.ensure((n_words + 4) * 4)) return -1;
    B200_CHECK(cudaMemsetAsync(d_tx.p, 0, (n_words + 4) * 4, st));
    B200_LAUNCH(text2_kernel, grid_for(n_words, 256, sms), 256, 0, st, ctx->d_pac.as<u8>(), (u64)l_pac, d_tx.as<u32>(), n_words);
    B200_CHECK(cudaGetLastError());

    // 2. groups of 6-base bins
    if (d_hist.ensure(4096 * 8 + 64)) return -1;
    B200_CHECK(cudaMemsetAsync(d_hist.p, 0, 4096 * 8 + 64, st));
    B200_LAUNCH(hist_kernel, grid_for(n, 256, sms), 256, 0, st, d_tx.as<u32>(), n, d_hist.as<unsigned long long>());
    B200_CHECK(cudaGetLastError());
    std::vector<unsigned long long> hist(4096);
    B200_CHECK(cudaMemcpyAsync(hist.data(), d_hist.p, 4096 * 8, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    const u64 cap = ctx->idx_group_cap > 0 ? (u64)ctx->idx_group_cap : ((u64)1 << 29);
    u64 largest = 0, total = 0;
    for (int b = 0; b < 4096; ++b) { largest = std::max<u64>(largest, hist[b]); total += hist[b]; }
    if (total != n) { b200_set_error("index build: histogram covers %llu of %llu suffixes", (unsigned long long)total, (unsigned long long)n); return -1; }
    if (largest > cap || largest >= 0x7fffffffull) {
        b200_set_error("index build: one 6-base bin holds %llu suffixes (> group cap %llu); use the host builder for such texts",
                       (unsigned long long)largest, (unsigned long long)cap);
        return -1;
    }
    const u64 buf_items = std::min<u64>(std::max<u64>(cap, largest), n);
    if (d_k0.ensure(buf_items * 8) || d_k1.ensure(buf_items * 8) || d_v0.ensure(buf_items * sizeof(PT)) || d_v1.ensure(buf_items * sizeof(PT))) return -1;
    const u64 tie_cap = ctx->idx_tie_cap > 0 ? (u64)ctx->idx_tie_cap : ((u64)1 << 24);
    if (d_ties.ensure(tie_cap * sizeof(TieRec))) return -1;
    unsigned long long *d_cnt = d_hist.as<unsigned long long>() + 4096;         // [0] compaction cursor, [1] tie cursor, [2] primary
    PT *sa_full = ctx->d_sa_full.as<PT>();
    HostText T{pac, (u64)l_pac, n};

    u64 done = 0, n_ties_total = 0;
    int n_groups = 0;
    for (int lo = 0; lo < 4096;) {
        int hi = lo;
        u64 m = 0;
        while (hi < 4096 && m + hist[hi] <= buf_items) m += hist[hi++];
        if (hi == lo) { b200_set_error("index build: internal grouping error"); return -1; }
        if (m) {
            ++n_groups;
            B200_CHECK(cudaMemsetAsync(d_cnt, 0, 16, st));
            B200_LAUNCH((compact_kernel<PT>), grid_for(n, 256, sms), 256, 0, st, d_tx.as<u32>(), n, (unsigned)lo, (unsigned)hi,
                        d_k0.as<u64>(), d_v0.as<PT>(), d_cnt);
            B200_CHECK(cudaGetLastError());
            u64 *ks = nullptr; PT *vs = nullptr;
            if (sort_pairs<PT>(ctx, d_k0.as<u64>(), d_k1.as<u64>(), d_v0.as<PT>(), d_v1.as<PT>(), m, d_tmp, &ks, &vs)) return -1;
            B200_LAUNCH((tie_mark_kernel<PT>), grid_for(m, 256, sms), 256, 0, st, ks, vs, m, d_ties.as<TieRec>(), tie_cap, d_cnt + 1);
            B200_CHECK(cudaGetLastError());
            unsigned long long cur[2] = {0, 0};
            B200_CHECK(cudaMemcpyAsync(cur, d_cnt, 16, cudaMemcpyDeviceToHost, st));
            B200_CHECK(cudaStreamSynchronize(st));
            if (cur[0] != m) { b200_set_error("index build: group [%d,%d) compacted %llu of %llu suffixes", lo, hi, cur[0], (unsigned long long)m); return -1; }
            if (cur[1] > tie_cap) {
                b200_set_error("index build: %llu suffixes share their first 32 bases in one group (> tie cap %llu); use the host builder for repeat-rich texts",
                               cur[1], (unsigned long long)tie_cap);
                return -1;
            }
            if (cur[1]) {                      // order colliding suffixes on the host by direct comparison
                const u64 nt = cur[1];
                std::vector<TieRec> ties(nt);
                B200_CHECK(cudaMemcpyAsync(ties.data(), d_ties.p, nt * sizeof(TieRec), cudaMemcpyDeviceToHost, st));
                B200_CHECK(cudaStreamSynchronize(st));
                std::sort(ties.begin(), ties.end(), [](const TieRec &a, const TieRec &b) { return a.j < b.j; });
                std::vector<u64> idx(nt), pos(nt);
                for (u64 a = 0; a < nt;) {
                    u64 b = a + 1;
                    while (b < nt && (ties[b].j & 1)) ++b;
                    for (u64 t = a; t < b; ++t) { idx[t] = ties[t].j >> 1; pos[t] = ties[t].pos; }
                    std::sort(pos.begin() + a, pos.begin() + b, [&T](u64 x, u64 y) { return T.less(x, y); });
                    a = b;
                }
                if (d_ti.ensure(nt * 8) || d_tp.ensure(nt * 8)) return -1;
                B200_CHECK(cudaMemcpyAsync(d_ti.p, idx.data(), nt * 8, cudaMemcpyHostToDevice, st));
                B200_CHECK(cudaMemcpyAsync(d_tp.p, pos.data(), nt * 8, cudaMemcpyHostToDevice, st));
                B200_LAUNCH((scatter_kernel<PT>), grid_for(nt, 256, sms), 256, 0, st, vs, d_ti.as<u64>(), d_tp.as<u64>(), nt);
                B200_CHECK(cudaGetLastError());
                B200_CHECK(cudaStreamSynchronize(st));          // idx / pos go out of scope
                n_ties_total += nt;
            }
            B200_CHECK(cudaMemcpyAsync(sa_full + 1 + done, vs, m * sizeof(PT), cudaMemcpyDeviceToDevice, st));
            done += m;
        }
        lo = hi;
    }
    if (done != n) { b200_set_error("index build: sorted %llu of %llu suffixes", (unsigned long long)done, (unsigned long long)n); return -1; }
    {   // row 0 is the sentinel suffix; the reference's bwt_sa(0) is sa[0] = -1 (bwt.c:75)
        const PT m1 = (PT)~(PT)0;
        B200_CHECK(cudaMemcpyAsync(sa_full, &m1, sizeof(PT), cudaMemcpyHostToDevice, st));
        B200_CHECK(cudaStreamSynchronize(st));
    }
    d_k0.release(); d_k1.release(); d_v0.release(); d_v1.release(); d_tmp.release(); d_ties.release();
    ctx->idx_groups = n_groups; ctx->idx_ties = (i64)n_ties_total;

    // 3. primary, BWT words
    B200_LAUNCH((find_primary_kernel<PT>), grid_for(n, 256, sms), 256, 0, st, sa_full, n, d_cnt + 2);
    B200_CHECK(cudaGetLastError());
    unsigned long long primary = 0;
    B200_CHECK(cudaMemcpyAsync(&primary, d_cnt + 2, 8, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    if (primary == 0 || primary > n) { b200_set_error("index build: no primary row found"); return -1; }
    *primary_out = primary;
    if (d_raw.ensure((n_words + 8) * 4)) return -1;
    B200_LAUNCH((bwt_raw_kernel<PT>), grid_for(n_words, 128, sms), 128, 0, st, sa_full, d_tx.as<u32>(), n, (u64)primary, d_raw.as<u32>(), n_words);
    B200_CHECK(cudaGetLastError());

    // 4. Occ checkpoints: per-chunk symbol counts, prefix sums on the host (a few 100 k values), interleaved layout
    const u64 n_blocks = (n + 127) / 128, n_occ = n_blocks + 1;
    const u64 n_chunks = (n_blocks + CHUNK_BLOCKS - 1) / CHUNK_BLOCKS;
    if (n_chunks >= 0x7fffffffull) { b200_set_error("index build: text too long"); return -1; }
    if (d_chunk.ensure(n_chunks * 32)) return -1;
    B200_LAUNCH(chunk_count_kernel, (int)n_chunks, CHUNK_BLOCKS, 0, st, d_raw.as<u32>(), n, n_blocks, d_chunk.as<unsigned long long>());
    B200_CHECK(cudaGetLastError());
    std::vector<unsigned long long> cc(n_chunks * 4);
    B200_CHECK(cudaMemcpyAsync(cc.data(), d_chunk.p, n_chunks * 32, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    unsigned long long run[4] = {0, 0, 0, 0};
    for (int s = 0; s < 8; ++s) for (int c = 0; c < 4; ++c) sb_cnt[s][c] = 0;
    const u64 chunk_rows = (u64)CHUNK_BLOCKS * 128;
    for (u64 k = 0; k < n_chunks; ++k) {
        const u64 row0 = k * chunk_rows;
        if (row0 && (row0 & 0xffffffffull) == 0 && (row0 >> 32) < 8) for (int c = 0; c < 4; ++c) sb_cnt[row0 >> 32][c] = run[c];
        for (int c = 0; c < 4; ++c) { const unsigned long long v = cc[k * 4 + c]; cc[k * 4 + c] = run[c]; run[c] += v; }
    }
    if (run[0] + run[1] + run[2] + run[3] != n) { b200_set_error("index build: BWT symbol counts do not add up"); return -1; }
    B200_CHECK(cudaMemcpyAsync(d_chunk.p, cc.data(), n_chunks * 32, cudaMemcpyHostToDevice, st));
    B200_LAUNCH(layout_kernel, (int)n_chunks, CHUNK_BLOCKS, 0, st, d_raw.as<u32>(), n, n_blocks, n_words,
                d_chunk.as<unsigned long long>(), ctx->d_bwt.as<u32>());
    B200_CHECK(cudaGetLastError());
    {   // closing checkpoint after the last symbol (bwtindex.c:170)
        u64 fin[4] = {run[0], run[1], run[2], run[3]};
        B200_CHECK(cudaMemcpyAsync(ctx->d_bwt.as<u32>() + (n_occ - 1) * 8 + n_words, fin, 32, cudaMemcpyHostToDevice, st));
        B200_CHECK(cudaStreamSynchronize(st));
    }
    L2_out[0] = 0;
    for (int c = 0; c < 4; ++c) L2_out[c + 1] = L2_out[c] + run[c];

    // 5. sampled SA
    B200_LAUNCH((sample_sa_kernel<PT>), grid_for(ctx->ix.n_sa, 256, sms), 256, 0, st, sa_full, ctx->ix.n_sa, ctx->ix.sa_intv, ctx->d_sa.as<u64>());
    B200_CHECK(cudaGetLastError());
    B200_CHECK(cudaStreamSynchronize(st));
    return 0;
}

}  // namespace



extern "C" int b200_index_build_device(b200_ctx_t *ctx, const uint8_t *pac, int64_t l_pac, int n_seqs, const int64_t *ann_offset,
                                       const int32_t *ann_len, const int32_t *ann_is_alt, int flags)
{
    if (!ctx || !pac || l_pac <= 0) { b200_set_error("null / empty argument"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->st;
    const u64 n = (u64)l_pac * 2;
    if (n >> 35) { b200_set_error("seq_len >= 2^35 not supported by the occ32 layout"); return -1; }
    const u64 n_words = (n + 15) >> 4, n_occ = (n + 127) / 128 + 1;
    const u64 bwt_size = n_words + n_occ * 8;
    const size_t bwt_bytes = ((size_t)bwt_size * 4 + 63) / 64 * 64 + 128;
    const int sa_intv = 32;
    const u64 n_sa = (n + sa_intv) / sa_intv;
    const bool narrow = n < 0xffffffffull;
    const size_t pac_bytes = (size_t)(l_pac / 4 + 1);
    if (ctx->d_bwt.ensure(bwt_bytes) || ctx->d_sa.ensure((size_t)n_sa * 8) || ctx->d_pac.ensure(pac_bytes) ||
        ctx->d_sa_full.ensure_exact((size_t)(n + 1) * (narrow ? 4 : 8)) ||
        ctx->d_ann_off.ensure((size_t)n_seqs * 8) || ctx->d_ann_len.ensure((size_t)n_seqs * 4) || ctx->d_ann_alt.ensure((size_t)n_seqs * 4)) return -1;
    B200_CHECK(cudaMemsetAsync(ctx->d_bwt.p, 0, bwt_bytes, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_pac.p, pac, pac_bytes, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_ann_off.p, ann_offset, (size_t)n_seqs * 8, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_ann_len.p, ann_len, (size_t)n_seqs * 4, cudaMemcpyHostToDevice, st));
    B200_CHECK(cudaMemcpyAsync(ctx->d_ann_alt.p, ann_is_alt, (size_t)n_seqs * 4, cudaMemcpyHostToDevice, st));
    DevIndex &ix = ctx->ix;
    memset(&ix, 0, sizeof(ix));
    ctx->has_index = false;
    ix.bwt = ctx->d_bwt.as<uint4>();
    ix.seq_len = n;
    ix.sa = ctx->d_sa.as<u64>(); ix.n_sa = n_sa; ix.sa_intv = sa_intv;
    ix.pac = ctx->d_pac.as<u8>(); ix.l_pac = l_pac; ix.n_seqs = n_seqs;
    ix.ann_off = ctx->d_ann_off.as<i64>(); ix.ann_len = ctx->d_ann_len.as<i32>(); ix.ann_alt = ctx->d_ann_alt.as<i32>();
    u64 L2[5], primary = 0;
    const int rc = narrow ? build_device<u32>(ctx, pac, l_pac, L2, &primary, ix.sb_cnt) : build_device<u64>(ctx, pac, l_pac, L2, &primary, ix.sb_cnt);
    if (rc) return rc;
    ix.primary = primary;
    for (int i = 0; i < 5; ++i) ix.L2[i] = L2[i];
    if (narrow) ix.sa_full32 = ctx->d_sa_full.as<u32>(); else ix.sa_full64 = ctx->d_sa_full.as<u64>();
    ctx->idx_bwt_size = bwt_size;
    if (b200_index_finish(ctx, flags)) return -1;
    ctx->bwt_identity = nullptr;
    ctx->has_index = true;
    return 0;
}

// out[0]=primary [1..5]=L2 [6]=seq_len [7]=bwt_size [8]=n_sa [9]=sa_intv [10]=l_pac [11]=sort groups [12]=suffixes ordered on the host
extern "C" int b200_index_info(b200_ctx_t *ctx, uint64_t *out)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    out[0] = ctx->ix.primary;
    for (int i = 0; i < 5; ++i) out[1 + i] = ctx->ix.L2[i];
    out[6] = ctx->ix.seq_len; out[7] = ctx->idx_bwt_size; out[8] = ctx->ix.n_sa; out[9] = (u64)ctx->ix.sa_intv; out[10] = (u64)ctx->ix.l_pac;
    out[11] = (u64)ctx->idx_groups; out[12] = (u64)ctx->idx_ties;
    return 0;
}

// the reference-layout arrays of the resident index, for a host program that wants to keep / write them (bwt_size words, n_sa values)
extern "C" int b200_index_download(b200_ctx_t *ctx, uint32_t *bwt, uint64_t *sa)
{
    if (!ctx || !ctx->has_index) { b200_set_error("no index uploaded"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    if (bwt) B200_CHECK(cudaMemcpyAsync(bwt, ctx->d_bwt.p, (size_t)ctx->idx_bwt_size * 4, cudaMemcpyDeviceToHost, ctx->st));
    if (sa) B200_CHECK(cudaMemcpyAsync(sa, ctx->d_sa.p, (size_t)ctx->ix.n_sa * 8, cudaMemcpyDeviceToHost, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    return 0;
}
