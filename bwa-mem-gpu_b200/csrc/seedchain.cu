// Stage 2: seed expansion + SA location (bwamem.c:272-294, bwt_sa bwt.c:86-96) and seed chaining +
// chain filtering (mem_chain bwamem.c:258-322, test_and_merge :197-218, mem_chain_weight :220-239,
// mem_chain_flt :334-392).
//
// Layout.  Seeds of read r live at seeds[seed_off[r] .. +seed_n[r]) in exactly the order the reference
// visits them (interval order, then SA row order with the max_occ stride), so the order-dependent
// chaining below sees the same sequence.  SA location is one thread per seed: with the device-expanded
// full suffix array it is a single 4/8-byte gather instead of the reference's ~31-step LF walk.
// Chaining is inherently sequential per read (each seed looks up the closest chain built so far), so it
// is one thread per read over a B-tree that reproduces klib's node layout (kbtree.cuh); chain weights
// are sorted with the order-exact introsort (ksort.cuh).
#include "b200_internal.h"
#include "kbtree.cuh"
#include "ksort.cuh"

// ---------------------------------------------------------------- seed counting / filling
__device__ __forceinline__ Intv ld_intv(const Intv *p)
{
    const uint4 *q = (const uint4 *)p;
    uint4 a = q[0], b = q[1];
    Intv r;
    r.x0 = (u64)a.y << 32 | a.x; r.x1 = (u64)a.w << 32 | a.z;
    r.x2 = (u64)b.y << 32 | b.x; r.info = (u64)b.w << 32 | b.z;
    return r;
}

__device__ __forceinline__ int seeds_of_intv(u64 x2, int max_occ)
{
    // for (k = count = 0; k < x2 && count < max_occ; k += step, ++count)   (bwamem.c:285-286)
    u64 step = x2 > (u64)max_occ ? x2 / (u64)max_occ : 1;
    u64 cnt = (x2 + step - 1) / step;
    return (int)(cnt < (u64)max_occ ? cnt : (u64)max_occ);
}

__global__ void seed_count_kernel(SeedParams P)
{
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_reads) return;
    const Intv *iv = P.intv + (size_t)r * P.cap_intv;
    int n = P.intv_n[r], tot = 0;
    int b = 0, e = 0, l_rep = 0;                        // bwamem.c:272-279
    for (int i = 0; i < n; ++i) {
        Intv p = ld_intv(&iv[i]);
        tot += seeds_of_intv(p.x2, P.max_occ);
        if (p.x2 <= (u64)P.max_occ) continue;
        int sb = (int)(p.info >> 32), se = (int)(u32)p.info;
        if (sb > e) { l_rep += e - b; b = sb; e = se; }
        else e = e > se ? e : se;
    }
    l_rep += e - b;
    P.seed_n[r] = tot;
    P.l_rep[r] = l_rep;
}

__global__ void seed_fill_kernel(SeedParams P)
{
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_reads) return;
    const Intv *iv = P.intv + (size_t)r * P.cap_intv;
    int n = P.intv_n[r];
    Seed *s = P.seeds + P.seed_off[r];
    for (int i = 0; i < n; ++i) {
        Intv p = ld_intv(&iv[i]);
        int slen = (int)(u32)p.info - (int)(p.info >> 32);
        u64 step = p.x2 > (u64)P.max_occ ? p.x2 / (u64)P.max_occ : 1;
        int count = 0;
        for (u64 k = 0; k < p.x2 && count < P.max_occ; k += step, ++count, ++s) {
            s->rbeg = (i64)(p.x0 + k);
            s->qbeg = (int)(p.info >> 32);
            s->len = slen; s->score = slen; s->rid = 0;
        }
    }
}

__global__ void sa_locate_kernel(SeedParams P)
{
    i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    int steps = 0;
    if (t < P.n_seeds) {
        Seed *s = &P.seeds[t];
        u64 pos = dev_sa(P.ix, (u64)s->rbeg, &steps);                 // bwamem.c:290
        s->rbeg = (i64)pos;
        s->rid = dev_intv2rid(P.ix, (i64)pos, (i64)pos + s->len);      // bwamem.c:293
    }
    int tot = __reduce_add_sync(B200_FULL_MASK, steps);
    if ((threadIdx.x & 31) == 0 && tot) atomicAdd(&P.stats[CNT_LF], (unsigned long long)tot);
}

// ---------------------------------------------------------------- chaining
struct ChainCtx {
    const Seed *S;
    int *next, *c_first, *c_last, *c_n, *c_w, *c_kept, *c_shadow;
};
struct PosOf {
    const Seed *S; const int *c_first;
    __device__ __forceinline__ i64 operator()(int id) const { return S[c_first[id]].rbeg; }
};
struct WeightGreater {
    const int *c_w;
    __device__ __forceinline__ bool operator()(int a, int b) const { return c_w[a] > c_w[b]; }   // flt_lt, bwamem.c:331
};

// bwamem.c:197-218
__device__ __forceinline__ int test_and_merge(const mem_opt_t &opt, i64 l_pac, ChainCtx &C, int c, int s)
{
    const Seed p = C.S[s];
    const Seed last = C.S[C.c_last[c]];
    const Seed first = C.S[C.c_first[c]];
    i64 qend = last.qbeg + last.len, rend = last.rbeg + last.len;
    if (p.rid != first.rid) return 0;
    if (p.qbeg >= first.qbeg && p.qbeg + p.len <= qend && p.rbeg >= first.rbeg && p.rbeg + p.len <= rend) return 1;
    if ((last.rbeg < l_pac || first.rbeg < l_pac) && p.rbeg >= l_pac) return 0;
    i64 x = p.qbeg - last.qbeg, y = p.rbeg - last.rbeg;
    if (y >= 0 && x - y <= opt.w && y - x <= opt.w && x - last.len < opt.max_chain_gap && y - last.len < opt.max_chain_gap) {
        C.next[C.c_last[c]] = s; C.next[s] = -1; C.c_last[c] = s; ++C.c_n[c];
        return 1;
    }
    return 0;
}

// bwamem.c:220-239
__device__ __forceinline__ int chain_weight(const ChainCtx &C, int c)
{
    i64 end = 0; int w = 0, tmp;
    for (int s = C.c_first[c]; s >= 0; s = C.next[s]) {
        const Seed t = C.S[s];
        if (t.qbeg >= end) w += t.len;
        else if (t.qbeg + t.len > end) w += (int)(t.qbeg + t.len - end);
        end = end > t.qbeg + t.len ? end : t.qbeg + t.len;
    }
    tmp = w; w = 0; end = 0;
    for (int s = C.c_first[c]; s >= 0; s = C.next[s]) {
        const Seed t = C.S[s];
        if (t.rbeg >= end) w += t.len;
        else if (t.rbeg + t.len > end) w += (int)(t.rbeg + t.len - end);
        end = end > t.rbeg + t.len ? end : t.rbeg + t.len;
    }
    w = w < tmp ? w : tmp;
    return w < (1 << 30) ? w : (1 << 30) - 1;
}

__global__ void __launch_bounds__(128) chain_kernel(ChainParams P)
{
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_reads) return;
    const int ns = P.seed_n[r];
    const i64 soff = P.seed_off[r];
    const int len = (int)(P.off[r + 1] - P.off[r]);
    P.chn_n[r] = 0; P.cseed_n[r] = 0;
    if (ns == 0) return;
    int *W = P.scratch + 16 * soff + 128 * (i64)r;
    ChainCtx C;
    C.S = P.seeds + soff;
    C.next = W; C.c_first = W + ns; C.c_last = W + 2 * ns; C.c_n = W + 3 * ns; C.c_w = W + 4 * ns;
    C.c_kept = W + 5 * ns; C.c_shadow = W + 6 * ns;
    int *ord = W + 7 * ns, *keep = W + 8 * ns;
    KbTree T;
    T.init(W + 9 * ns, (7 * ns + 128) / KBT_STRIDE);
    PosOf pos{C.S, C.c_first};
    const mem_opt_t &opt = P.opt;
    int nc = 0;
    for (int s = 0; s < ns; ++s) {                                     // bwamem.c:286-307
        if (C.S[s].rid < 0) continue;
        int to_add = 0;
        if (T.n_keys) {
            int low = T.lower(C.S[s].rbeg, pos);
            if (low < 0 || !test_and_merge(opt, P.ix.l_pac, C, low, s)) to_add = 1;
        } else to_add = 1;
        if (to_add) {
            int id = nc++;
            C.c_first[id] = C.c_last[id] = s; C.c_n[id] = 1; C.next[s] = -1;
            T.put(id, pos);
            if (T.overflow) { atomicMax(P.overflow, 1); return; }
        }
    }
    int n = T.traverse(ord);                                           // bwamem.c:311-315
    const float frac_rep = (float)P.l_rep[r] / len;                     // bwamem.c:317

    if (P.stage >= 1 && n > 0) {                                        // mem_chain_flt, bwamem.c:334-392
        int k = 0;
        for (int i = 0; i < n; ++i) {
            int id = ord[i];
            C.c_shadow[id] = -1; C.c_kept[id] = 0;
            C.c_w[id] = chain_weight(C, id);
            if (C.c_w[id] >= opt.min_chain_weight) ord[k++] = id;
        }
        n = k;
        if (n > 0) {
            ks_introsort(ord, n, WeightGreater{C.c_w});
            int nk = 0;
            C.c_kept[ord[0]] = 3; keep[nk++] = 0;
            for (int i = 1; i < n; ++i) {
                const int ci = ord[i];
                const int beg_i = C.S[C.c_first[ci]].qbeg;
                const int end_i = C.S[C.c_last[ci]].qbeg + C.S[C.c_last[ci]].len;
                const int alt_i = P.ix.ann_alt[C.S[C.c_first[ci]].rid] != 0;
                int large_ovlp = 0, kk;
                for (kk = 0; kk < nk; ++kk) {
                    const int j = keep[kk], cj = ord[j];
                    const int beg_j = C.S[C.c_first[cj]].qbeg;
                    const int end_j = C.S[C.c_last[cj]].qbeg + C.S[C.c_last[cj]].len;
                    const int alt_j = P.ix.ann_alt[C.S[C.c_first[cj]].rid] != 0;
                    int b_max = beg_j > beg_i ? beg_j : beg_i;
                    int e_min = end_j < end_i ? end_j : end_i;
                    if (e_min > b_max && (!alt_j || alt_i)) {
                        int li = end_i - beg_i, lj = end_j - beg_j;
                        int min_l = li < lj ? li : lj;
                        if ((float)(e_min - b_max) >= (float)min_l * opt.mask_level && min_l < opt.max_chain_gap) {
                            large_ovlp = 1;
                            if (C.c_shadow[cj] < 0) C.c_shadow[cj] = i;
                            if ((float)C.c_w[ci] < (float)C.c_w[cj] * opt.drop_ratio && C.c_w[cj] - C.c_w[ci] >= opt.min_seed_len << 1)
                                break;
                        }
                    }
                }
                if (kk == nk) { keep[nk++] = i; C.c_kept[ci] = large_ovlp ? 2 : 3; }
            }
            for (int i = 0; i < nk; ++i) {
                int cj = ord[keep[i]];
                if (C.c_shadow[cj] >= 0) C.c_kept[ord[C.c_shadow[cj]]] = 1;
            }
            int i2, k2;
            for (i2 = k2 = 0; i2 < n; ++i2) {                          // bwamem.c:380-385
                int kp = C.c_kept[ord[i2]];
                if (kp == 0 || kp == 3) continue;
                if (++k2 >= opt.max_chain_extend) break;
            }
            for (; i2 < n; ++i2) if (C.c_kept[ord[i2]] < 3) C.c_kept[ord[i2]] = 0;
        }
    }

    // emit (chains and their seeds share the read's seed offset: both are bounded by seed_n)
    ChainRec *out = P.chains + soff;
    Seed *cs = P.cseeds + soff;
    int nout = 0, nseed = 0;
    for (int i = 0; i < n; ++i) {
        int id = ord[i];
        if (P.stage >= 1 && C.c_kept[id] == 0) continue;
        ChainRec c;
        c.pos = C.S[C.c_first[id]].rbeg;
        c.seed_off = soff + nseed;
        c.n = C.c_n[id];
        c.rid = C.S[C.c_first[id]].rid;
        c.w = P.stage >= 1 ? C.c_w[id] : 0;
        c.kept = P.stage >= 1 ? C.c_kept[id] : 0;
        c.first = P.stage >= 1 ? C.c_shadow[id] : 0;
        c.is_alt = P.ix.ann_alt[c.rid] != 0;
        c.frac_rep = frac_rep;
        c.read = r;
        out[nout++] = c;
        for (int s = C.c_first[id]; s >= 0; s = C.next[s]) cs[nseed++] = C.S[s];
    }
    P.chn_n[r] = nout; P.cseed_n[r] = nseed;
}

// ---------------------------------------------------------------- exclusive scan (i32 -> i64 offsets)
#define SCAN_BLOCK 256
#define SCAN_ITEMS 4
__global__ void scan_block_sums(const i32 *in, int n, i64 *sums)
{
    __shared__ i64 sh[SCAN_BLOCK];
    int base = blockIdx.x * SCAN_BLOCK * SCAN_ITEMS + threadIdx.x * SCAN_ITEMS;
    i64 s = 0;
    for (int i = 0; i < SCAN_ITEMS; ++i) if (base + i < n) s += in[base + i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int d = SCAN_BLOCK / 2; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d) sh[threadIdx.x] += sh[threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[blockIdx.x] = sh[0];
}
__global__ void scan_sums_serial(i64 *sums, int nb, i64 *total)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        i64 acc = 0;
        for (int i = 0; i < nb; ++i) { i64 v = sums[i]; sums[i] = acc; acc += v; }
        *total = acc;
    }
}
__global__ void scan_apply(const i32 *in, int n, const i64 *sums, i64 *out)
{
    __shared__ i64 sh[SCAN_BLOCK];
    int base = blockIdx.x * SCAN_BLOCK * SCAN_ITEMS + threadIdx.x * SCAN_ITEMS;
    i64 v[SCAN_ITEMS], s = 0;
    for (int i = 0; i < SCAN_ITEMS; ++i) { v[i] = base + i < n ? in[base + i] : 0; s += v[i]; }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int d = 1; d < SCAN_BLOCK; d <<= 1) {
        i64 t = (int)threadIdx.x >= d ? sh[threadIdx.x - d] : 0;
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    i64 acc = sums[blockIdx.x] + sh[threadIdx.x] - s;
    for (int i = 0; i < SCAN_ITEMS; ++i) { if (base + i < n) out[base + i] = acc; acc += v[i]; }
}

int launch_exclusive_scan(const i32 *in, i64 *out, int n, i64 *tmp, cudaStream_t st)
{
    if (n <= 0) { B200_CHECK(cudaMemsetAsync(out, 0, sizeof(i64), st)); return 0; }
    int nb = (n + SCAN_BLOCK * SCAN_ITEMS - 1) / (SCAN_BLOCK * SCAN_ITEMS);
    B200_LAUNCH(scan_block_sums, nb, SCAN_BLOCK, 0, st, in, n, tmp);
    B200_LAUNCH(scan_sums_serial, 1, 32, 0, st, tmp, nb, out + n);
    B200_LAUNCH(scan_apply, nb, SCAN_BLOCK, 0, st, in, n, (const i64 *)tmp, out);
    B200_CHECK(cudaGetLastError());
    return 0;
}

int launch_seed_count(const SeedParams &p, cudaStream_t st)
{
    if (p.n_reads == 0) return 0;
    B200_LAUNCH(seed_count_kernel, (p.n_reads + 127) / 128, 128, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_seed_fill(const SeedParams &p, cudaStream_t st)
{
    if (p.n_reads == 0) return 0;
    B200_LAUNCH(seed_fill_kernel, (p.n_reads + 127) / 128, 128, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_sa_locate(const SeedParams &p, cudaStream_t st)
{
    if (p.n_seeds == 0) return 0;
    B200_LAUNCH(sa_locate_kernel, (int)((p.n_seeds + 255) / 256), 256, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}
int launch_chain(const ChainParams &p, int emit, cudaStream_t st)
{
    (void)emit;
    if (p.n_reads == 0) return 0;
    B200_LAUNCH(chain_kernel, (p.n_reads + 127) / 128, 128, 0, st, p);
    B200_CHECK(cudaGetLastError());
    return 0;
}
