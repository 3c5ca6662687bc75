// k-mer interval table for the seeding kernel.
//
// The bi-interval (x0, x1, x2) bwt_extend (bwt.c:262-275) arrives at after reading a string W depends on W only: x0 is
// the suffix-array interval of W, x1 that of its reverse complement, x2 their common size.  About half of all bwt_extend
// calls of mem_collect_intv are made on strings of at most ten bases (the first steps of every forward sweep, and the
// many short intervals a backward sweep carries along), and those are exactly the calls whose two interval ends lie
// far apart, i.e. the ones that cost TWO index sectors.  Their results are tabulated once per index: entry
// (kmer_off[k] + W) holds the bi-interval of the k-base string with code W (first base most significant), so the seeding
// kernel answers such a call with one 16-byte load from a table small enough to stay in L2 (22 MB for k <= 10).  The
// entries are produced by the same device code the kernel would have run (forward extension base by base, occ32_child),
// so they are bit-identical to it by construction.  That includes EMPTY intervals: bwt_seed_strategy1 (bwt.c:358-379)
// keeps extending an interval that has become empty until the seed is long enough, and the (path-defined) x0 / x1 it
// carries meanwhile are those of a forward extension from the first base, which is how the entries are built; a
// backward sweep never uses more than x2 of an empty child.
//
// Every record also carries the size-change mask of its string's prefixes (bit t set when the prefixes of t and t + 1
// bases have different interval sizes; sizes only shrink with length, so two prefixes have equal sizes exactly when no
// bit lies between them).  With the mask of the depth-k record of a window the seeding kernel settles the duplicate test
// of bwt.c:339 for ALL its table-range list entries of one backward position at once (smem3.cu).
#include "b200_internal.h"
#include "occ32.cuh"

template <class PT, bool PACK33>
__global__ void build_kmer_tab_kernel(DevIndex ix, uint4 *tab, int k, u32 off)
{
    const u64 W = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (W >> (2 * k)) return;
    int b = (int)((W >> (2 * (k - 1))) & 3);
    PT x0 = (PT)ix.L2[b] + 1, x2 = (PT)(ix.L2[b + 1] - ix.L2[b]), x1 = (PT)ix.L2[3 - b] + 1;    // bwt.c:301 (bwt_set_intv)
    u32 chg = 0;                                                 // bit t: the prefixes of t and t + 1 bases differ in size
    for (int t = 1; t < k; ++t) {
        b = (int)((W >> (2 * (k - 1 - t))) & 3);
        PT n0, n1, n2; int two;
        occ32_child<PT>(ix, x0, x1, x2, 0, 3 - b, n0, n1, n2, two);                             // forward extension, bwt.c:310
        chg |= (u32)(n2 != x2) << t;
        x0 = n0; x1 = n1; x2 = n2;
    }
    if (sizeof(PT) == 4) tab[off + W] = make_uint4((u32)x0, (u32)x1, (u32)x2, chg);
    else if (PACK33)                                             // Ent33 (smem3.cu): bit 32 of x0 / x1 / x2 in the top bits of the last word
        tab[off + W] = make_uint4((u32)x0, (u32)x1, (u32)x2, chg | (u32)((u64)x0 >> 32) << 31 | ((u32)((u64)x1 >> 32) & 1u) << 30 | ((u32)((u64)x2 >> 32) & 1u) << 29);
    else {
        uint4 *p = tab + ((size_t)off + W) * 2;
        p[0] = make_uint4((u32)x0, (u32)((u64)x0 >> 32), (u32)x1, (u32)((u64)x1 >> 32));
        p[1] = make_uint4((u32)x2, (u32)((u64)x2 >> 32), chg, 0u);
    }
}

// fills ix.kmer_off; the caller sets ix.kmer_tab / ix.kmer_k after the build
// wide: 0 = 16-byte entries of u32, 1 = 32-byte entries of u64, 2 = 16-byte packed entries of 33-bit values
int launch_build_kmer_tab(DevIndex &ix, uint4 *tab, int kmax, int wide, cudaStream_t st)
{
    u32 off = 0;
    for (int k = 1; k <= kmax; ++k) { ix.kmer_off[k] = off; off += 1u << (2 * k); }
    for (int k = 1; k <= kmax; ++k) {
        const u64 n = 1ull << (2 * k);
        const int grid = (int)((n + 255) / 256);
        if (wide == 2) B200_LAUNCH((build_kmer_tab_kernel<u64, true>), grid, 256, 0, st, ix, tab, k, ix.kmer_off[k]);
        else if (wide) B200_LAUNCH((build_kmer_tab_kernel<u64, false>), grid, 256, 0, st, ix, tab, k, ix.kmer_off[k]);
        else B200_LAUNCH((build_kmer_tab_kernel<u32, false>), grid, 256, 0, st, ix, tab, k, ix.kmer_off[k]);
        B200_CHECK(cudaGetLastError());
    }
    return 0;
}

size_t kmer_tab_entries(int kmax)
{
    size_t n = 0;
    for (int k = 1; k <= kmax; ++k) n += (size_t)1 << (2 * k);
    return n;
}
