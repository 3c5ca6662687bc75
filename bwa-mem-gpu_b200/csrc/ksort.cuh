// Order-exact sorting.
//
// Several seams of BWA-MEM sort with klib's *unstable* introsort (ksort.h:174-227: median-of-3
// quicksort, sub-ranges of <= 16 left for a final insertion sort, combsort when the depth budget
// runs out) and the resulting order of equal keys is observable in the output (chains by weight,
// bwamem.c:348; regions by end / by score, bwamem.c:448,485).  To stay bit-exact we run the same
// algorithm -- same pivot rule, same scan/swap sequence, same thresholds -- restated here as an
// index-based template usable from one device thread.  The element type only needs copy/assign.
#pragma once
#include "simt.h"

template <class T, class Lt>
__device__ __forceinline__ void ks_insertion(T *a, int lo, int hi /* exclusive */, Lt lt)
{
    for (int i = lo + 1; i < hi; ++i)
        for (int j = i; j > lo && lt(a[j], a[j - 1]); --j) { T tmp = a[j]; a[j] = a[j - 1]; a[j - 1] = tmp; }
}

// ksort.h:153-173
template <class T, class Lt>
__device__ inline void ks_comb(T *a, int n, Lt lt)
{
    const double shrink = 1.2473309501039786540366528676643;
    int gap = n, swapped;
    do {
        if (gap > 2) {
            gap = (int)((double)gap / shrink);
            if (gap == 9 || gap == 10) gap = 11;
        }
        swapped = 0;
        for (int i = 0; i + gap < n; ++i) {
            int j = i + gap;
            if (lt(a[j], a[i])) { T tmp = a[i]; a[i] = a[j]; a[j] = tmp; swapped = 1; }
        }
    } while (swapped || gap > 2);
    if (gap != 1) ks_insertion(a, 0, n, lt);
}

// ksort.h:174-227
template <class T, class Lt>
__device__ inline void ks_introsort(T *a, int n, Lt lt)
{
    if (n < 1) return;
    if (n == 2) {
        if (lt(a[1], a[0])) { T tmp = a[0]; a[0] = a[1]; a[1] = tmp; }
        return;
    }
    int d = 2;
    while ((1ul << d) < (unsigned long)n) ++d;
    struct Frame { int l, r, d; } stack[72];
    int top = 0, s = 0, t = n - 1;
    d <<= 1;
    for (;;) {
        if (s < t) {
            if (--d == 0) {
                ks_comb(a + s, t - s + 1, lt);
                t = s;
                continue;
            }
            int i = s, j = t, k = i + ((j - i) >> 1) + 1;
            if (lt(a[k], a[i])) {
                if (lt(a[k], a[j])) k = j;
            } else k = lt(a[j], a[i]) ? i : j;
            T rp = a[k];
            if (k != t) { T tmp = a[k]; a[k] = a[t]; a[t] = tmp; }
            for (;;) {
                do ++i; while (lt(a[i], rp));
                do --j; while (i <= j && lt(rp, a[j]));
                if (j <= i) break;
                T tmp = a[i]; a[i] = a[j]; a[j] = tmp;
            }
            { T tmp = a[i]; a[i] = a[t]; a[t] = tmp; }
            if (i - s > t - i) {
                if (i - s > 16) { stack[top].l = s; stack[top].r = i - 1; stack[top].d = d; ++top; }
                s = t - i > 16 ? i + 1 : t;
            } else {
                if (t - i > 16) { stack[top].l = i + 1; stack[top].r = t; stack[top].d = d; ++top; }
                t = i - s > 16 ? i - 1 : s;
            }
        } else {
            if (top == 0) { ks_insertion(a, 0, n, lt); return; }
            --top; s = stack[top].l; t = stack[top].r; d = stack[top].d;
        }
    }
}
