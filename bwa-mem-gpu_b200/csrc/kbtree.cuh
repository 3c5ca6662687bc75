// Chain B-tree with the reference's exact shape.
//
// mem_chain (bwamem.c:258-322) keeps chains in klib's B-tree (kbtree.h) keyed by chain position and,
// for every seed, asks for the chain with the greatest pos <= rbeg (kb_intervalp, kbtree.h:152-168).
// Duplicate positions are legal, and which duplicate is returned -- and the final in-order sequence
// -- depends on the node layout produced by klib's split rule.  To be bit-exact in every case we
// rebuild the same tree: minimum degree t = 6 (KB_DEFAULT_SIZE 512, 32-byte mem_chain_t keys:
// ((512-4-8)/(8+32)+1)>>1, kbtree.h:53), lower-bound search inside a node (kbtree.h:117-130),
// pre-emptive split on the way down (kbtree.h:170-226).  Keys are chain ids; positions come from a
// caller-supplied accessor.  Nodes live in a caller-provided int pool (no allocation).
#pragma once
#include "simt.h"

#define KBT_T 6
#define KBT_MAXK (2 * KBT_T - 1)
#define KBT_STRIDE (2 + KBT_MAXK + KBT_MAXK + 1)   // n, is_internal, keys[11], child[12]

struct KbTree {
    int *pool;
    int cap_nodes, n_nodes, root, n_keys, overflow;

    __device__ __forceinline__ int &N(int x) { return pool[x * KBT_STRIDE]; }
    __device__ __forceinline__ int &INTERNAL(int x) { return pool[x * KBT_STRIDE + 1]; }
    __device__ __forceinline__ int &KEY(int x, int i) { return pool[x * KBT_STRIDE + 2 + i]; }
    __device__ __forceinline__ int &CHILD(int x, int i) { return pool[x * KBT_STRIDE + 2 + KBT_MAXK + i]; }

    __device__ __forceinline__ int new_node(int internal)
    {
        if (n_nodes >= cap_nodes) { overflow = 1; return 0; }
        int x = n_nodes++;
        N(x) = 0; INTERNAL(x) = internal;
        return x;
    }
    __device__ __forceinline__ void init(int *pool_, int cap_nodes_)
    {
        pool = pool_; cap_nodes = cap_nodes_; n_nodes = 0; n_keys = 0; overflow = 0;
        root = new_node(0);
    }

    // __kb_getp_aux (kbtree.h:117-130): index of the first key equal to k, else of the greatest key < k
    // (-1 if none); *r = sign(k - key[returned index]) as the reference leaves it.
    template <class Pos>
    __device__ __forceinline__ int find_in_node(int x, i64 k, int *r, Pos pos)
    {
        int n = N(x);
        if (n == 0) return -1;
        int begin = 0, end = n;
        while (begin < end) {
            int mid = (begin + end) >> 1;
            if (pos(KEY(x, mid)) < k) begin = mid + 1; else end = mid;
        }
        if (begin == n) { *r = 1; return n - 1; }
        i64 kp = pos(KEY(x, begin));
        *r = (kp < k) - (k < kp);
        if (*r < 0) --begin;
        return begin;
    }

    // kb_intervalp (kbtree.h:152-168), lower bound only: chain id or -1
    template <class Pos>
    __device__ __forceinline__ int lower(i64 k, Pos pos)
    {
        int x = root, low = -1, r = 0;
        for (;;) {
            int i = find_in_node(x, k, &r, pos);
            if (i >= 0 && r == 0) return KEY(x, i);
            if (i >= 0) low = KEY(x, i);
            if (!INTERNAL(x)) return low;
            x = CHILD(x, i + 1);
        }
    }

    // __kb_split (kbtree.h:172-188)
    __device__ __forceinline__ void split(int x, int i, int y)
    {
        int z = new_node(INTERNAL(y));
        if (overflow) return;
        N(z) = KBT_T - 1;
        for (int j = 0; j < KBT_T - 1; ++j) KEY(z, j) = KEY(y, j + KBT_T);
        if (INTERNAL(y)) for (int j = 0; j < KBT_T; ++j) CHILD(z, j) = CHILD(y, j + KBT_T);
        N(y) = KBT_T - 1;
        int xn = N(x);
        for (int j = xn; j >= i + 1; --j) CHILD(x, j + 1) = CHILD(x, j);
        CHILD(x, i + 1) = z;
        for (int j = xn - 1; j >= i; --j) KEY(x, j + 1) = KEY(x, j);
        KEY(x, i) = KEY(y, KBT_T - 1);
        N(x) = xn + 1;
    }

    // kb_putp / __kb_putp_aux (kbtree.h:189-226)
    template <class Pos>
    __device__ __forceinline__ void put(int id, Pos pos)
    {
        i64 k = pos(id);
        ++n_keys;
        int r = root;
        if (N(r) == KBT_MAXK) {
            int s = new_node(1);
            if (overflow) return;
            root = s;
            CHILD(s, 0) = r;
            split(s, 0, r);
            if (overflow) return;
            r = s;
        }
        int x = r, dummy = 0;
        for (;;) {
            if (!INTERNAL(x)) {
                int i = find_in_node(x, k, &dummy, pos);
                int n = N(x);
                for (int j = n - 1; j >= i + 1; --j) KEY(x, j + 1) = KEY(x, j);
                KEY(x, i + 1) = id;
                N(x) = n + 1;
                return;
            }
            int i = find_in_node(x, k, &dummy, pos) + 1;
            if (N(CHILD(x, i)) == KBT_MAXK) {
                split(x, i, CHILD(x, i));
                if (overflow) return;
                if (k > pos(KEY(x, i))) ++i;
            }
            x = CHILD(x, i);
        }
    }

    // __kb_traverse (kbtree.h:336-358): in-order; returns the number of ids written
    __device__ __forceinline__ int traverse(int *out)
    {
        int sx[16], si[16], sp = 0, n = 0;
        sx[0] = root; si[0] = 0;
        for (;;) {
            // descend to the left-most unvisited leaf position
            while (INTERNAL(sx[sp]) && si[sp] <= N(sx[sp])) {
                int c = CHILD(sx[sp], si[sp]);
                ++sp; sx[sp] = c; si[sp] = 0;
            }
            if (!INTERNAL(sx[sp])) {
                for (int j = 0; j < N(sx[sp]); ++j) out[n++] = KEY(sx[sp], j);
            }
            // pop until a parent with a key to emit
            for (;;) {
                if (sp == 0) return n;
                --sp;
                if (si[sp] < N(sx[sp])) { out[n++] = KEY(sx[sp], si[sp]); ++si[sp]; break; }
                ++si[sp];
            }
        }
    }
};
