// Host-side index construction for synthetic / in-memory genomes (the workload side of bench.py and of
// the parity tests).  Produces exactly what the reference's `bwa index` produces for the same sequence
// (bwa_idx_build, bwtindex.c:253-323): forward 2-bit pac; BWT of forward+reverse-complement text with
// the sentinel removed and `primary` recorded (bwt_pac2bwt, bwtindex.c:63-126); Occ checkpoints every
// 128 bases interleaved into 64-byte blocks (bwt_bwtupdate_core, bwtindex.c:149-172); SA sampled every
// 32 rows with sa[0] = -1 (bwt_cal_sa, bwt.c:62-84).  The BWT and SA are mathematically determined, so
// any correct suffix sorter gives byte-identical arrays; ours is a plain SA-IS (induced sorting).
// Index construction stays on the host (north_star); nothing here runs on the hot path.
#include "../../include/bwamem_b200.h"
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

namespace {

template <class I>
struct Sais {
    // t: 1 = S-type, 0 = L-type
    static inline bool tget(const std::vector<uint8_t> &t, I i) { return t[(size_t)i >> 3] >> (i & 7) & 1; }
    static inline void tset(std::vector<uint8_t> &t, I i, bool b)
    {
        if (b) t[(size_t)i >> 3] |= (uint8_t)(1u << (i & 7)); else t[(size_t)i >> 3] &= (uint8_t)~(1u << (i & 7));
    }
    static inline bool is_lms(const std::vector<uint8_t> &t, I i) { return i > 0 && tget(t, i) && !tget(t, i - 1); }

    template <class S>
    static void buckets(const S *s, std::vector<I> &bkt, I n, I K, bool end)
    {
        std::fill(bkt.begin(), bkt.end(), 0);
        for (I i = 0; i < n; ++i) ++bkt[(size_t)s[i]];
        I sum = 0;
        for (I i = 0; i < K; ++i) { sum += bkt[i]; bkt[i] = end ? sum : sum - bkt[i]; }
    }
    template <class S>
    static void induce_l(const std::vector<uint8_t> &t, I *SA, const S *s, std::vector<I> &bkt, I n, I K)
    {
        buckets(s, bkt, n, K, false);
        for (I i = 0; i < n; ++i) {
            I j = SA[i] - 1;
            if (SA[i] > 0 && !tget(t, j)) SA[bkt[(size_t)s[j]]++] = j;
        }
    }
    template <class S>
    static void induce_s(const std::vector<uint8_t> &t, I *SA, const S *s, std::vector<I> &bkt, I n, I K)
    {
        buckets(s, bkt, n, K, true);
        for (I i = n - 1; i >= 0; --i) {
            I j = SA[i] - 1;
            if (SA[i] > 0 && tget(t, j)) SA[--bkt[(size_t)s[j]]] = j;
        }
    }

    // s[n-1] must be the unique smallest symbol (sentinel); symbols in [0, K)
    template <class S>
    static void run(const S *s, I *SA, I n, I K)
    {
        std::vector<uint8_t> t((size_t)n / 8 + 1, 0);
        tset(t, n - 1, true);
        if (n >= 2) tset(t, n - 2, false);
        for (I i = n - 3; i >= 0; --i)
            tset(t, i, s[i] < s[i + 1] || (s[i] == s[i + 1] && tget(t, i + 1)));
        std::vector<I> bkt((size_t)K);
        // stage 1: sort LMS substrings
        buckets(s, bkt, n, K, true);
        for (I i = 0; i < n; ++i) SA[i] = -1;
        for (I i = 1; i < n; ++i) if (is_lms(t, i)) SA[--bkt[(size_t)s[i]]] = i;
        induce_l(t, SA, s, bkt, n, K);
        induce_s(t, SA, s, bkt, n, K);
        I n1 = 0;
        for (I i = 0; i < n; ++i) if (is_lms(t, SA[i])) SA[n1++] = SA[i];
        for (I i = n1; i < n; ++i) SA[i] = -1;
        I name = 0, prev = -1;
        for (I i = 0; i < n1; ++i) {
            I pos = SA[i];
            bool diff = false;
            for (I d = 0; d < n; ++d) {
                if (prev == -1 || s[pos + d] != s[prev + d] || tget(t, pos + d) != tget(t, prev + d)) { diff = true; break; }
                else if (d > 0 && (is_lms(t, pos + d) || is_lms(t, prev + d))) break;
            }
            if (diff) { ++name; prev = pos; }
            SA[n1 + pos / 2] = name - 1;
        }
        for (I i = n - 1, j = n - 1; i >= n1; --i) if (SA[i] >= 0) SA[j--] = SA[i];
        // stage 2: solve the reduced problem
        I *SA1 = SA, *s1 = SA + n - n1;
        if (name < n1) run<I>(s1, SA1, n1, name);
        else for (I i = 0; i < n1; ++i) SA1[s1[i]] = i;
        // stage 3: induce the result
        buckets(s, bkt, n, K, true);
        for (I i = 1, j = 0; i < n; ++i) if (is_lms(t, i)) s1[j++] = i;
        for (I i = 0; i < n1; ++i) SA1[i] = s1[SA1[i]];
        for (I i = n1; i < n; ++i) SA[i] = -1;
        for (I i = n1 - 1; i >= 0; --i) {
            I j = SA[i];
            SA[i] = -1;
            SA[--bkt[(size_t)s[j]]] = j;
        }
        induce_l(t, SA, s, bkt, n, K);
        induce_s(t, SA, s, bkt, n, K);
    }
};

}  // namespace

struct b200_hostidx {
    uint64_t primary, L2[5], seq_len, bwt_size, n_sa;
    int sa_intv;
    int64_t l_pac;
    std::vector<uint32_t> bwt;
    std::vector<uint64_t> sa;
    std::vector<uint8_t> pac;
};

template <class I>
static void build_from_text(b200_hostidx *h, std::vector<uint8_t> &text /* seq_len + 1, symbols 1..4, sentinel 0 */)
{
    const uint64_t n = h->seq_len;
    std::vector<I> SA(n + 1);
    Sais<I>::template run<uint8_t>(text.data(), SA.data(), (I)(n + 1), (I)5);
    // BWT with the sentinel row removed (is_bwt, is.c:209-223), packed 16 bases per word
    std::vector<uint32_t> raw((n + 15) >> 4, 0);
    uint64_t primary = 0, out = 0;
    for (uint64_t i = 0; i <= n; ++i) {
        if (SA[i] == 0) { primary = i; continue; }
        uint32_t c = (uint32_t)text[(size_t)SA[i] - 1] - 1;
        raw[out >> 4] |= c << ((15 - (out & 15)) << 1);
        ++out;
    }
    h->primary = primary;
    // sampled SA (bwt_cal_sa): row k % 32 == 0 -> sa[k/32]; sa[0] = -1
    h->sa_intv = 32;
    h->n_sa = (n + 32) / 32;
    h->sa.assign(h->n_sa, 0);
    for (uint64_t k = 0; k <= n; k += 32) h->sa[k / 32] = (uint64_t)SA[k];
    h->sa[0] = (uint64_t)-1;
    std::vector<I>().swap(SA);
    // interleave Occ checkpoints (bwt_bwtupdate_core)
    const uint64_t n_occ = (n + 127) / 128 + 1;
    h->bwt_size = ((n + 15) >> 4) + n_occ * 8;
    h->bwt.assign(h->bwt_size, 0);
    uint64_t c[4] = {0, 0, 0, 0}, k = 0;
    for (uint64_t i = 0; i < n; ++i) {
        if ((i & 127) == 0) { memcpy(&h->bwt[k], c, 32); k += 8; }
        if ((i & 15) == 0) h->bwt[k++] = raw[i >> 4];
        ++c[raw[i >> 4] >> ((~i & 15) << 1) & 3];
    }
    memcpy(&h->bwt[k], c, 32);
}

extern "C" {

// fwd: nt4 codes (0..3) of the concatenated contigs, l_pac bases.  Codes > 3 are not accepted here (the
// reference replaces N by a random base while packing, bntseq.c:295-296; callers do that beforehand).
b200_hostidx *b200_hostidx_build(const uint8_t *fwd, int64_t l_pac)
{
    if (l_pac <= 0) return nullptr;
    for (int64_t i = 0; i < l_pac; ++i) if (fwd[i] > 3) return nullptr;
    b200_hostidx *h = new b200_hostidx();
    h->l_pac = l_pac;
    h->seq_len = (uint64_t)l_pac * 2;
    h->pac.assign((size_t)(l_pac / 4 + 1), 0);
    for (int64_t i = 0; i < l_pac; ++i) h->pac[i >> 2] |= (uint8_t)(fwd[i] << ((~i & 3) << 1));
    std::vector<uint8_t> text(h->seq_len + 1);
    memset(h->L2, 0, sizeof(h->L2));
    for (int64_t i = 0; i < l_pac; ++i) {
        text[i] = fwd[i] + 1;
        text[h->seq_len - 1 - i] = (3 - fwd[i]) + 1;
    }
    text[h->seq_len] = 0;
    for (uint64_t i = 0; i < h->seq_len; ++i) ++h->L2[text[i]];     // L2[1 + base]
    for (int i = 2; i <= 4; ++i) h->L2[i] += h->L2[i - 1];
    if (h->seq_len + 1 < 0x7fffffffull) build_from_text<int32_t>(h, text);
    else build_from_text<int64_t>(h, text);
    return h;
}

void b200_hostidx_free(b200_hostidx *h) { delete h; }

// out[0]=primary [1..5]=L2 [6]=seq_len [7]=bwt_size [8]=n_sa [9]=sa_intv [10]=l_pac
void b200_hostidx_info(const b200_hostidx *h, uint64_t *out)
{
    out[0] = h->primary;
    for (int i = 0; i < 5; ++i) out[1 + i] = h->L2[i];
    out[6] = h->seq_len; out[7] = h->bwt_size; out[8] = h->n_sa; out[9] = (uint64_t)h->sa_intv; out[10] = (uint64_t)h->l_pac;
}
const uint32_t *b200_hostidx_bwt(const b200_hostidx *h) { return h->bwt.data(); }
const uint64_t *b200_hostidx_sa(const b200_hostidx *h) { return h->sa.data(); }
const uint8_t *b200_hostidx_pac(const b200_hostidx *h) { return h->pac.data(); }

int b200_index_upload_host(b200_ctx_t *ctx, const b200_hostidx *h, int n_seqs, const int64_t *ann_offset,
                           const int32_t *ann_len, const int32_t *ann_is_alt, int flags)
{
    return b200_index_upload_raw(ctx, h->bwt.data(), h->bwt_size, h->primary, h->L2, h->seq_len, h->sa.data(), h->n_sa,
                                 h->sa_intv, h->pac.data(), h->l_pac, n_seqs, ann_offset, ann_len, ann_is_alt, flags);
}

}  // extern "C"
