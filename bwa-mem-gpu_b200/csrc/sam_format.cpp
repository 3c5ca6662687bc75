// SAM text of single-end reads from the arrays b200_reg2sam_se_batch returns: mem_aln2sam (bwamem.c:832-956) with
// add_cigar (:819-830), the unmapped record of mem_reg2sam (:1043-1047) and the XA strings of mem_gen_alt
// (bwamem_extra.c:143-161).  Host code: this is string formatting, done by a pool of threads over blocks of reads;
// every number in the text was computed on the device.  Two passes over the same templated formatter: the first only
// counts each read's bytes (no characters are produced), a prefix sum places the reads, the second writes every read
// straight into its place in the caller's buffer -- no intermediate strings, no concatenation.
#include <stdio.h>
#include <string.h>
#include <string>
#include <thread>
#include <vector>
#include <algorithm>
#include <functional>
#include <stdint.h>
#include "../../include/bwamem_b200.h"

#define SF_SOFTCLIP 0x200      /* MEM_F_SOFTCLIP, bwa.h:80 */
#define SF_REF_HDR 0x100       /* MEM_F_REF_HDR,  bwa.h:79 */
#define SF_XB 0x2000           /* MEM_F_XB,       bwa.h:84 */

namespace {

struct CountSink {                                       // pass 1: lengths only
    size_t l = 0;
    void push_back(char) { ++l; }
    CountSink &operator+=(const char *z) { l += strlen(z); return *this; }
    CountSink &operator+=(const std::string &z) { l += z.size(); return *this; }
    void append(const char *, size_t n) { l += n; }
    char *grow(size_t m) { l += m; return nullptr; }
    size_t size() const { return l; }
    void tabs_to_spaces(size_t) {}
};
struct WriteSink {                                       // pass 2: into the caller's buffer
    char *p; size_t l = 0;
    explicit WriteSink(char *q) : p(q) {}
    void push_back(char c) { p[l++] = c; }
    WriteSink &operator+=(const char *z) { const size_t n = strlen(z); memcpy(p + l, z, n); l += n; return *this; }
    WriteSink &operator+=(const std::string &z) { memcpy(p + l, z.data(), z.size()); l += z.size(); return *this; }
    void append(const char *z, size_t n) { memcpy(p + l, z, n); l += n; }
    char *grow(size_t m) { char *d = p + l; l += m; return d; }
    size_t size() const { return l; }
    void tabs_to_spaces(size_t from) { for (size_t i = from; i < l; ++i) if (p[i] == '\t') p[i] = ' '; }
};
struct StrSink {                                         // XA strings
    std::string s;
    void push_back(char c) { s.push_back(c); }
    StrSink &operator+=(const char *z) { s += z; return *this; }
};

template <class S>
inline void put_int(S &s, long long v)                  // kputw / kputl
{
    char buf[24]; int n = 0;
    unsigned long long x = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
    do { buf[n++] = (char)('0' + x % 10); x /= 10; } while (x);
    if (v < 0) s.push_back('-');
    while (n) s.push_back(buf[--n]);
}

inline int code_of(uint8_t c)                            // reads may be given as nt4 codes or as letters
{
    if (c < 4) return c;
    switch (c & 0xdf) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; }
    return 4;
}

struct Rec { const b200_aln_t *a; const uint32_t *cig; const char *md; int mapq; };

template <class S>
void put_cigar(S &s, const mem_opt_t *opt, const Rec &p, int which)        // add_cigar
{
    if (!p.a->n_cigar) { s.push_back('*'); return; }
    for (int i = 0; i < p.a->n_cigar; ++i) {
        int c = (int)(p.cig[i] & 0xf);
        if (!(opt->flag & SF_SOFTCLIP) && !p.a->is_alt && (c == 3 || c == 4)) c = which ? 4 : 3;
        put_int(s, p.cig[i] >> 4); s.push_back("MIDSH"[c]);
    }
}

struct Job {
    const mem_opt_t *opt; const char *const *ann_names; const char *const *ann_annos;
    const char *const *names; const char *const *comments; const uint8_t *seqs; const uint8_t *quals; const int64_t *off;
    const int64_t *roff; const b200_sel_t *sel; const b200_aln_t *aln; const uint32_t *cigar; const char *md; const char *rg;
};

template <class S>
void one_record(S &s, const Job &J, int r, const std::vector<Rec> &list, int which, const std::string *xa)
{
    const mem_opt_t *opt = J.opt;
    static const b200_aln_t unmapped = { -1, -1, 0x4, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0 };
    const Rec none = { &unmapped, nullptr, nullptr, 0 };
    const Rec &p = list.empty() ? none : list[which];
    const b200_aln_t &a = *p.a;
    const int l_seq = (int)(J.off[r + 1] - J.off[r]);
    const uint8_t *seq = J.seqs + J.off[r], *qual = J.quals ? J.quals + J.off[r] : nullptr;
    int flag = a.flag;
    if (a.rid < 0) flag |= 0x4;
    if (a.is_rev) flag |= 0x10;
    s += J.names[r]; s.push_back('\t');
    put_int(s, (flag & 0xffff) | (flag & 0x10000 ? 0x100 : 0)); s.push_back('\t');
    if (a.rid >= 0) {
        s += J.ann_names[a.rid]; s.push_back('\t');
        put_int(s, a.pos + 1); s.push_back('\t');
        put_int(s, p.mapq); s.push_back('\t');
        put_cigar(s, opt, p, which);
    } else s += "*\t0\t0\t*";
    s.push_back('\t');
    s += "*\t0\t0";                                      // single-end: no mate
    s.push_back('\t');
    if (flag & 0x100) s += "*\t*";
    else {
        int qb = 0, qe = l_seq;
        if (a.n_cigar && which && !(opt->flag & SF_SOFTCLIP) && !a.is_alt) {
            const int c0 = (int)(p.cig[0] & 0xf), c1 = (int)(p.cig[a.n_cigar - 1] & 0xf);
            if (!a.is_rev) {
                if (c0 == 4 || c0 == 3) qb += (int)(p.cig[0] >> 4);
                if (c1 == 4 || c1 == 3) qe -= (int)(p.cig[a.n_cigar - 1] >> 4);
            } else {
                if (c0 == 4 || c0 == 3) qe -= (int)(p.cig[0] >> 4);
                if (c1 == 4 || c1 == 3) qb += (int)(p.cig[a.n_cigar - 1] >> 4);
            }
        }
        if (!a.is_rev) {
            const int m = std::max(0, qe - qb);
            char *d = s.grow((size_t)m);
            if (d) for (int i = 0; i < m; ++i) d[i] = "ACGTN"[code_of(seq[qb + i])];
            s.push_back('\t');
            if (qual) s.append((const char *)qual + qb, (size_t)m); else s.push_back('*');
        } else {
            const int m = std::max(0, qe - qb);
            char *d = s.grow((size_t)m);
            if (d) for (int i = 0; i < m; ++i) d[i] = "TGCAN"[code_of(seq[qe - 1 - i])];
            s.push_back('\t');
            if (qual) {
                char *e = s.grow((size_t)m);
                if (e) for (int i = 0; i < m; ++i) e[i] = (char)qual[qe - 1 - i];
            } else s.push_back('*');
        }
    }
    if (a.n_cigar) {
        s += "\tNM:i:"; put_int(s, a.NM);
        s += "\tMD:Z:"; s.append(p.md, (size_t)a.l_md);
    }
    if (a.score >= 0) { s += "\tAS:i:"; put_int(s, a.score); }
    if (a.sub >= 0) { s += "\tXS:i:"; put_int(s, a.sub); }
    if (J.rg && J.rg[0]) { s += "\tRG:Z:"; s += J.rg; }
    if (!(flag & 0x100)) {
        const int n = (int)list.size();
        int i;
        for (i = 0; i < n; ++i) if (i != which && !(list[i].a->flag & 0x100)) break;
        if (i < n) {
            s += "\tSA:Z:";
            for (i = 0; i < n; ++i) {
                const Rec &q = list[i];
                if (i == which || (q.a->flag & 0x100)) continue;
                s += J.ann_names[q.a->rid]; s.push_back(',');
                put_int(s, q.a->pos + 1); s.push_back(',');
                s.push_back("+-"[q.a->is_rev]); s.push_back(',');
                for (int k = 0; k < q.a->n_cigar; ++k) { put_int(s, q.cig[k] >> 4); s.push_back("MIDSH"[q.cig[k] & 0xf]); }
                s.push_back(','); put_int(s, q.mapq);
                s.push_back(','); put_int(s, q.a->NM);
                s.push_back(';');
            }
        }
        if (a.alt_sc > 0) { char buf[48]; snprintf(buf, sizeof(buf), "\tpa:f:%.3f", (double)a.score / a.alt_sc); s += buf; }
    }
    if (xa && !xa->empty()) { s += (opt->flag & SF_XB) ? "\tXB:Z:" : "\tXA:Z:"; s += *xa; }
    if (J.comments && J.comments[r]) { s.push_back('\t'); s += J.comments[r]; }
    if ((opt->flag & SF_REF_HDR) && a.rid >= 0 && J.ann_annos && J.ann_annos[a.rid] && J.ann_annos[a.rid][0]) {
        s += "\tXR:Z:";
        const size_t t = s.size();
        s += J.ann_annos[a.rid];
        s.tabs_to_spaces(t);
    }
    s.push_back('\n');
}

struct Scratch { std::vector<Rec> list; std::vector<int> reg_of; std::vector<StrSink> xa; };      // reused over a thread's reads

template <class S>
void one_read(S &s, const Job &J, int r, Scratch &W)
{
    const int64_t g0 = J.roff[r];
    const int n = (int)(J.roff[r + 1] - g0);
    std::vector<Rec> &list = W.list;
    std::vector<int> &reg_of = W.reg_of;
    list.clear(); reg_of.clear();
    for (int k = 0; k < n; ++k)
        if (J.sel[g0 + k].kind & 1) {
            const b200_aln_t *a = J.aln + g0 + k;
            list.push_back(Rec{ a, J.cigar + a->cigar_off, J.md + a->md_off, J.sel[g0 + k].mapq });
            reg_of.push_back(k);
        }
    if (list.empty()) { one_record(s, J, r, list, 0, nullptr); return; }
    if (n == 1) { one_record(s, J, r, list, 0, nullptr); return; }      // one region, one record: no XA, no SA
    std::vector<StrSink> &xa = W.xa;
    if (xa.size() < list.size()) xa.resize(list.size());
    for (size_t w = 0; w < list.size(); ++w) xa[w].s.clear();
    for (int i = 0; i < n; ++i) {                        // mem_gen_alt: entries appended in region order
        const b200_sel_t &e = J.sel[g0 + i];
        if (!(e.kind & 2)) continue;
        const auto it = std::find(reg_of.begin(), reg_of.end(), e.xa_of);
        if (it == reg_of.end()) continue;                // its record is not printed
        StrSink &x = xa[(size_t)(it - reg_of.begin())];
        const b200_aln_t &t = J.aln[g0 + i];
        const uint32_t *cg = J.cigar + t.cigar_off;
        x += J.ann_names[t.rid]; x.push_back(','); x.push_back("+-"[t.is_rev]); put_int(x, t.pos + 1); x.push_back(',');
        for (int k = 0; k < t.n_cigar; ++k) { put_int(x, cg[k] >> 4); x.push_back("MIDSHN"[cg[k] & 0xf]); }
        x.push_back(','); put_int(x, t.NM);
        if (J.opt->flag & SF_XB) { x.push_back(','); put_int(x, t.score); }
        x.push_back(';');
    }
    for (int w = 0; w < (int)list.size(); ++w) one_record(s, J, r, list, w, &xa[w].s);
}

}  // namespace

extern "C" void *b200_host_alloc(size_t bytes);
extern "C" void b200_host_free(void *p);

extern "C" int64_t b200_sam_se_format(const mem_opt_t *opt, const char *const *ann_names, const char *const *ann_annos, int n_reads,
                                      const char *const *names, const char *const *comments, const uint8_t *seqs,
                                      const uint8_t *quals, const int64_t *off, const int32_t *cnt, const b200_sel_t *sel,
                                      const b200_aln_t *aln, const uint32_t *cigar, const char *md, const char *rg_id,
                                      int n_threads, int64_t cap, char *sam, int64_t *sam_off)
{
    std::vector<int64_t> roff((size_t)n_reads + 1, 0);
    for (int r = 0; r < n_reads; ++r) roff[r + 1] = roff[r] + cnt[r];
    const Job J = { opt, ann_names, ann_annos, names, comments, seqs, quals, off, roff.data(), sel, aln, cigar, md, rg_id };
    const int nt = std::max(1, std::min(n_threads, std::max(1, n_reads / 256)));
    std::vector<int64_t> own;
    int64_t *pos = sam_off;                                // n_reads + 1 offsets of the reads' records
    if (!pos) { own.resize((size_t)n_reads + 1); pos = own.data(); }
    auto run = [&](const std::function<void(int, int, int)> &fn) {
        if (nt == 1) { fn(0, 0, n_reads); return; }
        std::vector<std::thread> th;
        for (int t = 0; t < nt; ++t)
            th.emplace_back(fn, t, (int)((int64_t)n_reads * t / nt), (int)((int64_t)n_reads * (t + 1) / nt));
        for (auto &x : th) x.join();
    };
    run([&](int, int lo, int hi) {                         // pass 1: pos[r + 1] = bytes of read r
        Scratch W;
        for (int r = lo; r < hi; ++r) { CountSink c; one_read(c, J, r, W); pos[r + 1] = (int64_t)c.l; }
    });
    pos[0] = 0;
    for (int r = 0; r < n_reads; ++r) pos[r + 1] += pos[r];
    const int64_t total = pos[n_reads];
    if (total > cap) return -total;
    std::vector<int> bad((size_t)nt, 0);
    run([&](int t, int lo, int hi) {                       // pass 2: every read written in place
        Scratch W;
        for (int r = lo; r < hi; ++r) {
            WriteSink w(sam + pos[r]);
            one_read(w, J, r, W);
            if ((int64_t)w.l != pos[r + 1] - pos[r]) bad[t] = 1;
        }
    });
    for (int b : bad) if (b) return INT64_MIN;             // the two passes disagree: a bug, never silent
    return total;
}

// ---- reads in host memory -> SAM text in host memory, single-end, as one call ------------------------------------
// The batch is cut into sub-batches; for each: b200_align_batch (worker1; only region counts come back),
// b200_batch_reg2sam_se (worker2 on the resident regions), then b200_sam_se_format on host threads -- which runs while
// the device is already busy with the next sub-batch (two sets of page-locked record buffers).
extern "C" void *b200_ctx_pinned(b200_ctx_t *ctx, int slot, size_t bytes);

extern "C" int64_t b200_align_sam_se_batch(b200_ctx_t *ctx, const mem_opt_t *opt, const char *const *ann_names, const char *const *ann_annos,
                                           int n_reads, int64_t n_processed, const char *const *names, const char *const *comments,
                                           const uint8_t *seqs, const uint8_t *quals, const int64_t *off, const char *rg_id, int n_threads,
                                           int sub_reads, int64_t cap, char *sam, int64_t *sam_off)
{
    const int sub = sub_reads > 0 ? sub_reads : (1 << 20);
    std::thread fmt;
    int64_t written = 0, need = 0;                         // bytes placed so far / bytes the whole text needs
    bool fmt_failed = false;
    auto join = [&]() { if (fmt.joinable()) fmt.join(); };
    if (sam_off) sam_off[0] = 0;
    for (int r0 = 0, k = 0; r0 < n_reads; r0 += sub, ++k) {
        const int n = std::min(sub, n_reads - r0);
        const int s0 = (k & 1) * 5;                        // buffer set; the formatter of sub-batch k-2 was joined before k-1's started
        int32_t *cnt = (int32_t *)b200_ctx_pinned(ctx, s0, (size_t)n * 4);
        if (!cnt) { join(); return B200_E_FAIL; }
        int64_t nr = b200_align_batch(ctx, opt, n, seqs, off + r0, 0, nullptr, cnt);
        if (B200_IS_ERROR(nr)) { join(); return nr; }      // a real error (message set), not "-(regions)"
        if (nr > 0) { join(); return B200_E_FAIL; }        // cannot happen with cap 0
        nr = -nr;
        if (nr >= 0x7fffffff) { join(); return B200_E_FAIL; }
        b200_sel_t *sel = (b200_sel_t *)b200_ctx_pinned(ctx, s0 + 1, (size_t)(nr + 1) * sizeof(b200_sel_t));
        b200_aln_t *aln = (b200_aln_t *)b200_ctx_pinned(ctx, s0 + 2, (size_t)(nr + 1) * sizeof(b200_aln_t));
        const int64_t bases = off[r0 + n] - off[r0];             // indel-rich long reads: CIGAR / MD grow with the bases, not the regions
        int64_t cap_c = 4 * nr + bases / 8 + 64, cap_m = 16 * nr + bases / 4 + 64, totals[2];
        uint32_t *cig = nullptr; char *md = nullptr;
        if (!sel || !aln) { join(); return B200_E_FAIL; }
        for (;;) {
            cig = (uint32_t *)b200_ctx_pinned(ctx, s0 + 3, (size_t)cap_c * 4);
            md = (char *)b200_ctx_pinned(ctx, s0 + 4, (size_t)cap_m);
            if (!cig || !md) { join(); return B200_E_FAIL; }
            const int64_t rc = nr ? b200_batch_reg2sam_se(ctx, opt, n_processed + r0, cnt, nullptr, sel, aln, cap_c, cig, cap_m, md, totals) : 0;
            if (rc == -2) { cap_c = totals[0] + 64; cap_m = totals[1] + 64; continue; }
            if (rc < 0) { join(); return B200_E_FAIL; }
            break;
        }
        join();                                            // the previous sub-batch's text is in place: `written` is final
        if (fmt_failed) return B200_E_FAIL;
        fmt = std::thread([&, r0, n, cnt, sel, aln, cig, md]() {
            int64_t *so = sam_off ? sam_off + r0 : nullptr;
            const int64_t room = need > cap ? 0 : cap - written;
            const int64_t rc = b200_sam_se_format(opt, ann_names, ann_annos, n, names + r0, comments ? comments + r0 : nullptr, seqs, quals,
                                                  off + r0, cnt, sel, aln, cig, md, rg_id, n_threads, room, sam + written, so);
            if (rc == INT64_MIN) { fmt_failed = true; return; }
            const int64_t bytes = rc < 0 ? -rc : rc;
            if (so) for (int i = 0; i <= n; ++i) so[i] += need;      // offsets of the whole text (so[n] is the next sub-batch's so[0])
            need += bytes;
            if (rc >= 0 && need <= cap) written += bytes;
        });
    }
    join();
    if (fmt_failed) { return B200_E_FAIL; }
    return need > cap ? -need : need;
}
