// Drop-in mem_process_seqs (bwamem.h:69, bwamem.c:1215-1244).  By default the whole batch -- worker1 (mem_align1_core for
// every read, bwamem.c:1183-1195), the batch's mem_pestat (bwamem_pair.c:72), mate rescue (:291-302), worker2 (mem_mark_primary_se,
// mem_pair, the MAPQ arithmetic and flags of mem_sam_pe, mem_reg2aln, bwamem.c:1197-1213) and the SAM text (mem_aln2sam) -- runs
// on the device(s) through b200_reads_to_sam (csrc/reads2sam.cu); no function of the host program is used.
// b200_process_seqs_host_worker2(1) selects the older route, kept as a cross-check: worker1 and mate rescue on the device, then the
// REFERENCE's own host functions
// (mem_pestat, mem_mark_primary_se / mem_reg2sam / mem_sam_pe with MEM_F_NO_RESCUE), which this library does not contain: those
// symbols are resolved from the host program the library is linked into (weak references, or the global symbol scope).
// Nothing here computes alignments on the CPU; if that route is selected and the host functions are missing the call fails loudly.
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <thread>
#include <vector>
#include <atomic>
#include <functional>
#include <string>
#include "../../include/bwamem_b200.h"

#define MEM_F_PRIMARY5_ 0x800        /* bwamem.h:47 */
#define MEM_F_NO_RESCUE_ 0x20        /* bwa.h:78 */

extern "C" {
int  mem_mark_primary_se(const mem_opt_t *opt, int n, mem_alnreg_t *a, int64_t id) __attribute__((weak));
void mem_reg2sam(const mem_opt_t *opt, const bntseq_t *bns, const uint8_t *pac, bseq1_t *s, mem_alnreg_v *a,
                 int extra_flag, const void *m) __attribute__((weak));
int  mem_sam_pe(const mem_opt_t *opt, const bntseq_t *bns, const uint8_t *pac, const mem_pestat_t pes[4], uint64_t id,
                bseq1_t s[2], mem_alnreg_v a[2]) __attribute__((weak));
void mem_pestat(const mem_opt_t *opt, int64_t l_pac, int n, const mem_alnreg_v *regs, mem_pestat_t pes[4]) __attribute__((weak));
void mem_reorder_primary5(int T, mem_alnreg_v *a) __attribute__((weak));
}

typedef int  (*mark_fn)(const mem_opt_t *, int, mem_alnreg_t *, int64_t);
typedef void (*reg2sam_fn)(const mem_opt_t *, const bntseq_t *, const uint8_t *, bseq1_t *, mem_alnreg_v *, int, const void *);
typedef int  (*sam_pe_fn)(const mem_opt_t *, const bntseq_t *, const uint8_t *, const mem_pestat_t *, uint64_t, bseq1_t *, mem_alnreg_v *);
typedef void (*reorder5_fn)(int, mem_alnreg_v *);
typedef void (*pestat_fn)(const mem_opt_t *, int64_t, int, const mem_alnreg_v *, mem_pestat_t *);

static void *host_symbol(void *weak_addr, const char *name)
{
    if (weak_addr) return weak_addr;
    void *p = dlsym(RTLD_DEFAULT, name);
    if (!p) {
        fprintf(stderr, "[b200::mem_process_seqs] host function %s not found: link this library into the reference's "
                        "host code (see INTEGRATION.md); there is no CPU path in libbwamem_b200\n", name);
        exit(1);
    }
    return p;
}

// nst_nt4_table (bntseq.c:40-57) restricted to what mem_align1_core needs (bwamem.c:1067-1068)
static inline uint8_t nt4_of(uint8_t c)
{
    if (c < 4) return c;
    switch (c & 0xdf) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; }
    return c == '-' ? 5 : 4;
}

// One context per device; a batch is cut into contiguous blocks of reads (whole pairs), one per context, each driven
// by its own host thread.  Nothing is exchanged between devices: the only batch-level quantity, mem_pestat, is computed
// on the host over all blocks (as the reference does over all reads) and handed to every block's rescue stage.
static std::vector<b200_ctx_t *> g_ps_ctxs;
static const bwt_t *g_ps_bwt = nullptr;
static int g_ps_want_devices = 1;
// worker2: 0 or -1 (default) = on the device -- long regions too since reg2aln_warp_kernel (one region per warp;
// profiles/r5d_long_sam.jsonl: 50 000 reads of 1-10 kbp in 0.8 s, was 36 s with one region per lane); 1 = the host
// program's functions (kept as a cross-check)
static int g_ps_host_worker2 = -1;
extern "C" void b200_process_seqs_host_worker2(int on) { g_ps_host_worker2 = on; }
// read group id for the RG:Z tag (the host program's bwa_rg_id, bwa.c:44); unset: looked up as a dynamic symbol of the host program
static const char *g_ps_rg = nullptr;
extern "C" void b200_process_seqs_set_rg_id(const char *rg_id) { g_ps_rg = rg_id; }

extern "C" void b200_process_seqs_set_ctx(b200_ctx_t *ctx) { g_ps_ctxs.assign(1, ctx); g_ps_bwt = nullptr; }
extern "C" void b200_process_seqs_set_ctxs(int n, b200_ctx_t **ctxs) { g_ps_ctxs.assign(ctxs, ctxs + n); g_ps_bwt = nullptr; }
extern "C" void b200_process_seqs_use_devices(int n_devices) { g_ps_want_devices = n_devices > 0 ? n_devices : 1; }

struct Shard {
    b200_ctx_t *ctx; int r0, n;
    std::vector<mem_alnreg_t> regs; std::vector<int32_t> cnt;
    std::string err;
    std::vector<char> sam; std::vector<int64_t> sam_off;
};

static void shard_download(Shard &sh)
{
    for (;;) {
        int64_t got = b200_batch_download(sh.ctx, (int64_t)sh.regs.size(), sh.regs.data(), sh.cnt.data());
        if (got >= 0) return;
        if (!B200_IS_ERROR(got) && -got > (int64_t)sh.regs.size()) { sh.regs.resize((size_t)-got); continue; }
        sh.err = b200_last_error();
        return;
    }
}

static void run_shards(std::vector<Shard> &shards, const std::function<void(Shard &)> &fn)
{
    std::vector<std::thread> th;
    for (size_t d = 1; d < shards.size(); ++d) th.emplace_back([&, d]() { fn(shards[d]); });
    fn(shards[0]);
    for (auto &t : th) t.join();
    for (auto &sh : shards)
        if (!sh.err.empty()) { fprintf(stderr, "[b200::mem_process_seqs] %s\n", sh.err.c_str()); exit(1); }
}

extern "C" void b200_mem_process_seqs(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac,
                                      int64_t n_processed, int n, bseq1_t *seqs, const mem_pestat_t *pes0)
{
    const bool pe = (opt->flag & MEM_F_PE) != 0;
    const bool dev_worker2 = g_ps_host_worker2 <= 0;       // everything on the device(s) unless the host program's worker2 was asked for explicitly
    mark_fn f_mark = dev_worker2 ? nullptr : (mark_fn)host_symbol((void *)mem_mark_primary_se, "mem_mark_primary_se");
    reg2sam_fn f_reg2sam = dev_worker2 ? nullptr : (reg2sam_fn)host_symbol((void *)mem_reg2sam, "mem_reg2sam");
    sam_pe_fn f_sam_pe = (pe && !dev_worker2) ? (sam_pe_fn)host_symbol((void *)mem_sam_pe, "mem_sam_pe") : nullptr;
    // -5: single-end batches reorder between marking and mem_reg2sam (bwamem.c:1205, mem_reorder_primary5 bwamem.c:988 is a
    // non-static function of the host program); mem_sam_pe applies it itself (bwamem_pair.c:305-307)
    reorder5_fn f_reorder5 = (!pe && !dev_worker2 && (opt->flag & MEM_F_PRIMARY5_))
                                 ? (reorder5_fn)host_symbol((void *)mem_reorder_primary5, "mem_reorder_primary5") : nullptr;
    pestat_fn f_pestat = (pe && !pes0 && !dev_worker2) ? (pestat_fn)host_symbol((void *)mem_pestat, "mem_pestat") : nullptr;

    if (g_ps_ctxs.empty() || (g_ps_bwt && g_ps_bwt != bwt)) {         // lazy contexts from the index identity
        g_ps_ctxs.clear();
        for (int d = 0; d < g_ps_want_devices; ++d) {
            b200_ctx_t *c = b200_ctx_create(d);
            if (!c || b200_index_upload(c, bwt, bns, pac, B200_IDX_FULL_SA)) {
                fprintf(stderr, "[b200::mem_process_seqs] %s\n", b200_last_error());
                exit(1);
            }
            g_ps_ctxs.push_back(c);
        }
        g_ps_bwt = bwt;
    }

    // ---- worker1 on the device(s)
    std::vector<int64_t> off((size_t)n + 1);
    int64_t tot = 0;
    for (int i = 0; i < n; ++i) { off[i] = tot; tot += seqs[i].l_seq; }
    off[n] = tot;
    std::vector<uint8_t> flat((size_t)tot + 1);
    for (int i = 0; i < n; ++i) memcpy(flat.data() + off[i], seqs[i].seq, (size_t)seqs[i].l_seq);
    const int G = (int)g_ps_ctxs.size();
    std::vector<Shard> shards((size_t)G);
    for (int d = 0; d < G; ++d) {
        int lo = (int)((int64_t)n * d / G), hi = (int)((int64_t)n * (d + 1) / G);
        if (pe) { lo &= ~1; hi = d == G - 1 ? n : hi & ~1; }          // mates stay together
        shards[d].ctx = g_ps_ctxs[d]; shards[d].r0 = lo; shards[d].n = hi - lo;
    }
    if (dev_worker2) {                                                // worker1, mem_pestat, worker2 and the SAM text (bwamem.c:1183-1244) on the device(s)
        int n_qual = 0;
        for (int i = 0; i < n; ++i) n_qual += seqs[i].qual != nullptr;
        if (n_qual && n_qual != n) { fprintf(stderr, "[b200::mem_process_seqs] reads with and without qualities in one batch\n"); exit(1); }
        std::vector<uint8_t> qflat(n_qual ? (size_t)tot + 1 : 0);
        std::vector<int64_t> name_off((size_t)n + 1, 0), com_off((size_t)n + 1, 0);
        bool any_comment = false;
        for (int i = 0; i < n; ++i) {
            name_off[i + 1] = name_off[i] + (int64_t)strlen(seqs[i].name);
            const size_t lc = seqs[i].comment ? strlen(seqs[i].comment) : 0;
            com_off[i + 1] = com_off[i] + (int64_t)lc;
            any_comment |= lc != 0;
        }
        std::vector<char> nflat((size_t)name_off[n] + 1), cflat((size_t)com_off[n] + 1);
        for (int i = 0; i < n; ++i) {
            if (n_qual) memcpy(qflat.data() + off[i], seqs[i].qual, (size_t)seqs[i].l_seq);
            memcpy(nflat.data() + name_off[i], seqs[i].name, (size_t)(name_off[i + 1] - name_off[i]));
            if (com_off[i + 1] > com_off[i]) memcpy(cflat.data() + com_off[i], seqs[i].comment, (size_t)(com_off[i + 1] - com_off[i]));
        }
        std::vector<const char *> ann_names((size_t)bns->n_seqs + 1), ann_annos((size_t)bns->n_seqs + 1);
        for (int i = 0; i < bns->n_seqs; ++i) { ann_names[i] = bns->anns[i].name; ann_annos[i] = bns->anns[i].anno; }
        const char *rg = g_ps_rg ? g_ps_rg : (const char *)dlsym(RTLD_DEFAULT, "bwa_rg_id");      // bwa.c:44, set by -R; absent = no RG tag
        b200_reads_t R;
        R.n_reads = n; R.seq = flat.data(); R.seq_off = off.data(); R.qual = n_qual ? qflat.data() : nullptr;
        R.name = nflat.data(); R.name_off = name_off.data();
        R.comment = any_comment ? cflat.data() : nullptr; R.comment_off = any_comment ? com_off.data() : nullptr;
        std::vector<int64_t> sam_off((size_t)n + 1, 0);
        std::vector<char> sam((size_t)(2 * tot + (int64_t)n * 200 + name_off[n] + com_off[n] + 4096));
        for (;;) {
            const int64_t rc = b200_reads_to_sam(g_ps_ctxs.data(), G, opt, ann_names.data(), ann_annos.data(), &R, n_processed, pes0, rg, 0,
                                                 (int64_t)sam.size(), sam.data(), sam_off.data(), nullptr);
            if (B200_IS_ERROR(rc)) { fprintf(stderr, "[b200::mem_process_seqs] %s\n", b200_last_error()); exit(1); }      // err_fatal style (bwamem_pair.c:386)
            if (rc < 0) { sam.resize((size_t)-rc + 64); continue; }
            break;
        }
        for (int i = 0; i < n; ++i) {                                 // seqs[i].sam: heap memory owned by the caller (bwamem.c:1054)
            const size_t l = (size_t)(sam_off[i + 1] - sam_off[i]);
            char *o = (char *)malloc(l + 1);
            memcpy(o, sam.data() + sam_off[i], l); o[l] = 0;
            seqs[i].sam = o;
        }
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < seqs[i].l_seq; ++j) seqs[i].seq[j] = (char)nt4_of((uint8_t)seqs[i].seq[j]);   // bwamem.c:1067-1068
        return;
    }
    run_shards(shards, [&](Shard &sh) {
        sh.cnt.assign((size_t)sh.n + 1, 0);
        sh.regs.resize((size_t)(2 * (int64_t)sh.n + 64));
        if (b200_batch_upload(sh.ctx, sh.n, flat.data(), off.data() + sh.r0) || b200_batch_run(sh.ctx, opt, nullptr)) { sh.err = b200_last_error(); return; }
        shard_download(sh);
    });
    mem_alnreg_v *rv = (mem_alnreg_v *)malloc((size_t)(n > 0 ? n : 1) * sizeof(mem_alnreg_v));
    auto fill_rv = [&](bool first) {
        for (auto &sh : shards)
            for (int64_t i = 0, k = 0; i < sh.n; ++i) {
                mem_alnreg_v &v = rv[sh.r0 + i];
                if (!first) free(v.a);
                v.n = v.m = sh.cnt[i];
                v.a = (mem_alnreg_t *)malloc((size_t)(sh.cnt[i] > 0 ? sh.cnt[i] : 1) * sizeof(mem_alnreg_t));
                memcpy(v.a, sh.regs.data() + k, (size_t)sh.cnt[i] * sizeof(mem_alnreg_t));
                k += sh.cnt[i];
            }
    };
    fill_rv(true);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < seqs[i].l_seq; ++j) seqs[i].seq[j] = (char)nt4_of((uint8_t)seqs[i].seq[j]);   // bwamem.c:1067-1068

    // ---- batch-level insert-size statistics (the reference's own host code), then mate rescue on the device
    mem_pestat_t pes[4];
    mem_opt_t opt2 = *opt;
    if (pe) {
        if (pes0) memcpy(pes, pes0, 4 * sizeof(mem_pestat_t));        // bwamem.c:1237
        else f_pestat(opt, bns->l_pac, n, rv, pes);                   // bwamem.c:1238
        if (!(opt->flag & MEM_F_NO_RESCUE_)) {                        // bwamem_pair.c:291-302 for every pair
            run_shards(shards, [&](Shard &sh) {
                if (b200_batch_rescue(sh.ctx, opt, pes, nullptr, nullptr)) { sh.err = b200_last_error(); return; }
                shard_download(sh);
            });
            fill_rv(false);
            opt2.flag |= MEM_F_NO_RESCUE_;                            // worker2 must not repeat it
        }
    }
    opt = &opt2;
    const int n_units = pe ? n >> 1 : n;
    int n_threads = opt->n_threads > 0 ? opt->n_threads : 1;
    if (n_threads > n_units) n_threads = n_units > 0 ? n_units : 1;
    std::atomic<int> next(0);
    auto work = [&]() {
        for (;;) {
            const int i = next.fetch_add(1);
            if (i >= n_units) break;
            if (!pe) {                                                // bwamem.c:1202-1207
                f_mark(opt, rv[i].n, rv[i].a, n_processed + i);
                if (f_reorder5) f_reorder5(opt->T, &rv[i]);          // bwamem.c:1205
                f_reg2sam(opt, bns, pac, &seqs[i], &rv[i], 0, nullptr);
                free(rv[i].a);
            } else {                                                  // bwamem.c:1208-1212
                f_sam_pe(opt, bns, pac, pes, (uint64_t)((n_processed >> 1) + i), &seqs[i << 1], &rv[i << 1]);
                free(rv[i << 1 | 0].a); free(rv[i << 1 | 1].a);
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < n_threads; ++t) pool.emplace_back(work);
    work();
    for (auto &th : pool) th.join();
    free(rv);
}

#ifndef BWAMEM_B200_NO_DROPIN
// the reference's own name (bwamem.h:69)
extern "C" void mem_process_seqs(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac,
                                 int64_t n_processed, int n, bseq1_t *seqs, const mem_pestat_t *pes0)
{
    b200_mem_process_seqs(opt, bwt, bns, pac, n_processed, n, seqs, pes0);
}
#endif
