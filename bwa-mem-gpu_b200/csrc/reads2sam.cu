// Reads in host memory -> SAM text in host memory on ONE OR SEVERAL devices, as one call (b200_reads_to_sam).
//
// This is the per-GPU shard driver of the path north_star describes: the read set is cut into sub-batches of whole reads
// (whole pairs), the sub-batches are handed to the devices' host threads from one counter, each goes through worker1
// (b200_align_batch: seeding, chaining, extension), worker2 (finalize + mem_reg2aln on the resident regions) and the
// device SAM formatter (samdev.cu), and the text is placed in INPUT ORDER in the caller's buffer: a sub-batch's offset is
// the sum of the byte counts of all earlier sub-batches, known as soon as those have been COUNTED (not written), so the
// device-to-host copy of a sub-batch goes straight to its final place while later sub-batches are still on the devices.
// No collective and no device-to-device traffic: the FM-index is replicated, the reads shard.  Replaces,
// for the batch seam, superBatchMain / miniBatchMain (cuda/superbatch_process.cpp:133, cuda/minibatch_process.cpp:188)
// and worker1 + worker2 of mem_process_seqs (bwamem.c:1183-1244); the host does no formatting at all.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <string.h>
#include <thread>
#include <vector>
#include <math.h>
#include <unistd.h>
#include "samdev.cuh"        // after the standard headers: the emulator build redefines __noinline__

extern "C" void *b200_ctx_pinned(b200_ctx_t *ctx, int slot, size_t bytes);

namespace {

struct Placer {                                            // running prefix of the sub-batches' byte counts, in input order
    std::mutex mu; std::condition_variable cv;
    int next = 0; i64 offset = 0; bool failed = false; std::string err;
    i64 place(int j, i64 bytes)                            // blocks until every earlier sub-batch has been placed; -1 after a failure
    {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&]() { return next == j || failed; });
        if (failed) return -1;
        const i64 o = offset;
        offset += bytes; ++next;
        cv.notify_all();
        return o;
    }
    void fail(const char *msg)
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!failed) { failed = true; err = msg ? msg : "?"; }
        cv.notify_all();
    }
};

struct Job {                                               // one call
    const mem_opt_t *opt; const char *const *ann_names; const char *const *ann_annos; int n_seqs;
    const b200_reads_t *reads; i64 n_processed; const char *rg; int sub; int n_sub;
    i64 cap; char *sam; i64 *sam_off;
    Placer placer; std::atomic<int> next_sub{0};
    double t_align = 0, t_worker2 = 0, t_wait = 0; std::mutex tmu;
    // paired-end: worker1 of every sub-batch first (phase A: regions saved on their device, per-pair insert sizes to the
    // host), then mem_pestat over the whole call, then rescue + worker2 + text per sub-batch on the device that holds it
    bool pe = false; mem_pestat_t pes[4];
    std::vector<u64> isz;                                  // per pair: 0, or (insert size << 2 | orientation) + 1
    struct Saved { int owner = -1; DevBuf regs, cnt; i64 total = 0; };
    std::vector<Saved> saved;
};

struct Sub {                                               // one sub-batch on one device
    b200_ctx *ctx; const Job *job; SamDevParams P; int r0, n;
    DevBuf *text; i64 local_bytes = 0;
    i64 *h_pos = nullptr;                                  // page-locked: per read its offset inside the sub-batch's text
    std::vector<std::pair<int, i64> > chunk_base;          // (first read, text offset) of every chunk
};

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int grow_keep(b200_ctx *ctx, DevBuf &b, size_t keep, size_t need)       // grow-only, contents [0, keep) preserved
{
    if (need <= b.bytes) return 0;
    DevBuf nb;
    if (nb.ensure(need + need / 4)) return -1;
    if (keep) B200_CHECK(cudaMemcpyAsync(nb.p, b.p, keep, cudaMemcpyDeviceToDevice, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    b.release();
    b = nb;
    return 0;
}

// called by worker2 for every chunk of whole reads while its records are on the device: count, place inside the
// sub-batch, write
int sam_chunk(void *user, const R2AChunk &c)
{
    Sub &S = *(Sub *)user;
    b200_ctx *ctx = S.ctx;
    cudaStream_t st = ctx->st;
    const int n = c.r_hi - c.r_lo;
    if (n <= 0) return 0;
    DevBuf &d_len = ctx->d_sam[10], &d_pos = ctx->d_sam[11], &d_tmp = ctx->d_sam[12];
    if (d_len.ensure((size_t)n * 4) || d_pos.ensure(((size_t)n + 1) * 8) || d_tmp.ensure(((size_t)n / 1024 + 2) * 8) || ctx->d_ctl.ensure(CTL_SLOTS * 8)) return -1;
    B200_CHECK(cudaMemsetAsync(ctx->d_ctl.as<u64>() + CTL_OVF2, 0, 8, st));
    SamDevParams P = S.P;
    P.r_lo = c.r_lo; P.n_reads = n; P.roff = c.d_roff; P.g0 = c.g0;
    P.sel = c.sel; P.aln = c.aln; P.cig = c.cig; P.md = c.md; P.cig_base = c.cig_base; P.md_base = c.md_base;
    P.mate_reg = c.mate_reg; P.read_flag = c.read_flag;
    P.len = d_len.as<i32>(); P.pos = d_pos.as<i64>(); P.overflow = (int *)(ctx->d_ctl.as<u64>() + CTL_OVF2);
    if (launch_sam_count(P, ctx->n_sms, st)) return -1;
    if (launch_exclusive_scan(P.len, d_pos.as<i64>(), n, d_tmp.as<i64>(), st)) return -1;
    i64 total = 0;
    B200_CHECK(cudaMemcpyAsync(&total, d_pos.as<i64>() + n, 8, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    if (grow_keep(ctx, *S.text, (size_t)S.local_bytes, (size_t)(S.local_bytes + total) + 64)) return -1;
    P.out = S.text->as<char>() + S.local_bytes;
    if (launch_sam_write(P, ctx->n_sms, st)) return -1;
    if (S.h_pos) B200_CHECK(cudaMemcpyAsync(S.h_pos + c.r_lo, d_pos.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    u64 ovf = 0;
    B200_CHECK(cudaMemcpyAsync(&ovf, ctx->d_ctl.as<u64>() + CTL_OVF2, 8, cudaMemcpyDeviceToHost, st));
    B200_CHECK(cudaStreamSynchronize(st));
    if ((int)ovf) { b200_set_error("SAM formatter: %s", (int)ovf == 2 ? "count and write passes disagree" : "record longer than 2^31 bytes"); return -1; }
    S.chunk_base.push_back(std::make_pair(c.r_lo, S.local_bytes));
    S.local_bytes += total;
    return 0;
}

int upload_strings(b200_ctx *ctx, const char *const *v, int n, DevBuf &d_chars, DevBuf &d_off)       // n 0-terminated strings -> flat + n + 1 offsets
{
    std::vector<i32> off((size_t)n + 1, 0);
    std::string flat;
    for (int i = 0; i < n; ++i) { if (v && v[i]) flat += v[i]; off[i + 1] = (i32)flat.size(); }
    if (d_chars.ensure(flat.size() + 16) || d_off.ensure(off.size() * 4)) return -1;
    if (!flat.empty()) B200_CHECK(cudaMemcpyAsync(d_chars.p, flat.data(), flat.size(), cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(d_off.p, off.data(), off.size() * 4, cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaStreamSynchronize(ctx->st));
    return 0;
}

// a slice [r0, r0 + n) of a flat text field (names, comments): chars to d_chars, the n + 1 ABSOLUTE offsets to d_off;
// the kernel indexes (chars - off[r0]) + off[r]
int upload_slice(b200_ctx *ctx, const char *chars, const i64 *off, int r0, int n, DevBuf &d_chars, DevBuf &d_off, const char **biased)
{
    const i64 b0 = off[r0], b1 = off[r0 + n];
    if (d_chars.ensure((size_t)(b1 - b0) + 16) || d_off.ensure(((size_t)n + 1) * 8)) return -1;
    if (b1 > b0) B200_CHECK(cudaMemcpyAsync(d_chars.p, chars + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, ctx->st));
    B200_CHECK(cudaMemcpyAsync(d_off.p, off + r0, ((size_t)n + 1) * 8, cudaMemcpyHostToDevice, ctx->st));
    *biased = d_chars.as<char>() - b0;
    return 0;
}

// ---- paired-end, phase A: worker1 of this device's share of the sub-batches; regions stay on the device, the per-pair
// summary mem_pestat needs (bwamem_pair.c:76-90) goes to the host
void worker_pe_a(b200_ctx *ctx, int dev_index, Job *job)
{
    Placer &pl = job->placer;
    auto die = [&]() { pl.fail(b200_last_error()); };
    if (cudaSetDevice(ctx->device) != cudaSuccess) { b200_set_error("cudaSetDevice(%d) failed", ctx->device); die(); return; }
    const b200_reads_t &R = *job->reads;
    double t_align = 0;
    for (;;) {
        const int j = job->next_sub.fetch_add(1);
        if (j >= job->n_sub || pl.failed) break;
        const int r0 = (int)((i64)j * job->sub), n = (int)std::min<i64>(job->sub, R.n_reads - r0);
        const double t0 = now_s();
        i32 *cnt = (i32 *)b200_ctx_pinned(ctx, 0, (size_t)n * 4 + 4);
        if (!cnt) { die(); return; }
        const i64 nr = b200_align_batch(ctx, job->opt, n, R.seq, R.seq_off + r0, 0, nullptr, cnt);
        if (B200_IS_ERROR(nr) || nr > 0) { die(); return; }
        Job::Saved &sv = job->saved[j];
        sv.owner = dev_index; sv.total = ctx->total_regs;
        cudaStream_t st = ctx->st;
        const int n_pairs = n >> 1;
        auto fail = [&](const char *what) { b200_set_error("paired-end phase A: %s", what); die(); };
        if (sv.regs.ensure((size_t)(sv.total + 1) * sizeof(mem_alnreg_t)) || sv.cnt.ensure((size_t)n * 4 + 4) ||
            ctx->d_pa.ensure(((size_t)n + 1) * 8) || ctx->d_pb.ensure((size_t)n_pairs * 8 + 8) || ctx->d_scan_tmp.ensure(((size_t)n / 1024 + 2) * 8)) { die(); return; }
        if (sv.total && cudaMemcpyAsync(sv.regs.p, ctx->d_final.p, (size_t)sv.total * sizeof(mem_alnreg_t), cudaMemcpyDeviceToDevice, st) != cudaSuccess) { fail("copy"); return; }
        if (cudaMemcpyAsync(sv.cnt.p, ctx->d_final_n.p, (size_t)n * 4, cudaMemcpyDeviceToDevice, st) != cudaSuccess) { fail("copy"); return; }
        if (launch_exclusive_scan(ctx->d_final_n.as<i32>(), ctx->d_pa.as<i64>(), n, ctx->d_scan_tmp.as<i64>(), st)) { die(); return; }
        if (n_pairs) {
            if (launch_pestat_pairs(job->opt, ctx->ix.l_pac, (const mem_alnreg_t *)ctx->d_final.p, (const i64 *)ctx->d_pa.p, n_pairs,
                                    ctx->d_pb.as<u64>(), st)) { die(); return; }
            if (cudaMemcpyAsync(job->isz.data() + (r0 >> 1), ctx->d_pb.p, (size_t)n_pairs * 8, cudaMemcpyDeviceToHost, st) != cudaSuccess) { fail("summary download"); return; }
        }
        if (cudaStreamSynchronize(st) != cudaSuccess) { fail("synchronize"); return; }
        t_align += now_s() - t0;
    }
    std::lock_guard<std::mutex> lk(job->tmu);
    job->t_align += t_align;
}

// mem_pestat (bwamem_pair.c:72-135) from the per-pair summaries: insert-size quartiles, mean / std.dev inside the outlier
// bounds, the proper-pair window per orientation.  Host code: a reduction over the whole batch, a few
// microseconds per 10^6 pairs once the sizes are sorted.
void pestat_host(const mem_opt_t *opt, const std::vector<u64> &isz, mem_pestat_t pes[4])
{
    // insert sizes are integers in [1, max_ins] (the kernel applied bwamem_pair.c:89), so "sorted order" is a histogram walk:
    // the sums below add the values one by one in ascending order exactly as the reference does over its sorted array
    const size_t nbin = (size_t)(opt->max_ins > 0 ? opt->max_ins : 0) + 2;
    std::vector<u32> hist[4];
    size_t cnt[4] = { 0, 0, 0, 0 };
    memset(pes, 0, 4 * sizeof(mem_pestat_t));
    for (int d = 0; d < 4; ++d) hist[d].assign(nbin, 0);
    for (u64 v : isz) if (v) { const u64 w = v - 1, is = w >> 2; if (is < nbin) { ++hist[w & 3][is]; ++cnt[w & 3]; } }
    const bool verbose = getenv("B200_VERBOSE") != nullptr;
    for (int d = 0; d < 4; ++d) {
        mem_pestat_t *r = &pes[d];
        const std::vector<u32> &h = hist[d];
        const size_t n = cnt[d];
        if (n < 10) { r->failed = 1; continue; }                           // MIN_DIR_CNT
        auto at = [&](size_t idx) -> int {                                 // element idx of the sorted array
            size_t cum = 0;
            for (size_t v = 0; v < nbin; ++v) { cum += h[v]; if (cum > idx) return (int)v; }
            return (int)nbin - 1;
        };
        const int p25 = at((size_t)(int)(.25 * n + .499));
        const int p50 = at((size_t)(int)(.50 * n + .499));
        const int p75 = at((size_t)(int)(.75 * n + .499));
        r->low = (int)(p25 - 2.0 * (p75 - p25) + .499);                     // OUTLIER_BOUND
        if (r->low < 1) r->low = 1;
        r->high = (int)(p75 + 2.0 * (p75 - p25) + .499);
        size_t x = 0;
        r->avg = 0;
        for (size_t v = 0; v < nbin; ++v)
            if (h[v] && (i64)v >= r->low && (i64)v <= r->high) { for (u32 c = 0; c < h[v]; ++c) r->avg += (double)v; x += h[v]; }
        r->avg /= x;
        r->std = 0;
        for (size_t v = 0; v < nbin; ++v)
            if (h[v] && (i64)v >= r->low && (i64)v <= r->high) { const double t = ((double)v - r->avg) * ((double)v - r->avg); for (u32 c = 0; c < h[v]; ++c) r->std += t; }
        r->std = sqrt(r->std / x);
        r->low = (int)(p25 - 3.0 * (p75 - p25) + .499);                     // MAPPING_BOUND
        r->high = (int)(p75 + 3.0 * (p75 - p25) + .499);
        if (r->low > r->avg - 4.0 * r->std) r->low = (int)(r->avg - 4.0 * r->std + .499);      // MAX_STDDEV
        if (r->high < r->avg + 4.0 * r->std) r->high = (int)(r->avg + 4.0 * r->std + .499);
        if (r->low < 1) r->low = 1;
        if (verbose) fprintf(stderr, "[b200::mem_pestat] %c%c: %zu pairs, (25, 50, 75) percentile (%d, %d, %d), mean %.2f std %.2f, proper pairs (%d, %d)\n",
                             "FR"[d >> 1 & 1], "FR"[d & 1], n, p25, p50, p75, r->avg, r->std, r->low, r->high);
    }
    size_t mx = 0;
    for (int d = 0; d < 4; ++d) mx = std::max(mx, cnt[d]);
    for (int d = 0; d < 4; ++d)
        if (pes[d].failed == 0 && (double)cnt[d] < mx * 0.05) pes[d].failed = 1;              // MIN_DIR_RATIO
}

// ---- worker2 + text of one sub-batch whose worker1 regions are resident in ctx (single-end: straight after
// b200_align_batch; paired-end phase B: after the restore + mate rescue), then placement and the copy to the host
struct DevLoop {
    b200_ctx *ctx; Job *job; SamDevParams base; cudaEvent_t ev_written = nullptr; bool pending[2] = { false, false }; int iter = 0;
    double t_w2 = 0, t_wait = 0;
};

int finish_sub(DevLoop &L, int j, int r0, int n, const i32 *cnt)
{
    b200_ctx *ctx = L.ctx; Job *job = L.job;
    const b200_reads_t &R = *job->reads;
    const double t1 = now_s();
    Sub S;
    S.ctx = ctx; S.job = job; S.r0 = r0; S.n = n; S.P = L.base;
    S.P.seq = ctx->d_seq.as<u8>(); S.P.off = ctx->d_off.as<i64>();
    if (upload_slice(ctx, R.name, R.name_off, r0, n, ctx->d_sam[5], ctx->d_sam[6], &S.P.names)) return -1;
    S.P.name_off = ctx->d_sam[6].as<i64>();
    if (R.comment && R.comment_off) {
        if (upload_slice(ctx, R.comment, R.comment_off, r0, n, ctx->d_sam[8], ctx->d_sam[9], &S.P.comments)) return -1;
        S.P.comment_off = ctx->d_sam[9].as<i64>();
    }
    if (R.qual) {
        const i64 b0 = R.seq_off[r0], b1 = R.seq_off[r0 + n];
        if (ctx->d_sam[7].ensure((size_t)(b1 - b0) + 16)) return -1;
        if (b1 > b0) B200_CHECK(cudaMemcpyAsync(ctx->d_sam[7].p, R.qual + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, ctx->st));
        S.P.qual = ctx->d_sam[7].as<u8>();
    }
    const int tb = L.iter & 1;                             // text buffer; its previous copy (two sub-batches ago) must have left
    ++L.iter;
    if (L.pending[tb]) { cudaEventSynchronize(ctx->sam_ev[tb]); L.pending[tb] = false; }
    S.text = &ctx->d_sam[14 + tb];
    if (S.text->ensure((size_t)n * 64 + 4096)) return -1;
    if (job->sam_off) { S.h_pos = (i64 *)b200_ctx_pinned(ctx, 1 + tb, ((size_t)n + 1) * 8); if (!S.h_pos) return -1; }
    if (b200_batch_reg2sam_dev(ctx, job->opt, job->n_processed + r0, cnt, sam_chunk, &S, job->pe ? job->pes : nullptr) < 0) return -1;
    B200_CHECK(cudaEventRecord(L.ev_written, ctx->st));
    const double t2 = now_s();
    const i64 O = job->placer.place(j, S.local_bytes);     // input order: after every earlier sub-batch has been counted
    const double t3 = now_s();
    if (O < 0) return -2;                                  // another device failed
    if (O + S.local_bytes <= job->cap) {
        cudaStreamWaitEvent(ctx->st_d2h, L.ev_written, 0);
        if (S.local_bytes) B200_CHECK(cudaMemcpyAsync(job->sam + O, S.text->p, (size_t)S.local_bytes, cudaMemcpyDeviceToHost, ctx->st_d2h));
        cudaEventRecord(ctx->sam_ev[tb], ctx->st_d2h);
        L.pending[tb] = true;
    }
    if (job->sam_off) {                                    // offsets of the reads' records in the whole text
        size_t c = 0;
        for (int r = 0; r < n; ++r) {
            while (c + 1 < S.chunk_base.size() && S.chunk_base[c + 1].first <= r) ++c;
            job->sam_off[r0 + r] = O + (S.chunk_base.empty() ? 0 : S.chunk_base[c].second) + S.h_pos[r];
        }
        if (r0 + n == R.n_reads) job->sam_off[R.n_reads] = O + S.local_bytes;
    }
    L.t_w2 += t2 - t1; L.t_wait += t3 - t2;
    return 0;
}

// single-end: all of it; paired-end: phase B (dev_index >= 0: only the sub-batches this device holds)
void worker(b200_ctx *ctx, int dev_index, Job *job)
{
    Placer &pl = job->placer;
    auto die = [&]() { pl.fail(b200_last_error()); };
    if (cudaSetDevice(ctx->device) != cudaSuccess) { b200_set_error("cudaSetDevice(%d) failed", ctx->device); die(); return; }
    if (!ctx->st_d2h) {
        if (cudaStreamCreateWithFlags(&ctx->st_h2d, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&ctx->st_d2h, cudaStreamNonBlocking) != cudaSuccess) { b200_set_error("cudaStreamCreate failed"); die(); return; }
    }
    for (cudaEvent_t &e : ctx->sam_ev)
        if (!e && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { b200_set_error("cudaEventCreate failed"); die(); return; }
    DevLoop L;
    L.ctx = ctx; L.job = job;
    if (cudaEventCreateWithFlags(&L.ev_written, cudaEventDisableTiming) != cudaSuccess) { b200_set_error("cudaEventCreate failed"); die(); return; }
    const b200_reads_t &R = *job->reads;
    const mem_opt_t *opt = job->opt;
    memset(&L.base, 0, sizeof(L.base));
    L.base.opt = *opt;
    if (upload_strings(ctx, job->ann_names, job->n_seqs, ctx->d_sam[0], ctx->d_sam[1])) { die(); return; }
    L.base.ann_names = ctx->d_sam[0].as<char>(); L.base.ann_name_off = ctx->d_sam[1].as<i32>();
    if (job->ann_annos && (opt->flag & 0x100)) {
        if (upload_strings(ctx, job->ann_annos, job->n_seqs, ctx->d_sam[2], ctx->d_sam[3])) { die(); return; }
        L.base.ann_annos = ctx->d_sam[2].as<char>(); L.base.ann_anno_off = ctx->d_sam[3].as<i32>();
    }
    if (job->rg && job->rg[0]) {
        const char *one[1] = { job->rg };
        if (upload_strings(ctx, one, 1, ctx->d_sam[4], ctx->d_sam[13])) { die(); return; }
        L.base.rg = ctx->d_sam[4].as<char>(); L.base.rg_len = (int)strlen(job->rg);
    }
    double t_align = 0;
    for (int jj = 0; ; ++jj) {
        int j;
        if (job->pe) { j = jj; if (j >= job->n_sub) break; if (job->saved[j].owner != dev_index) continue; }
        else { j = job->next_sub.fetch_add(1); if (j >= job->n_sub) break; }
        if (pl.failed) break;
        const int r0 = (int)((i64)j * job->sub), n = (int)std::min<i64>(job->sub, R.n_reads - r0);
        const double t0 = now_s();
        i32 *cnt = (i32 *)b200_ctx_pinned(ctx, 0, (size_t)n * 4 + 4);
        if (!cnt) { die(); return; }
        if (!job->pe) {
            const i64 nr = b200_align_batch(ctx, opt, n, R.seq, R.seq_off + r0, 0, nullptr, cnt);       // worker1; regions stay on the device
            if (B200_IS_ERROR(nr) || nr > 0) { if (nr > 0) b200_set_error("internal: b200_align_batch with cap 0 returned %lld", (long long)nr); die(); return; }
        } else {                                           // reads back on the device, worker1's regions restored, mate rescue, new counts
            Job::Saved &sv = job->saved[j];
            if (b200_batch_upload(ctx, n, R.seq, R.seq_off + r0)) { die(); return; }
            if (ctx->d_final.ensure((size_t)(sv.total + 1) * sizeof(mem_alnreg_t)) || ctx->d_final_n.ensure((size_t)n * 4 + 4)) { die(); return; }
            bool ok = true;
            if (sv.total) ok = ok && cudaMemcpyAsync(ctx->d_final.p, sv.regs.p, (size_t)sv.total * sizeof(mem_alnreg_t), cudaMemcpyDeviceToDevice, ctx->st) == cudaSuccess;
            ok = ok && cudaMemcpyAsync(ctx->d_final_n.p, sv.cnt.p, (size_t)n * 4, cudaMemcpyDeviceToDevice, ctx->st) == cudaSuccess;
            ok = ok && (ctx->n_bytes == 0 || launch_pack_reads(ctx->d_seq.as<u8>(), ctx->n_bytes, ctx->st) == 0);
            ok = ok && cudaStreamSynchronize(ctx->st) == cudaSuccess;
            if (!ok) { b200_set_error("paired-end phase B: restoring the regions failed"); die(); return; }
            ctx->total_regs = sv.total;
            sv.regs.release(); sv.cnt.release();
            if (!(opt->flag & 0x20) && b200_batch_rescue(ctx, opt, job->pes, nullptr, nullptr)) { die(); return; }      // MEM_F_NO_RESCUE
            const i64 got = b200_batch_download(ctx, 0, nullptr, cnt);                                               // counts only
            if (B200_IS_ERROR(got)) { die(); return; }
        }
        t_align += now_s() - t0;
        const int rc = finish_sub(L, j, r0, n, cnt);
        if (rc == -2) break;
        if (rc) { die(); return; }
    }
    cudaStreamSynchronize(ctx->st_d2h);
    cudaStreamSynchronize(ctx->st);
    cudaEventDestroy(L.ev_written);
    std::lock_guard<std::mutex> lk(job->tmu);
    job->t_align += t_align; job->t_worker2 += L.t_w2; job->t_wait += L.t_wait;
}

}  // namespace

extern "C" int64_t b200_reads_to_sam(b200_ctx_t *const *ctxs, int n_ctx, const mem_opt_t *opt, const char *const *ann_names,
                                     const char *const *ann_annos, const b200_reads_t *reads, int64_t n_processed,
                                     const mem_pestat_t *pes0, const char *rg_id, int sub_reads, int64_t cap, char *sam,
                                     int64_t *sam_off, double *stats)
{
    if (!ctxs || n_ctx < 1 || !opt || !reads || !ann_names) { b200_set_error("b200_reads_to_sam: null argument"); return B200_E_FAIL; }
    for (int d = 0; d < n_ctx; ++d)
        if (!ctxs[d] || !ctxs[d]->has_index) { b200_set_error("b200_reads_to_sam: context %d has no index", d); return B200_E_FAIL; }
    const bool pe = (opt->flag & MEM_F_PE) != 0;
    if (reads->n_reads < 0 || reads->n_reads > 0x7fffffff) { b200_set_error("b200_reads_to_sam: %lld reads in one call", (long long)reads->n_reads); return B200_E_FAIL; }
    if (pe && ((reads->n_reads & 1) || (n_processed & 1))) { b200_set_error("b200_reads_to_sam: paired-end batches hold whole pairs (interleaved mates)"); return B200_E_FAIL; }
    if (sam_off) sam_off[0] = 0;
    if (reads->n_reads == 0) return 0;
    if (!reads->seq || !reads->seq_off || !reads->name || !reads->name_off) { b200_set_error("b200_reads_to_sam: reads need seq and name fields"); return B200_E_FAIL; }
    if (pe)                                                // mem_sam_pe refuses mates with different names (bwamem_pair.c:386)
        for (int64_t r = 0; r < reads->n_reads; r += 2) {
            const int64_t a0 = reads->name_off[r], a1 = reads->name_off[r + 1], a2 = reads->name_off[r + 2];
            if (a1 - a0 != a2 - a1 || memcmp(reads->name + a0, reads->name + a1, (size_t)(a1 - a0)) != 0) {
                b200_set_error("paired reads have different names (reads %lld and %lld)", (long long)r, (long long)r + 1);
                return B200_E_FAIL;
            }
        }
    Job job;
    job.opt = opt; job.ann_names = ann_names; job.ann_annos = ann_annos; job.n_seqs = ctxs[0]->ix.n_seqs;
    job.reads = reads; job.n_processed = n_processed; job.rg = rg_id; job.pe = pe;
    // sub-batches of about 2^20 reads, their number a multiple of the device count so that the last round is full
    const i64 want = sub_reads > 0 ? sub_reads : (1 << 20);
    i64 n_sub = (reads->n_reads + want - 1) / want;
    if (sub_reads <= 0) n_sub = (n_sub + n_ctx - 1) / n_ctx * n_ctx;
    i64 sub = (reads->n_reads + n_sub - 1) / n_sub;
    sub = (sub + 1) & ~(i64)1;                             // whole pairs, whatever the caller interleaved
    job.sub = (int)sub; job.n_sub = (int)((reads->n_reads + sub - 1) / sub);
    job.cap = cap; job.sam = sam; job.sam_off = sam_off;
    const double t0 = now_s();
    auto run_all = [&](void (*fn)(b200_ctx *, int, Job *)) {
        std::vector<std::thread> th;
        for (int d = 1; d < n_ctx; ++d) th.emplace_back(fn, ctxs[d], d, &job);
        fn(ctxs[0], 0, &job);
        for (auto &t : th) t.join();
    };
    if (pe) {
        job.isz.assign((size_t)(reads->n_reads >> 1), 0);
        job.saved.resize((size_t)job.n_sub);
        run_all(worker_pe_a);                              // worker1 everywhere, then the batch-level statistic, then the rest
        if (!job.placer.failed) {
            if (pes0) memcpy(job.pes, pes0, 4 * sizeof(mem_pestat_t));      // bwamem.c:1237
            else pestat_host(opt, job.isz, job.pes);                        // bwamem.c:1238
            job.next_sub = 0;
            run_all(worker);
        }
        for (auto &sv : job.saved) {
            if (sv.owner >= 0 && (sv.regs.p || sv.cnt.p)) { cudaSetDevice(ctxs[sv.owner]->device); sv.regs.release(); sv.cnt.release(); }
        }
    } else run_all(worker);
    if (stats) { stats[0] = now_s() - t0; stats[1] = job.t_align; stats[2] = job.t_worker2; stats[3] = job.t_wait; stats[4] = (double)job.n_sub; }
    if (job.placer.failed) { b200_set_error("%s", job.placer.err.c_str()); return B200_E_FAIL; }
    return job.placer.offset > cap ? -job.placer.offset : job.placer.offset;
}

// ---- FASTQ files -> SAM text on a file descriptor: the whole `bwa mem` data path (reader -> devices -> writer) in one call.
// Three stages run concurrently, as upstream's kt_pipeline does with its three steps (fastmap.c:305): a reader thread parses
// batch k + 1 (b200_fastq_read: chunk_bases bases per batch = bseq_read's rule, so mem_pestat sees the reference's batches),
// the calling thread aligns batch k on the devices (b200_reads_to_sam), a writer thread writes batch k - 1.
extern "C" int64_t b200_fastq_to_sam(b200_ctx_t *const *ctxs, int n_ctx, const mem_opt_t *opt, const char *const *ann_names,
                                     const char *const *ann_annos, const char *fn1, const char *fn2, int64_t chunk_bases,
                                     const mem_pestat_t *pes0, const char *rg_id, int n_threads, int copy_comment, int out_fd, double *stats)
{
    b200_fastq_t *f = b200_fastq_open(fn1, fn2, n_threads);
    if (!f) return B200_E_FAIL;
    struct Slot { b200_reads_t R; i64 n = 0; };
    Slot cur, next;
    std::mutex mu; std::condition_variable cv;
    bool next_ready = false, next_taken = true, reader_done = false, failed = false;
    std::string err;
    double t_read = 0, t_align = 0, t_write = 0;
    std::thread reader([&]() {
        for (;;) {
            {   std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&]() { return next_taken || failed; }); if (failed) return; }
            const double t0 = now_s();
            b200_reads_t R;
            const i64 n = b200_fastq_read(f, chunk_bases, &R);
            t_read += now_s() - t0;
            std::lock_guard<std::mutex> lk(mu);
            if (n < 0) { failed = true; err = b200_last_error(); cv.notify_all(); return; }
            next.R = R; next.n = n; next_ready = true; next_taken = false;
            if (n == 0) reader_done = true;
            cv.notify_all();
            if (n == 0) return;
        }
    });
    // two SAM buffers: the writer drains one while the devices fill the other (page-locked, kept by the first context across calls)
    char *sam[2] = { (char *)ctxs[0]->pin_p[10], (char *)ctxs[0]->pin_p[11] }; size_t cap[2] = { ctxs[0]->pin_cap[10], ctxs[0]->pin_cap[11] };
    std::thread writer;
    i64 w_bytes = 0; bool w_failed = false;
    i64 n_processed = 0, total_bytes = 0, n_batches = 0;
    const double t_all = now_s();
    for (int k = 0; ; ++k) {
        {   std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&]() { return next_ready || failed; });
            if (failed) break;
            cur = next; next_ready = false; next_taken = true; cv.notify_all(); }
        if (cur.n == 0) break;
        if (!copy_comment) { cur.R.comment = nullptr; cur.R.comment_off = nullptr; }                 // fastmap.c: comments only with -C
        mem_opt_t o = *opt;
        const int b = k & 1;
        const double t0 = now_s();
        i64 rc;
        for (;;) {
            const size_t want = cap[b] ? cap[b] : (size_t)(cur.R.seq_off[cur.n] * 2 + cur.n * 220 + cur.R.name_off[cur.n] + 4096);
            if (cap[b] < want) { sam[b] = (char *)b200_ctx_pinned(ctxs[0], 10 + b, want); cap[b] = sam[b] ? ctxs[0]->pin_cap[10 + b] : 0; if (!sam[b]) { rc = B200_E_FAIL; break; } }
            rc = b200_reads_to_sam(ctxs, n_ctx, &o, ann_names, ann_annos, &cur.R, n_processed, pes0, rg_id, 0, (i64)cap[b], sam[b], nullptr, nullptr);
            if (!B200_IS_ERROR(rc) && rc < 0) { sam[b] = (char *)b200_ctx_pinned(ctxs[0], 10 + b, (size_t)-rc + 4096); cap[b] = sam[b] ? ctxs[0]->pin_cap[10 + b] : 0; if (!sam[b]) { rc = B200_E_FAIL; break; } continue; }
            break;
        }
        t_align += now_s() - t0;
        if (B200_IS_ERROR(rc)) { std::lock_guard<std::mutex> lk(mu); failed = true; err = b200_last_error(); cv.notify_all(); break; }
        if (writer.joinable()) writer.join();                   // batch k - 1 is out: its buffer (the other one) is free again
        if (w_failed) { std::lock_guard<std::mutex> lk(mu); failed = true; err = "writing the SAM text failed"; cv.notify_all(); break; }
        const char *p = sam[b]; const i64 nb = rc;
        writer = std::thread([&, p, nb]() {
            const double tw = now_s();
            i64 done = 0;
            while (out_fd >= 0 && done < nb) { const ssize_t w = write(out_fd, p + done, (size_t)std::min<i64>(nb - done, (i64)1 << 30)); if (w <= 0) { w_failed = true; break; } done += w; }
            w_bytes += nb; t_write += now_s() - tw;
        });
        n_processed += cur.n; total_bytes += nb; ++n_batches;
    }
    if (writer.joinable()) writer.join();
    {   std::lock_guard<std::mutex> lk(mu); if (!reader_done) failed = true; cv.notify_all(); }     // an early exit: release the reader
    reader.join();
    b200_fastq_close(f);
    if (stats) { stats[0] = now_s() - t_all; stats[1] = t_read; stats[2] = t_align; stats[3] = t_write; stats[4] = (double)n_batches; stats[5] = (double)total_bytes; }
    if (!err.empty() || w_failed) { b200_set_error("%s", err.empty() ? "writing the SAM text failed" : err.c_str()); return B200_E_FAIL; }
    return n_processed;
}
