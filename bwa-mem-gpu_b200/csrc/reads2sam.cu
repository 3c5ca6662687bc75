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

void worker(b200_ctx *ctx, Job *job)
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
    cudaEvent_t ev_written = nullptr;
    if (cudaEventCreateWithFlags(&ev_written, cudaEventDisableTiming) != cudaSuccess) { b200_set_error("cudaEventCreate failed"); die(); return; }
    const b200_reads_t &R = *job->reads;
    const mem_opt_t *opt = job->opt;
    SamDevParams base;
    memset(&base, 0, sizeof(base));
    base.opt = *opt;
    if (upload_strings(ctx, job->ann_names, job->n_seqs, ctx->d_sam[0], ctx->d_sam[1])) { die(); return; }
    base.ann_names = ctx->d_sam[0].as<char>(); base.ann_name_off = ctx->d_sam[1].as<i32>();
    if (job->ann_annos && (opt->flag & 0x100)) {
        if (upload_strings(ctx, job->ann_annos, job->n_seqs, ctx->d_sam[2], ctx->d_sam[3])) { die(); return; }
        base.ann_annos = ctx->d_sam[2].as<char>(); base.ann_anno_off = ctx->d_sam[3].as<i32>();
    }
    if (job->rg && job->rg[0]) {
        const char *one[1] = { job->rg };
        if (upload_strings(ctx, one, 1, ctx->d_sam[4], ctx->d_sam[13])) { die(); return; }
        base.rg = ctx->d_sam[4].as<char>(); base.rg_len = (int)strlen(job->rg);
    }
    double t_align = 0, t_w2 = 0, t_wait = 0;
    int iter = 0;
    bool pending[2] = { false, false };
    for (;; ++iter) {
        const int j = job->next_sub.fetch_add(1);
        if (j >= job->n_sub) break;
        const int r0 = (int)((i64)j * job->sub), n = (int)std::min<i64>(job->sub, R.n_reads - r0);
        double t0 = now_s();
        i32 *cnt = (i32 *)b200_ctx_pinned(ctx, 0, (size_t)n * 4 + 4);
        if (!cnt) { die(); return; }
        const i64 nr = b200_align_batch(ctx, opt, n, R.seq, R.seq_off + r0, 0, nullptr, cnt);       // worker1; regions stay on the device
        if (B200_IS_ERROR(nr) || nr > 0) { if (nr > 0) b200_set_error("internal: b200_align_batch with cap 0 returned %lld", (long long)nr); die(); return; }
        double t1 = now_s();
        Sub S;
        S.ctx = ctx; S.job = job; S.r0 = r0; S.n = n; S.P = base;
        S.P.seq = ctx->d_seq.as<u8>(); S.P.off = ctx->d_off.as<i64>();
        if (upload_slice(ctx, R.name, R.name_off, r0, n, ctx->d_sam[5], ctx->d_sam[6], &S.P.names)) { die(); return; }
        S.P.name_off = ctx->d_sam[6].as<i64>();
        if (R.comment && R.comment_off) {
            if (upload_slice(ctx, R.comment, R.comment_off, r0, n, ctx->d_sam[8], ctx->d_sam[9], &S.P.comments)) { die(); return; }
            S.P.comment_off = ctx->d_sam[9].as<i64>();
        }
        if (R.qual) {
            const i64 b0 = R.seq_off[r0], b1 = R.seq_off[r0 + n];
            if (ctx->d_sam[7].ensure((size_t)(b1 - b0) + 16)) { die(); return; }
            if (b1 > b0 && cudaMemcpyAsync(ctx->d_sam[7].p, R.qual + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, ctx->st) != cudaSuccess) {
                b200_set_error("quality upload failed"); die(); return;
            }
            S.P.qual = ctx->d_sam[7].as<u8>();
        }
        const int tb = iter & 1;                           // text buffer; its previous copy (two sub-batches ago) must have left
        if (pending[tb]) { cudaEventSynchronize(ctx->sam_ev[tb]); pending[tb] = false; }
        S.text = &ctx->d_sam[14 + tb];
        if (S.text->ensure((size_t)n * 64 + 4096)) { die(); return; }
        if (job->sam_off) { S.h_pos = (i64 *)b200_ctx_pinned(ctx, 1 + tb, ((size_t)n + 1) * 8); if (!S.h_pos) { die(); return; } }
        if (b200_batch_reg2sam_dev(ctx, opt, job->n_processed + r0, cnt, sam_chunk, &S) < 0) { die(); return; }
        if (cudaEventRecord(ev_written, ctx->st) != cudaSuccess) { b200_set_error("cudaEventRecord failed"); die(); return; }
        double t2 = now_s();
        const i64 O = pl.place(j, S.local_bytes);          // input order: after every earlier sub-batch has been counted
        double t3 = now_s();
        if (O < 0) return;                                 // another device failed
        if (O + S.local_bytes <= job->cap) {
            cudaStreamWaitEvent(ctx->st_d2h, ev_written, 0);
            if (S.local_bytes && cudaMemcpyAsync(job->sam + O, S.text->p, (size_t)S.local_bytes, cudaMemcpyDeviceToHost, ctx->st_d2h) != cudaSuccess) {
                b200_set_error("SAM download failed"); die(); return;
            }
            cudaEventRecord(ctx->sam_ev[tb], ctx->st_d2h);
            pending[tb] = true;
        }
        if (job->sam_off) {                                // offsets of the reads' records in the whole text
            size_t c = 0;
            for (int r = 0; r < n; ++r) {
                while (c + 1 < S.chunk_base.size() && S.chunk_base[c + 1].first <= r) ++c;
                job->sam_off[r0 + r] = O + (S.chunk_base.empty() ? 0 : S.chunk_base[c].second) + S.h_pos[r];
            }
            if (r0 + n == R.n_reads) job->sam_off[R.n_reads] = O + S.local_bytes;
        }
        t_align += t1 - t0; t_w2 += t2 - t1; t_wait += t3 - t2;
    }
    cudaStreamSynchronize(ctx->st_d2h);
    cudaStreamSynchronize(ctx->st);
    cudaEventDestroy(ev_written);
    std::lock_guard<std::mutex> lk(job->tmu);
    job->t_align += t_align; job->t_worker2 += t_w2; job->t_wait += t_wait;
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
    if (opt->flag & MEM_F_PE) { b200_set_error("b200_reads_to_sam: paired-end batches go through b200_pairs_to_sam"); return B200_E_FAIL; }
    (void)pes0;
    if (reads->n_reads < 0 || reads->n_reads > 0x7fffffff) { b200_set_error("b200_reads_to_sam: %lld reads in one call", (long long)reads->n_reads); return B200_E_FAIL; }
    if (sam_off) sam_off[0] = 0;
    if (reads->n_reads == 0) return 0;
    if (!reads->seq || !reads->seq_off || !reads->name || !reads->name_off) { b200_set_error("b200_reads_to_sam: reads need seq and name fields"); return B200_E_FAIL; }
    Job job;
    job.opt = opt; job.ann_names = ann_names; job.ann_annos = ann_annos; job.n_seqs = ctxs[0]->ix.n_seqs;
    job.reads = reads; job.n_processed = n_processed; job.rg = rg_id;
    // sub-batches of about 2^20 reads, their number a multiple of the device count so that the last round is full
    const i64 want = sub_reads > 0 ? sub_reads : (1 << 20);
    i64 n_sub = (reads->n_reads + want - 1) / want;
    if (sub_reads <= 0) n_sub = (n_sub + n_ctx - 1) / n_ctx * n_ctx;
    i64 sub = (reads->n_reads + n_sub - 1) / n_sub;
    sub = (sub + 1) & ~(i64)1;                             // whole pairs, whatever the caller interleaved
    job.sub = (int)sub; job.n_sub = (int)((reads->n_reads + sub - 1) / sub);
    job.cap = cap; job.sam = sam; job.sam_off = sam_off;
    const double t0 = now_s();
    std::vector<std::thread> th;
    for (int d = 1; d < n_ctx; ++d) th.emplace_back(worker, ctxs[d], &job);
    worker(ctxs[0], &job);
    for (auto &t : th) t.join();
    if (stats) { stats[0] = now_s() - t0; stats[1] = job.t_align; stats[2] = job.t_worker2; stats[3] = job.t_wait; stats[4] = (double)job.n_sub; }
    if (job.placer.failed) { b200_set_error("%s", job.placer.err.c_str()); return B200_E_FAIL; }
    return job.placer.offset > cap ? -job.placer.offset : job.placer.offset;
}
