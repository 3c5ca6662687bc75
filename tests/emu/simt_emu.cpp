// TEST HARNESS ONLY (see simt_emu.h).  Block scheduler for the SIMT emulator.
#include "simt_emu.h"

namespace emu {

thread_local Block *g_block = nullptr;
thread_local Fiber *g_fiber = nullptr;

static void fiber_entry()
{
    Block *b = g_block;
    (*b->body)();
    g_fiber->done = true;
    b->idle_spins = 0;
    swapcontext(&g_fiber->ctx, &b->sched);
}

static int host_threads()
{
    const char *e = getenv("B200_EMU_THREADS");
    int n = e ? atoi(e) : (int)std::thread::hardware_concurrency();
    return n < 1 ? 1 : n;
}

static void run_blocks(std::atomic<unsigned> *next, unsigned grid, unsigned nthreads, size_t smem_bytes,
                       std::function<void()> *body)
{
    Block blk;
    blk.nthreads = nthreads;
    blk.grid = grid;
    blk.fibers.resize(nthreads);
    blk.warps.resize((nthreads + 31) / 32);
    blk.body = body;
    blk.dyn_smem = smem_bytes ? (unsigned char *)malloc(smem_bytes) : nullptr;
    for (unsigned t = 0; t < nthreads; ++t) {
        blk.fibers[t].stack = (char *)mmap(nullptr, kStackBytes, PROT_READ | PROT_WRITE,
                                           MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (blk.fibers[t].stack == MAP_FAILED) { perror("mmap"); abort(); }
    }
    g_block = &blk;
    for (;;) {
        unsigned bid = next->fetch_add(1);
        if (bid >= grid) break;
        blk.bid = bid;
        blk.bar_count = blk.bar_gen = 0;
        blk.idle_spins = 0;
        for (auto &w : blk.warps) memset(&w, 0, sizeof(Warp));
        for (unsigned t = 0; t < nthreads; ++t) {
            Fiber &f = blk.fibers[t];
            f.tid = t;
            f.done = false;
            getcontext(&f.ctx);
            f.ctx.uc_stack.ss_sp = f.stack;
            f.ctx.uc_stack.ss_size = kStackBytes;
            f.ctx.uc_link = nullptr;
            makecontext(&f.ctx, fiber_entry, 0);
        }
        unsigned done = 0;
        while (done < nthreads) {
            done = 0;
            for (unsigned t = 0; t < nthreads; ++t) {
                Fiber &f = blk.fibers[t];
                if (f.done) { ++done; continue; }
                g_fiber = &f;
                swapcontext(&blk.sched, &f.ctx);
                if (f.done) ++done;
            }
        }
    }
    for (unsigned t = 0; t < nthreads; ++t) munmap(blk.fibers[t].stack, kStackBytes);
    free(blk.dyn_smem);
    g_block = nullptr;
    g_fiber = nullptr;
}

void launch(dim3 grid, dim3 block, size_t smem_bytes, std::function<void()> body)
{
    if (grid.y != 1 || grid.z != 1 || block.y != 1 || block.z != 1) { fprintf(stderr, "[simt_emu] 1-D launches only\n"); abort(); }
    if (grid.x == 0) return;
    std::atomic<unsigned> next(0);
    int nt = std::min<int>(host_threads(), (int)grid.x);
    if (nt <= 1) { run_blocks(&next, grid.x, block.x, smem_bytes, &body); return; }
    std::vector<std::thread> th;
    for (int i = 0; i < nt; ++i) th.emplace_back(run_blocks, &next, grid.x, block.x, smem_bytes, &body);
    for (auto &t : th) t.join();
}

}  // namespace emu
