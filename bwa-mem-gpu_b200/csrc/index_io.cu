// On-disk device-ready index: the arrays a context holds after an upload / a device build -- the
// reference-layout BWT with Occ checkpoints (bwtindex.c:149-172), the sampled SA (bwt.c:62-84), the 2-bit .pac, the contig
// table -- AND the derived structures the kernels read (occ32 sectors, k-mer interval table, optionally the full suffix
// array) written as they lie in HBM, so that a later process maps the file and copies it to the device instead of
// re-laying bwa_idx_load's arrays (bwa.c:384-434, bwt.c:421-462) on every start.  One header, then every array at a
// 4096-byte boundary; loading is a sequence of large H2D copies straight from the page cache.
#include "ctx.h"
#include <fcntl.h>
#include <stdio.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <algorithm>

namespace {

enum { A_BWT = 0, A_SA, A_PAC, A_ANN_OFF, A_ANN_LEN, A_ANN_ALT, A_OCC32, A_KMER, A_SA_FULL, A_N };

struct IdxHeader {
    char magic[8];                       // "B200IDX6"
    u64 header_bytes, file_bytes;
    DevIndex ix;                         // scalars as they are; pointers are rewritten on load
    u64 off[A_N], bytes[A_N];            // file offset / length of every array (0 = absent)
    u64 ptr_off[A_N];                    // offset of the pointer the kernels use inside its array (arrays are padded in front of nothing today: 0)
    i32 force_wide, kmer_k; u64 idx_bwt_size;
    i32 sa_full_kind, pad;               // 0 = u32 per row, 1 = u64, 2 = 5 bytes per row
};

DevBuf *buf_of(b200_ctx *c, int a)
{
    DevBuf *b[A_N] = { &c->d_bwt, &c->d_sa, &c->d_pac, &c->d_ann_off, &c->d_ann_len, &c->d_ann_alt, &c->d_occ32, &c->d_kmer, &c->d_sa_full };
    return b[a];
}

}  // namespace

extern "C" int b200_index_save(b200_ctx_t *ctx, const char *path, int with_full_sa)
{
    if (!ctx || !ctx->has_index || !path) { b200_set_error("b200_index_save: no index"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    IdxHeader H;
    memset(&H, 0, sizeof(H));
    memcpy(H.magic, "B200IDX6", 8);
    H.header_bytes = sizeof(H);
    H.ix = ctx->ix;
    H.force_wide = ctx->force_wide; H.kmer_k = ctx->kmer_k; H.idx_bwt_size = ctx->idx_bwt_size;
    H.sa_full_kind = ctx->ix.sa_full40 ? 2 : ctx->ix.sa_full64 ? 1 : 0;
    // the bytes the kernels may touch of every array: what the context allocated for it (allocations are grow-only and
    // padded; the used part is recomputed here so that a re-used, larger buffer does not bloat the file)
    const DevIndex &ix = ctx->ix;
    const u64 n_blk = (ix.seq_len >> 6) + 2;
    u64 used[A_N];
    used[A_BWT] = ((u64)ctx->idx_bwt_size * 4 + 63) / 64 * 64 + 128;
    used[A_SA] = ix.n_sa * 8; used[A_PAC] = (u64)(ix.l_pac / 4 + 1);
    used[A_ANN_OFF] = (u64)ix.n_seqs * 8; used[A_ANN_LEN] = (u64)ix.n_seqs * 4; used[A_ANN_ALT] = (u64)ix.n_seqs * 4;
    used[A_OCC32] = n_blk * 32;
    used[A_KMER] = ix.kmer_tab ? (u64)kmer_tab_entries(ix.kmer_k) * (b200_wide_mode(ctx) == 1 ? 32 : 16) : 0;
    used[A_SA_FULL] = with_full_sa ? (ix.sa_full40 ? (u64)sa40_bytes(ix.seq_len + 1) : ix.sa_full64 ? (ix.seq_len + 1) * 8 : ix.sa_full32 ? (ix.seq_len + 1) * 4 : 0) : 0;
    u64 pos = (sizeof(H) + 4095) / 4096 * 4096;
    for (int a = 0; a < A_N; ++a) {
        DevBuf *b = buf_of(ctx, a);
        if (!b->p || !used[a]) continue;
        if (used[a] > b->bytes) { b200_set_error("b200_index_save: internal size mismatch for array %d", a); return -1; }
        H.off[a] = pos; H.bytes[a] = used[a];
        pos = (pos + used[a] + 4095) / 4096 * 4096;
    }
    H.file_bytes = pos;
    FILE *fp = fopen(path, "wb");
    if (!fp) { b200_set_error("b200_index_save: cannot create %s", path); return -1; }
    const size_t CH = (size_t)64 << 20;
    void *stage = nullptr;
    if (cudaHostAlloc(&stage, CH, cudaHostAllocDefault) != cudaSuccess) { fclose(fp); b200_set_error("b200_index_save: no staging memory"); return -1; }
    bool ok = fwrite(&H, sizeof(H), 1, fp) == 1;
    for (int a = 0; a < A_N && ok; ++a) {
        if (!H.bytes[a]) continue;
        ok = fseeko(fp, (off_t)H.off[a], SEEK_SET) == 0;
        const char *src = (const char *)buf_of(ctx, a)->p;
        for (u64 o = 0; o < H.bytes[a] && ok; o += CH) {
            const size_t n = (size_t)std::min<u64>(CH, H.bytes[a] - o);
            ok = cudaMemcpy(stage, src + o, n, cudaMemcpyDeviceToHost) == cudaSuccess && fwrite(stage, 1, n, fp) == n;
        }
    }
    if (ok) ok = ftruncate(fileno(fp), (off_t)H.file_bytes) == 0;
    cudaFreeHost(stage);
    ok = fclose(fp) == 0 && ok;
    if (!ok) { b200_set_error("b200_index_save: writing %s failed", path); return -1; }
    return 0;
}

extern "C" int b200_index_load(b200_ctx_t *ctx, const char *path, int flags)
{
    if (!ctx || !path) { b200_set_error("b200_index_load: null argument"); return -1; }
    B200_CHECK(cudaSetDevice(ctx->device));
    const int fd = open(path, O_RDONLY);
    if (fd < 0) { b200_set_error("b200_index_load: cannot open %s", path); return -1; }
    struct stat sb;
    IdxHeader H;
    if (fstat(fd, &sb) != 0 || (size_t)sb.st_size < sizeof(H) || pread(fd, &H, sizeof(H), 0) != (ssize_t)sizeof(H) ||
        memcmp(H.magic, "B200IDX6", 8) != 0 || H.header_bytes != sizeof(H) || H.file_bytes > (u64)sb.st_size) {
        close(fd); b200_set_error("b200_index_load: %s is not an index of this library version", path); return -1;
    }
    const char *map = (const char *)mmap(nullptr, (size_t)H.file_bytes, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (map == MAP_FAILED) { b200_set_error("b200_index_load: mmap failed"); return -1; }
    madvise((void *)map, (size_t)H.file_bytes, MADV_SEQUENTIAL);
    ctx->has_index = false;
    int rc = 0;
    for (int a = 0; a < A_N && !rc; ++a) {
        DevBuf *b = buf_of(ctx, a);
        if (!H.bytes[a]) continue;
        if (b->ensure_exact((size_t)H.bytes[a])) { rc = -1; break; }
        if (cudaMemcpyAsync(b->p, map + H.off[a], (size_t)H.bytes[a], cudaMemcpyHostToDevice, ctx->st) != cudaSuccess) { b200_set_error("b200_index_load: upload failed"); rc = -1; }
    }
    if (!rc && cudaStreamSynchronize(ctx->st) != cudaSuccess) { b200_set_error("b200_index_load: upload failed"); rc = -1; }
    munmap((void *)map, (size_t)H.file_bytes);
    if (rc) return rc;
    DevIndex ix = H.ix;
    ix.bwt = ctx->d_bwt.as<uint4>(); ix.sa = ctx->d_sa.as<u64>(); ix.pac = ctx->d_pac.as<u8>();
    ix.ann_off = ctx->d_ann_off.as<i64>(); ix.ann_len = ctx->d_ann_len.as<i32>(); ix.ann_alt = ctx->d_ann_alt.as<i32>();
    ix.occ32 = ctx->d_occ32.as<uint4>();
    ix.kmer_tab = H.bytes[A_KMER] ? ctx->d_kmer.as<uint4>() : nullptr;
    if (!ix.kmer_tab) ix.kmer_k = 0;
    ix.sa_full32 = nullptr; ix.sa_full64 = nullptr; ix.sa_full40 = nullptr;
    if (H.bytes[A_SA_FULL]) {
        if (H.sa_full_kind == 2) ix.sa_full40 = ctx->d_sa_full.as<u32>();
        else if (H.sa_full_kind == 1) ix.sa_full64 = ctx->d_sa_full.as<u64>();
        else ix.sa_full32 = ctx->d_sa_full.as<u32>();
    }
    ctx->ix = ix;
    ctx->force_wide = H.force_wide; ctx->kmer_k = H.kmer_k; ctx->idx_bwt_size = H.idx_bwt_size;
    ctx->bwt_identity = nullptr; ctx->stub_pac = false;
    if ((flags & B200_IDX_FULL_SA) && !ix.sa_full32 && !ix.sa_full64 && !ix.sa_full40) {   // the file was saved without it: expand the sampled SA on the device
        DevIndex &jx = ctx->ix;
        const bool narrow = jx.seq_len < 0xffffffffull;
        if (ctx->d_sa_full.ensure_exact((size_t)(jx.seq_len + 1) * (narrow ? 4 : 8))) return -1;
        if (launch_build_full_sa(jx, narrow ? ctx->d_sa_full.as<u32>() : nullptr, narrow ? nullptr : ctx->d_sa_full.as<u64>(), ctx->st)) return -1;
        if (narrow) jx.sa_full32 = ctx->d_sa_full.as<u32>(); else jx.sa_full64 = ctx->d_sa_full.as<u64>();
        B200_CHECK(cudaStreamSynchronize(ctx->st));
        if (b200_pack_full_sa(ctx)) return -1;
    }
    ctx->has_index = true;
    return 0;
}
