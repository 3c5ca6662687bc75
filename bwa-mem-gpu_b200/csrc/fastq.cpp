// Input side: FASTQ / FASTA text -> the flat, page-locked read buffers the device pipeline consumes
// (b200_reads_t), in batches cut exactly like bseq_read's (bwa.c:191-224: stop after the record that brings the batch to
// chunk_size bases, on an even record count), with kseq.h's record grammar (name = header up to the first blank, comment
// = the rest, sequence and quality possibly over several lines, '>' records without qualities) and trim_readno
// (bwa.c:64-68: a trailing "/1" or "/2" is dropped).  The reference reads with ONE thread through kseq + zlib and
// allocates four strings per read (kseq2bseq1, bwa.c:180-188); the fork's flat super-batch buffers
// (bseq_read2, bwa.c:144-178; cuda/superbatch_process.cpp:85-121) are the same idea as here, still filled by one thread.
//
// Here: plain files are memory-mapped and parsed in place; gzip files are inflated into a window (BGZF files -- gzip "by
// block", what bgzip writes -- block-parallel, ordinary gzip streams by one thread while the parsers work on the previous
// window); the window is cut at record boundaries into one piece per thread; pass 1 measures every record of a piece,
// a prefix sum places them, pass 2 writes sequence, quality, name and comment bytes straight into their final place.
// Two files (mates in separate files) are read by two such readers and interleaved record by record.
#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>
#include <algorithm>
#include <functional>
#include <string>
#include <thread>
#include <vector>
#include "../../include/bwamem_b200.h"

extern "C" void *b200_host_alloc(size_t bytes);
extern "C" void b200_host_free(void *p);
void b200_set_error(const char *fmt, ...);

namespace {

struct Rec { const char *name; int l_name; const char *comment; int l_comment; int l_seq, l_qual; const char *body; };

// one record starting at p (p[0] is '@' or '>'); returns the start of the next record or nullptr when the text ends
// inside this one (`complete` = false) -- the caller then waits for more text, or at end of input takes what is there
inline const char *eol(const char *p, const char *end) { const char *q = (const char *)memchr(p, '\n', (size_t)(end - p)); return q ? q : end; }

const char *scan_record(const char *p, const char *end, bool at_eof, Rec &r, bool &complete)
{
    complete = false;
    const char *h = p + 1, *he = eol(h, end);
    if (he == end && !at_eof) return nullptr;
    const char *hl = he;
    if (hl > h && hl[-1] == '\r') --hl;
    const char *sp = h;
    while (sp < hl && !(*sp == ' ' || *sp == '\t')) ++sp;
    r.name = h; r.l_name = (int)(sp - h);
    const char *c = sp < hl ? sp + 1 : hl;                              // kseq: everything after ONE separator character
    r.comment = c; r.l_comment = (int)(hl - c);
    if (r.l_name > 2 && r.name[r.l_name - 2] == '/' && r.name[r.l_name - 1] >= '0' && r.name[r.l_name - 1] <= '9') r.l_name -= 2;   // trim_readno
    const char *q = he < end ? he + 1 : end;
    r.body = q; r.l_seq = 0; r.l_qual = 0;
    // sequence lines until '+', '>' or '@' at a line start (kseq.h:183-195)
    for (;;) {
        if (q >= end) { if (!at_eof) return nullptr; complete = true; return end; }
        if (*q == '>' || *q == '@') { complete = true; return q; }     // FASTA record, or FASTQ without '+'
        if (*q == '+') break;
        const char *le = eol(q, end);
        if (le == end && !at_eof) return nullptr;
        const char *ll = le;
        if (ll > q && ll[-1] == '\r') --ll;
        r.l_seq += (int)(ll - q);
        q = le < end ? le + 1 : end;
    }
    const char *pe = eol(q, end);                                        // the '+' line
    if (pe == end && !at_eof) return nullptr;
    q = pe < end ? pe + 1 : end;
    while (r.l_qual < r.l_seq) {                                         // quality lines until as long as the sequence (kseq.h:203-206)
        if (q >= end) { if (!at_eof) return nullptr; break; }
        const char *le = eol(q, end);
        if (le == end && !at_eof) return nullptr;
        const char *ll = le;
        if (ll > q && ll[-1] == '\r') --ll;
        r.l_qual += (int)(ll - q);
        q = le < end ? le + 1 : end;
    }
    complete = true;
    return q;
}

// copies the record's sequence / quality characters (line breaks removed); `end` = end of the text
void copy_lines(const char *&q, const char *end, int want, char *dst)
{
    int n = 0;
    while (n < want && q < end) {
        const size_t span = std::min<size_t>((size_t)(want - n) + 2, (size_t)(end - q));
        const char *le = (const char *)memchr(q, '\n', span);
        int l = le ? (int)(le - q) : (int)std::min<size_t>((size_t)(want - n), (size_t)(end - q));
        if (l > 0 && q[l - 1] == '\r') --l;
        if (l > want - n) l = want - n;
        memcpy(dst + n, q, (size_t)l); n += l;
        q = le ? le + 1 : q + l;
    }
}

void copy_body(const Rec &r, const char *end, char *seq, char *qual)
{
    const char *q = r.body;
    copy_lines(q, end, r.l_seq, seq);
    if (!qual || r.l_qual == 0) return;
    while (q < end && *q != '+') { const char *nl = (const char *)memchr(q, '\n', (size_t)(end - q)); if (!nl) return; q = nl + 1; }
    { const char *nl = (const char *)memchr(q, '\n', (size_t)(end - q)); if (!nl) return; q = nl + 1; }     // past the '+' line
    copy_lines(q, end, r.l_qual, qual);
}

struct Stream {                                            // one input file
    int fd = -1; bool plain = false, bgzf = false, eof = false;
    const char *map = nullptr; size_t map_len = 0, map_pos = 0;       // plain: the mapping and the parse position
    gzFile gz = nullptr;                                                // ordinary gzip
    FILE *fp = nullptr;                                                 // BGZF: compressed blocks are read by hand
    std::vector<char> win; size_t win_lo = 0;                           // inflated text not yet consumed: win[win_lo ..)
    const char *cur = nullptr, *cur_end = nullptr;                      // the text the next batch is parsed from
    std::string err;

    bool open(const char *fn, int n_threads)
    {
        fd = ::open(fn, O_RDONLY);
        if (fd < 0) { err = std::string("cannot open ") + fn; return false; }
        unsigned char m[18] = { 0 };
        const ssize_t got = pread(fd, m, sizeof(m), 0);
        struct stat sb;
        if (fstat(fd, &sb) != 0) { err = "fstat failed"; return false; }
        if (got >= 2 && m[0] == 0x1f && m[1] == 0x8b) {
            bgzf = got >= 18 && (m[3] & 4) && m[12] == 'B' && m[13] == 'C';
            if (bgzf) { fp = fdopen(fd, "rb"); if (!fp) { err = "fdopen failed"; return false; } }
            else { gz = gzdopen(fd, "rb"); if (!gz) { err = "gzdopen failed"; return false; } gzbuffer(gz, 1 << 20); }
        } else {
            plain = true;
            map_len = (size_t)sb.st_size;
            if (map_len) {
                map = (const char *)mmap(nullptr, map_len, PROT_READ, MAP_PRIVATE, fd, 0);
                if (map == MAP_FAILED) { map = nullptr; err = "mmap failed"; return false; }
                madvise((void *)map, map_len, MADV_SEQUENTIAL);
            } else eof = true;
        }
        (void)n_threads;
        return true;
    }
    // make at least `want` bytes of text available in [cur, cur_end) (less only at the end of the input)
    bool fill(size_t want, int n_threads)
    {
        if (plain) { cur = map + map_pos; cur_end = map + map_len; eof = true; return true; }
        if (win_lo) { win.erase(win.begin(), win.begin() + (ptrdiff_t)win_lo); win_lo = 0; }
        while (win.size() < want && !eof) {
            if (!bgzf) {
                const size_t old = win.size(), step = std::max<size_t>(want - old, (size_t)8 << 20);
                win.resize(old + step);
                const int n = gzread(gz, win.data() + old, (unsigned)std::min<size_t>(step, 1u << 30));
                if (n < 0) { err = "gzread failed"; return false; }
                win.resize(old + (size_t)n);
                if ((size_t)n < step) eof = true;
            } else if (!inflate_blocks(std::max<size_t>(want - win.size(), (size_t)32 << 20), n_threads)) return false;
        }
        cur = win.data(); cur_end = win.data() + win.size();
        return true;
    }
    void consumed(const char *upto)
    {
        if (plain) map_pos = (size_t)(upto - map);
        else win_lo = (size_t)(upto - win.data());
    }
    // BGZF: read compressed blocks worth about `want` bytes of text, inflate them with n_threads, append in order
    bool inflate_blocks(size_t want, int n_threads)
    {
        struct Blk { std::vector<unsigned char> c; size_t isize; size_t off; };
        std::vector<Blk> blks;
        size_t total = 0;
        while (total < want) {
            unsigned char h[18];
            const size_t g = fread(h, 1, 18, fp);
            if (g == 0) { eof = true; break; }
            if (g < 18 || h[0] != 0x1f || h[1] != 0x8b || !(h[3] & 4) || h[12] != 'B' || h[13] != 'C') { err = "not a BGZF block"; return false; }
            const size_t bsize = (size_t)(h[16] | h[17] << 8) + 1, xlen = (size_t)(h[10] | h[11] << 8);
            Blk b;
            b.c.resize(bsize - 18);
            if (fread(b.c.data(), 1, b.c.size(), fp) != b.c.size()) { err = "truncated BGZF block"; return false; }
            const size_t skip = xlen - 6;                                // extra subfields after BC
            const unsigned char *t = b.c.data() + b.c.size() - 4;
            b.isize = (size_t)t[0] | (size_t)t[1] << 8 | (size_t)t[2] << 16 | (size_t)t[3] << 24;
            b.off = total; total += b.isize;
            if (skip) b.c.erase(b.c.begin(), b.c.begin() + (ptrdiff_t)skip);
            blks.push_back(std::move(b));
        }
        const size_t old = win.size();
        win.resize(old + total);
        const int nt = std::max(1, std::min<int>(n_threads, (int)blks.size()));
        std::vector<int> bad((size_t)nt, 0);
        auto work = [&](int t) {
            for (size_t i = (size_t)t; i < blks.size(); i += (size_t)nt) {
                z_stream zs;
                memset(&zs, 0, sizeof(zs));
                if (inflateInit2(&zs, -15) != Z_OK) { bad[(size_t)t] = 1; return; }
                zs.next_in = blks[i].c.data(); zs.avail_in = (unsigned)(blks[i].c.size() - 8);
                zs.next_out = (unsigned char *)win.data() + old + blks[i].off; zs.avail_out = (unsigned)blks[i].isize;
                const int rc = inflate(&zs, Z_FINISH);
                inflateEnd(&zs);
                if (rc != Z_STREAM_END || zs.avail_out) { bad[(size_t)t] = 1; return; }
            }
        };
        std::vector<std::thread> th;
        for (int t = 1; t < nt; ++t) th.emplace_back(work, t);
        work(0);
        for (auto &x : th) x.join();
        for (int b : bad) if (b) { err = "BGZF inflate failed"; return false; }
        return true;
    }
    void close()
    {
        if (map) munmap((void *)map, map_len);
        if (gz) gzclose(gz); else if (fp) fclose(fp); else if (fd >= 0) ::close(fd);
        map = nullptr; gz = nullptr; fp = nullptr; fd = -1;
    }
};

struct Piece { const char *lo, *hi; std::vector<Rec> recs; };

}  // namespace

struct HBuf {                                              // grow-only host buffer, page-locked when a device is present
    char *p = nullptr; size_t cap = 0; bool pinned = false;
    bool grow(size_t need)
    {
        if (need <= cap) return true;
        release();
        cap = need + need / 4 + 4096;
        p = (char *)b200_host_alloc(cap);
        pinned = p != nullptr;
        if (!p) p = (char *)malloc(cap);                   // no device in this process (the parser on its own): pageable memory
        if (!p) cap = 0;
        return p != nullptr;
    }
    void release() { if (p) { if (pinned) b200_host_free(p); else free(p); } p = nullptr; cap = 0; }
};

struct b200_fastq {
    Stream s[2]; int n_files = 0, n_threads = 1;
    // two sets of output buffers, used alternately: a batch stays valid until the call AFTER the next one, so a consumer can
    // work on batch k while another thread already reads batch k + 1
    struct Out { HBuf seq, qual, name, comment, seq_off, name_off, com_off; } out[2];
    int parity = 0;
    bool warned = false;
    Rec *recs[2] = { nullptr, nullptr }; size_t recs_cap[2] = { 0, 0 }, n_recs[2] = { 0, 0 };     // records indexed for the current call
};

extern "C" b200_fastq_t *b200_fastq_open(const char *fn1, const char *fn2, int n_threads)
{
    b200_fastq *f = new b200_fastq();
    f->n_threads = n_threads > 0 ? (n_threads < 64 ? n_threads : 64) : 1;
    f->n_files = fn2 ? 2 : 1;
    const char *fn[2] = { fn1, fn2 };
    for (int k = 0; k < f->n_files; ++k)
        if (!f->s[k].open(fn[k], f->n_threads)) { b200_set_error("%s", f->s[k].err.c_str()); delete f; return nullptr; }
    return f;
}

extern "C" void b200_fastq_close(b200_fastq_t *f)
{
    if (!f) return;
    for (int k = 0; k < f->n_files; ++k) f->s[k].close();
    free(f->recs[0]); free(f->recs[1]);
    for (auto &o : f->out) {
        HBuf *all[] = { &o.seq, &o.qual, &o.name, &o.comment, &o.seq_off, &o.name_off, &o.com_off };
        for (HBuf *b : all) b->release();
    }
    delete f;
}

// records of stream s up to `max_recs` / `want_bases` (bseq_read's cut is applied by the caller over both streams)
static bool index_stream(b200_fastq *f, int which, Stream &s, size_t text_budget, const char **upto)
{
    if (!s.fill(text_budget, f->n_threads)) { b200_set_error("%s", s.err.c_str()); return false; }
    const char *lo = s.cur, *end = s.cur_end;
    const char *lim = lo + std::min<size_t>((size_t)(end - lo), text_budget);
    const bool whole = lim == end && s.eof;
    // pieces: cut [lo, lim) at record starts.  A record start inside the text is a line beginning with '@' or '>' whose
    // line after next begins with '+' (4-line FASTQ), or any '>' line (FASTA); multi-line FASTQ is parsed by one thread.
    const int nt = f->n_threads;
    std::vector<const char *> cuts(1, lo);
    const bool fastq = lo < end && *lo == '@';
    if (nt > 1 && (size_t)(lim - lo) > ((size_t)1 << 20)) {
        for (int t = 1; t < nt; ++t) {
            const char *p = lo + (size_t)(lim - lo) * (size_t)t / (size_t)nt;
            const char *found = nullptr;
            for (int tries = 0; tries < 64 && p < lim; ++tries) {
                const char *nl = (const char *)memchr(p, '\n', (size_t)(lim - p));
                if (!nl || nl + 1 >= lim) break;
                p = nl + 1;
                if (!fastq) { if (*p == '>') { found = p; break; } continue; }
                if (*p != '@') continue;
                const char *l2 = (const char *)memchr(p, '\n', (size_t)(lim - p));
                const char *l3 = l2 ? (const char *)memchr(l2 + 1, '\n', (size_t)(lim - l2 - 1)) : nullptr;
                if (l3 && l3 + 1 < lim && l3[1] == '+') { found = p; break; }
            }
            if (found && found > cuts.back()) cuts.push_back(found);
        }
    }
    cuts.push_back(lim);
    const int np = (int)cuts.size() - 1;
    std::vector<Piece> pieces((size_t)np);
    std::vector<const char *> stop((size_t)np, nullptr);
    auto work = [&](int k) {
        Piece &P = pieces[(size_t)k];
        const char *p = cuts[(size_t)k], *hi = cuts[(size_t)k + 1];
        const bool last = k == np - 1;
#ifdef MADV_POPULATE_READ
        if (s.plain && hi > p) {                             // fault the piece's pages in with one call instead of one fault per 4 KB
            const uintptr_t a0 = (uintptr_t)p & ~(uintptr_t)4095, a1 = ((uintptr_t)hi + 4095) & ~(uintptr_t)4095;
            const uintptr_t mend = ((uintptr_t)s.map + s.map_len + 4095) & ~(uintptr_t)4095;
            madvise((void *)a0, (size_t)((a1 < mend ? a1 : mend) - a0), MADV_POPULATE_READ);
        }
#endif
        while (p < hi) {
            while (p < hi && (*p == '\n' || *p == '\r')) ++p;
            if (p >= hi) break;
            Rec r; bool complete;
            // inner pieces end at a record start, so their records are complete inside [p, hi); the last piece may end mid-record
            const char *nx = scan_record(p, last ? end : hi, last ? whole : true, r, complete);
            if (!nx || !complete || (last && nx > lim && !whole)) { break; }
            P.recs.push_back(r);
            p = nx;
        }
        stop[(size_t)k] = p;
    };
    std::vector<std::thread> th;
    for (int k = 1; k < np; ++k) th.emplace_back(work, k);
    work(0);
    for (auto &x : th) x.join();
    // the pieces' records side by side in one array (copied by the pieces' threads)
    std::vector<size_t> base((size_t)np + 1, 0);
    for (int k = 0; k < np; ++k) base[(size_t)k + 1] = base[(size_t)k] + pieces[(size_t)k].recs.size();
    const size_t total = base[(size_t)np];
    if (total > f->recs_cap[which]) {
        free(f->recs[which]);
        f->recs_cap[which] = total + total / 4 + 1024;
        f->recs[which] = (Rec *)malloc(f->recs_cap[which] * sizeof(Rec));
        if (!f->recs[which]) { f->recs_cap[which] = 0; b200_set_error("out of memory"); return false; }
    }
    f->n_recs[which] = total;
    auto gather = [&](int k) { if (!pieces[(size_t)k].recs.empty()) memcpy(f->recs[which] + base[(size_t)k], pieces[(size_t)k].recs.data(), pieces[(size_t)k].recs.size() * sizeof(Rec)); };
    th.clear();
    for (int k = 1; k < np; ++k) th.emplace_back(gather, k);
    gather(0);
    for (auto &x : th) x.join();
    *upto = stop[(size_t)np - 1];
    return true;
}

extern "C" int64_t b200_fastq_read(b200_fastq_t *f, int64_t chunk_bases, b200_reads_t *out)
{
    if (!f || !out || chunk_bases <= 0) { b200_set_error("b200_fastq_read: bad argument"); return -1; }
    memset(out, 0, sizeof(*out));
    // text budget: a record of l bases takes about 2 l + name + 6 bytes; index a little more than one batch needs
    size_t budget = (size_t)chunk_bases * 9 / 4 + ((size_t)2 << 20);
    const char *upto[2] = { nullptr, nullptr };
    int64_t n = 0;
    const int two = f->n_files == 2;
    auto rec_of = [&](int64_t i) -> const Rec & { return two ? f->recs[i & 1][(size_t)(i >> 1)] : f->recs[0][(size_t)i]; };
    auto par = [&](int64_t count, const std::function<void(int, int64_t, int64_t)> &fn) {      // fn(thread, lo, hi) over [0, count)
        const int nt = (int)std::max<int64_t>(1, std::min<int64_t>(f->n_threads, count / 8192));
        std::vector<std::thread> th;
        for (int t = 1; t < nt; ++t) th.emplace_back(fn, t, count * t / nt, count * (t + 1) / nt);
        fn(0, 0, count / nt);
        for (auto &x : th) x.join();
        return nt;
    };
    b200_fastq::Out &O = f->out[f->parity];
    int64_t *seq_off = nullptr, *name_off = nullptr, *com_off = nullptr;
    for (;;) {
        for (int k = 0; k < f->n_files; ++k)
            if (!index_stream(f, k, f->s[k], two ? budget / 2 : budget, &upto[k])) return -1;
        // offsets of every field of every indexed record, in batch order (mates alternate): a two-level prefix sum
        const int64_t avail = two ? 2 * (int64_t)std::min(f->n_recs[0], f->n_recs[1]) : (int64_t)f->n_recs[0];
        const size_t need_n = ((size_t)avail + 1) * 8;
        if (!O.seq_off.grow(need_n) || !O.name_off.grow(need_n) || !O.com_off.grow(need_n)) { b200_set_error("out of memory"); return -1; }
        seq_off = (int64_t *)O.seq_off.p; name_off = (int64_t *)O.name_off.p; com_off = (int64_t *)O.com_off.p;
        int64_t part[64][3];
        const int nt = par(avail, [&](int t, int64_t lo, int64_t hi) {
            int64_t a = 0, b = 0, c = 0;
            for (int64_t i = lo; i < hi; ++i) { const Rec &r = rec_of(i); a += r.l_seq; b += r.l_name; c += r.l_comment; }
            part[t][0] = a; part[t][1] = b; part[t][2] = c;
        });
        int64_t run[3] = { 0, 0, 0 };
        for (int t = 0; t < nt; ++t) for (int q = 0; q < 3; ++q) { const int64_t v = part[t][q]; part[t][q] = run[q]; run[q] += v; }
        par(avail, [&](int t, int64_t lo, int64_t hi) {
            int64_t a = part[t][0], b = part[t][1], c = part[t][2];
            for (int64_t i = lo; i < hi; ++i) { const Rec &r = rec_of(i); seq_off[i] = a; name_off[i] = b; com_off[i] = c; a += r.l_seq; b += r.l_name; c += r.l_comment; }
        });
        seq_off[avail] = run[0]; name_off[avail] = run[1]; com_off[avail] = run[2];
        // bseq_read's cut (bwa.c:191-224): records in order until size >= chunk_size on an even count (sizes only grow, so
        // that is the first count with size >= chunk, rounded up to even)
        bool cut = false;
        n = avail;
        if (run[0] >= chunk_bases) {
            const int64_t n0 = std::lower_bound(seq_off, seq_off + avail + 1, chunk_bases) - seq_off;      // first count whose size reaches the chunk
            const int64_t ne = (n0 + 1) & ~(int64_t)1;
            if (ne <= avail) { n = ne; cut = true; }
        }
        bool all_in = true;                                  // every stream fully indexed to its end?
        for (int k = 0; k < f->n_files; ++k) all_in = all_in && f->s[k].eof && upto[k] == f->s[k].cur_end;
        if (cut || all_in) break;
        budget *= 2;                                         // long records: look at more text
    }
    f->parity ^= 1;
    if (two && f->n_recs[0] != f->n_recs[1] && !f->warned) {
        bool e0 = f->s[0].eof && upto[0] == f->s[0].cur_end, e1 = f->s[1].eof && upto[1] == f->s[1].cur_end;
        if (e0 && e1) { fprintf(stderr, "[W::b200_fastq_read] the %s file has fewer sequences.\n", f->n_recs[0] < f->n_recs[1] ? "1st" : "2nd"); f->warned = true; }
    }
    if (n == 0) return 0;
    const size_t per = two ? (size_t)n / 2 : (size_t)n;
    const int64_t so = seq_off[n], no = name_off[n], co = com_off[n];
    int fl[64][3];
    const int ntf = par(n, [&](int t, int64_t lo, int64_t hi) {
        int aq = 0, nq = 0, ac = 0;
        for (int64_t i = lo; i < hi; ++i) { const Rec &r = rec_of(i); aq |= r.l_qual > 0; nq |= r.l_qual != r.l_seq; ac |= r.l_comment > 0; }
        fl[t][0] = aq; fl[t][1] = nq; fl[t][2] = ac;
    });
    bool any_qual = false, all_qual = true, any_com = false;
    for (int t = 0; t < ntf; ++t) { any_qual |= fl[t][0] != 0; all_qual &= fl[t][1] == 0; any_com |= fl[t][2] != 0; }
    if (any_qual && !all_qual) { b200_set_error("FASTQ batch mixes reads with and without (complete) qualities"); return -1; }
    if (!O.seq.grow((size_t)so + 64) || (any_qual && !O.qual.grow((size_t)so + 64)) || !O.name.grow((size_t)no + 64) ||
        !O.comment.grow((size_t)co + 64)) { b200_set_error("out of memory"); return -1; }
    // ---- pass 2: every thread writes its share of the records into place
    const int nt = (int)std::max<int64_t>(1, std::min<int64_t>(f->n_threads, n / 4096));
    auto fillw = [&](int t) {
        const int64_t lo = n * t / nt, hi = n * (t + 1) / nt;
        for (int64_t i = lo; i < hi; ++i) {
            const Rec &r = rec_of(i);
            const Stream &S = f->s[two ? (int)(i & 1) : 0];
            copy_body(r, S.cur_end, O.seq.p + seq_off[i], any_qual ? O.qual.p + seq_off[i] : nullptr);
            memcpy(O.name.p + name_off[i], r.name, (size_t)r.l_name);
            if (r.l_comment) memcpy(O.comment.p + com_off[i], r.comment, (size_t)r.l_comment);
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(fillw, t);
    fillw(0);
    for (auto &x : th) x.join();
    // ---- consume the text of the records handed out
    for (int k = 0; k < f->n_files; ++k) {
        const char *next;
        if (per < f->n_recs[k]) next = f->recs[k][per].name - 1;      // the header character of the first record left over
        else next = upto[k];
        f->s[k].consumed(next);
        if (f->s[k].plain) f->s[k].eof = false;               // plain mappings are "filled" again on the next call
    }
    out->n_reads = n; out->seq = (const uint8_t *)O.seq.p; out->seq_off = seq_off; out->qual = any_qual ? (const uint8_t *)O.qual.p : nullptr;
    out->name = O.name.p; out->name_off = name_off; out->comment = any_com ? O.comment.p : nullptr; out->comment_off = any_com ? com_off : nullptr;
    return n;
}
