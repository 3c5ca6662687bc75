/* TEST INFRASTRUCTURE ONLY -- never linked into the product library.
 *
 * Flat-array (ctypes-friendly) wrappers around the UNMODIFIED reference CPU
 * implementation.  This file is compiled together with the reference's own
 * sources, taken where they lie under /root/reference, into
 * oracle/_ref/libbwaref.so (see oracle/Makefile).  No reference source is
 * copied into this repository; only its public headers are #included at build
 * time.
 *
 * Seams exposed:
 *   (1) bwt_occ4 / bwt_2occ4 / bwt_extend / bwt_sa on arrays of inputs
 *   (2) bwt_smem1a per (read, x); mem_collect_intv output (through mem_chain)
 *   (3) mem_chain, mem_chain_flt outputs
 *   (4) ksw_extend2 / ksw_align2 / ksw_global2 batches
 *   (5) mem_align1_core -> mem_alnreg_t per read
 *   (7) mem_process_seqs -> SAM text
 *   (8) mem_reg2aln per region; worker2 single-end (mem_mark_primary_se, mem_approx_mapq_se, mem_reg2sam) -> regions + SAM text
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <stdint.h>
#include <zlib.h>
#include "bwa.h"
#include "bwamem.h"
#include "bwt.h"
#include "bntseq.h"
#include "ksw.h"
#include "kvec.h"
#include "utils.h"

extern void kt_for(int n_threads, void (*func)(void*, long, int), void *data, long n);
extern mem_alnreg_v mem_align1_core(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns,
                                    const uint8_t *pac, int l_seq, char *seq, void *buf);
extern mem_chain_v mem_chain(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns, int len, const uint8_t *seq, void *buf);
extern int mem_chain_flt(const mem_opt_t *opt, int n_chn, mem_chain_t *a);
extern void mem_flt_chained_seeds(const mem_opt_t *opt, const bntseq_t *bns, const uint8_t *pac,
                                  int l_query, const uint8_t *query, int n_chn, mem_chain_t *a);

/* ---------------------------------------------------------------- index */

int ref_index_build(const char *fa, const char *prefix)
{
	return bwa_idx_build(fa, prefix, BWTALGO_AUTO, 10000000);
}

void *ref_index_load(const char *prefix)
{
	return bwa_idx_load(prefix, BWA_IDX_ALL);
}

void ref_index_destroy(void *p) { bwa_idx_destroy((bwaidx_t*)p); }

/* out[0]=l_pac [1]=seq_len [2]=primary [3..7]=L2 [8]=bwt_size(words) [9]=n_sa
 * [10]=sa_intv [11]=n_seqs [12]=n_holes */
void ref_index_info(void *p, int64_t *out)
{
	bwaidx_t *idx = (bwaidx_t*)p;
	int i;
	out[0] = idx->bns->l_pac; out[1] = idx->bwt->seq_len; out[2] = idx->bwt->primary;
	for (i = 0; i < 5; ++i) out[3+i] = idx->bwt->L2[i];
	out[8] = idx->bwt->bwt_size; out[9] = idx->bwt->n_sa; out[10] = idx->bwt->sa_intv;
	out[11] = idx->bns->n_seqs; out[12] = idx->bns->n_holes;
}
const uint32_t *ref_index_bwt(void *p) { return ((bwaidx_t*)p)->bwt->bwt; }
const uint64_t *ref_index_sa(void *p)  { return ((bwaidx_t*)p)->bwt->sa; }
const uint8_t  *ref_index_pac(void *p) { return ((bwaidx_t*)p)->pac; }
const void *ref_index_bwt_struct(void *p) { return ((bwaidx_t*)p)->bwt; }
const void *ref_index_bns_struct(void *p) { return ((bwaidx_t*)p)->bns; }
void ref_index_anns(void *p, int64_t *offset, int32_t *len, int32_t *is_alt)
{
	bwaidx_t *idx = (bwaidx_t*)p;
	int i;
	for (i = 0; i < idx->bns->n_seqs; ++i) {
		offset[i] = idx->bns->anns[i].offset; len[i] = idx->bns->anns[i].len;
		is_alt[i] = idx->bns->anns[i].is_alt;
	}
}
const char *ref_index_ann_name(void *p, int i) { return ((bwaidx_t*)p)->bns->anns[i].name; }

/* ---------------------------------------------------------------- options */

void *ref_opt_init(void) { return mem_opt_init(); }
void ref_free(void *p) { free(p); }
void ref_fill_scmat(int a, int b, int8_t *mat) { bwa_fill_scmat(a, b, mat); }
int ref_sizeof(int what)
{
	switch (what) {
	case 0: return sizeof(mem_opt_t);
	case 1: return sizeof(mem_seed_t);
	case 2: return sizeof(mem_chain_t);
	case 3: return sizeof(mem_alnreg_t);
	case 4: return sizeof(bwtintv_t);
	case 5: return sizeof(bwt_t);
	case 6: return sizeof(bntseq_t);
	case 7: return sizeof(bntann1_t);
	case 8: return sizeof(bseq1_t);
	case 9: return sizeof(mem_pestat_t);
	case 10: return sizeof(kswr_t);
	}
	return -1;
}

/* ---------------------------------------------------------------- FM-index primitives */

void ref_occ4(void *p, int64_t n, const uint64_t *k, uint64_t *out)
{
	const bwt_t *bwt = ((bwaidx_t*)p)->bwt; int64_t i;
	for (i = 0; i < n; ++i) bwt_occ4(bwt, k[i], &out[4*i]);
}
void ref_2occ4(void *p, int64_t n, const uint64_t *k, const uint64_t *l, uint64_t *outk, uint64_t *outl)
{
	const bwt_t *bwt = ((bwaidx_t*)p)->bwt; int64_t i;
	for (i = 0; i < n; ++i) bwt_2occ4(bwt, k[i], l[i], &outk[4*i], &outl[4*i]);
}
/* ik: n x 3 (x0,x1,x2); out: n x 4 x 3 */
void ref_extend(void *p, int64_t n, const uint64_t *ik, int is_back, uint64_t *out)
{
	const bwt_t *bwt = ((bwaidx_t*)p)->bwt; int64_t i; int c;
	for (i = 0; i < n; ++i) {
		bwtintv_t in, ok[4];
		in.x[0] = ik[3*i]; in.x[1] = ik[3*i+1]; in.x[2] = ik[3*i+2]; in.info = 0;
		bwt_extend(bwt, &in, ok, is_back);
		for (c = 0; c < 4; ++c) {
			out[12*i + 3*c + 0] = ok[c].x[0]; out[12*i + 3*c + 1] = ok[c].x[1]; out[12*i + 3*c + 2] = ok[c].x[2];
		}
	}
}
void ref_sa(void *p, int64_t n, const uint64_t *k, uint64_t *out)
{
	const bwt_t *bwt = ((bwaidx_t*)p)->bwt; int64_t i;
	for (i = 0; i < n; ++i) out[i] = bwt_sa(bwt, k[i]);
}

/* one bwt_smem1a call; out: cap x 4 (x0,x1,x2,info); returns n (or -needed if cap too small) */
int ref_smem1a(void *p, int len, const uint8_t *q, int x, int min_intv, uint64_t max_intv, int cap, uint64_t *out, int *ret)
{
	const bwt_t *bwt = ((bwaidx_t*)p)->bwt;
	bwtintv_v mem = {0,0,0};
	int i, n;
	*ret = bwt_smem1a(bwt, len, q, x, min_intv, max_intv, &mem, 0);
	n = mem.n;
	if (n > cap) { free(mem.a); return -n; }
	for (i = 0; i < n; ++i) {
		out[4*i] = mem.a[i].x[0]; out[4*i+1] = mem.a[i].x[1]; out[4*i+2] = mem.a[i].x[2]; out[4*i+3] = mem.a[i].info;
	}
	free(mem.a);
	return n;
}

int ref_seed_strategy1(void *p, int len, const uint8_t *q, int x, int min_len, int max_intv, uint64_t *out)
{
	const bwt_t *bwt = ((bwaidx_t*)p)->bwt;
	bwtintv_t m; int r;
	r = bwt_seed_strategy1(bwt, len, q, x, min_len, max_intv, &m);
	out[0] = m.x[0]; out[1] = m.x[1]; out[2] = m.x[2]; out[3] = m.info;
	return r;
}

/* ---------------------------------------------------------------- per-read helpers */

static smem_aux_t *aux_new(void)
{
	smem_aux_t *a = calloc(1, sizeof(smem_aux_t));
	a->tmpv[0] = calloc(1, sizeof(bwtintv_v));
	a->tmpv[1] = calloc(1, sizeof(bwtintv_v));
	return a;
}
static void aux_del(smem_aux_t *a)
{
	free(a->tmpv[0]->a); free(a->tmpv[0]); free(a->tmpv[1]->a); free(a->tmpv[1]);
	free(a->mem.a); free(a->mem1.a); free(a);
}
static uint8_t *nt4_copy(const uint8_t *s, int l)
{
	uint8_t *r = malloc(l + 1); int i;
	for (i = 0; i < l; ++i) r[i] = s[i] < 4 ? s[i] : nst_nt4_table[s[i]];
	return r;
}

/* mem_collect_intv output for each read (reached through mem_chain with our own aux).
 * out_intv: cap x 4 u64; out_n[n_reads]; returns total or -1 on overflow */
int64_t ref_collect_intv(void *p, const void *opt_, int n_reads, const uint8_t *seqs, const int64_t *off,
                         int64_t cap, uint64_t *out_intv, int32_t *out_n)
{
	bwaidx_t *idx = (bwaidx_t*)p; const mem_opt_t *opt = (const mem_opt_t*)opt_;
	smem_aux_t *aux = aux_new();
	int64_t tot = 0; int r, i;
	for (r = 0; r < n_reads; ++r) {
		int l = (int)(off[r+1] - off[r]);
		uint8_t *s = nt4_copy(seqs + off[r], l);
		mem_chain_v chn;
		aux->mem.n = 0;
		chn = mem_chain(opt, idx->bwt, idx->bns, l, s, aux);
		for (i = 0; i < chn.n; ++i) free(chn.a[i].seeds);
		free(chn.a);
		out_n[r] = l < opt->min_seed_len ? 0 : aux->mem.n;
		if (tot + out_n[r] > cap) { free(s); aux_del(aux); return -1; }
		for (i = 0; i < out_n[r]; ++i, ++tot) {
			out_intv[4*tot] = aux->mem.a[i].x[0]; out_intv[4*tot+1] = aux->mem.a[i].x[1];
			out_intv[4*tot+2] = aux->mem.a[i].x[2]; out_intv[4*tot+3] = aux->mem.a[i].info;
		}
		free(s);
	}
	aux_del(aux);
	return tot;
}

/* mem_chain (stage: 0), + mem_chain_flt (stage: 1), + mem_flt_chained_seeds (stage: 2).
 * Per chain (8 x int64): pos, n, rid, w, kept, first, is_alt, frac_rep (float bits)
 * Per seed  (4 x int64): rbeg, qbeg, len, score
 * returns 0 or -1 on overflow */
int ref_chains(void *p, const void *opt_, int stage, int n_reads, const uint8_t *seqs, const int64_t *off,
               int64_t cap_chn, int64_t *out_chn, int32_t *out_nchn,
               int64_t cap_seed, int64_t *out_seed, int64_t *totals)
{
	bwaidx_t *idx = (bwaidx_t*)p; const mem_opt_t *opt = (const mem_opt_t*)opt_;
	int64_t nc = 0, ns = 0; int r, i, j, ovf = 0;
	for (r = 0; r < n_reads; ++r) {
		int l = (int)(off[r+1] - off[r]);
		uint8_t *s = nt4_copy(seqs + off[r], l);
		mem_chain_v chn = mem_chain(opt, idx->bwt, idx->bns, l, s, 0);
		if (stage >= 1) chn.n = mem_chain_flt(opt, chn.n, chn.a);
		if (stage >= 2) mem_flt_chained_seeds(opt, idx->bns, idx->pac, l, s, chn.n, chn.a);
		out_nchn[r] = chn.n;
		for (i = 0; i < chn.n; ++i) {
			mem_chain_t *c = &chn.a[i];
			if (nc >= cap_chn || ns + c->n > cap_seed) ovf = 1;
			if (!ovf) {
				int64_t *o = &out_chn[8*nc]; float fr = c->frac_rep; uint32_t fb;
				memcpy(&fb, &fr, 4);
				o[0] = c->pos; o[1] = c->n; o[2] = c->rid; o[3] = stage >= 1 ? c->w : 0;
				o[4] = stage >= 1 ? c->kept : 0; o[5] = stage >= 1 ? c->first : 0; o[6] = c->is_alt; o[7] = fb;
				for (j = 0; j < c->n; ++j) {
					int64_t *q = &out_seed[4*(ns+j)];
					q[0] = c->seeds[j].rbeg; q[1] = c->seeds[j].qbeg; q[2] = c->seeds[j].len; q[3] = c->seeds[j].score;
				}
			}
			++nc; ns += c->n;
			free(c->seeds);
		}
		free(chn.a); free(s);
	}
	totals[0] = nc; totals[1] = ns;
	return ovf ? -1 : 0;
}

/* ---------------------------------------------------------------- mem_align1_core over a batch (worker1) */

typedef struct {
	bwaidx_t *idx; const mem_opt_t *opt; const uint8_t *seqs; const int64_t *off;
	mem_alnreg_v *regs; smem_aux_t **aux;
} aln_task_t;

static void aln_worker(void *data, long i, int tid)
{
	aln_task_t *t = (aln_task_t*)data;
	int l = (int)(t->off[i+1] - t->off[i]);
	char *s = malloc(l + 1);
	memcpy(s, t->seqs + t->off[i], l); s[l] = 0;
	t->regs[i] = mem_align1_core(t->opt, t->idx->bwt, t->idx->bns, t->idx->pac, l, s, t->aux[tid]);
	free(s);
}

/* out_regs: raw mem_alnreg_t records (88 B each); returns total regs, or -(needed) on overflow.
 * If out_regs == NULL only counts (used for timing the worker1 path). */
int64_t ref_align(void *p, const void *opt_, int n_reads, const uint8_t *seqs, const int64_t *off, int n_threads,
                  int64_t cap, void *out_regs, int32_t *out_n)
{
	aln_task_t t; int64_t tot = 0; int i;
	t.idx = (bwaidx_t*)p; t.opt = (const mem_opt_t*)opt_; t.seqs = seqs; t.off = off;
	t.regs = calloc(n_reads, sizeof(mem_alnreg_v));
	if (n_threads < 1) n_threads = 1;
	t.aux = malloc(n_threads * sizeof(void*));
	for (i = 0; i < n_threads; ++i) t.aux[i] = aux_new();
	kt_for(n_threads, aln_worker, &t, n_reads);
	for (i = 0; i < n_threads; ++i) aux_del(t.aux[i]);
	free(t.aux);
	for (i = 0; i < n_reads; ++i) tot += t.regs[i].n;
	if (out_regs && tot <= cap) {
		mem_alnreg_t *o = (mem_alnreg_t*)out_regs;
		for (i = 0; i < n_reads; ++i) {
			if (t.regs[i].n) memcpy(o, t.regs[i].a, t.regs[i].n * sizeof(mem_alnreg_t));
			o += t.regs[i].n;
		}
	}
	for (i = 0; i < n_reads; ++i) { if (out_n) out_n[i] = t.regs[i].n; free(t.regs[i].a); }
	free(t.regs);
	return (out_regs && tot > cap) ? -tot : tot;
}

/* ---------------------------------------------------------------- DP primitives */

/* out: n x 6: score, qle, tle, gtle, gscore, max_off */
void ref_ksw_extend2(int64_t n, const int32_t *qlen, const int64_t *qoff, const uint8_t *q,
                     const int32_t *tlen, const int64_t *toff, const uint8_t *t,
                     const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins,
                     const int32_t *w, int end_bonus, int zdrop, const int32_t *h0, int32_t *out)
{
	int64_t i;
	for (i = 0; i < n; ++i) {
		int qle, tle, gtle, gscore, max_off;
		int sc = ksw_extend2(qlen[i], q + qoff[i], tlen[i], t + toff[i], 5, mat, o_del, e_del, o_ins, e_ins,
		                     w[i], end_bonus, zdrop, h0[i], &qle, &tle, &gtle, &gscore, &max_off);
		int32_t *o = &out[6*i];
		o[0] = sc; o[1] = qle; o[2] = tle; o[3] = gtle; o[4] = gscore; o[5] = max_off;
	}
}

/* out: n x 7: score, te, qe, score2, te2, tb, qb */
void ref_ksw_align2(int64_t n, const int32_t *qlen, const int64_t *qoff, const uint8_t *q,
                    const int32_t *tlen, const int64_t *toff, const uint8_t *t,
                    const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins, const int32_t *xtra, int32_t *out)
{
	int64_t i;
	for (i = 0; i < n; ++i) {
		kswr_t r = ksw_align2(qlen[i], (uint8_t*)q + qoff[i], tlen[i], (uint8_t*)t + toff[i], 5, mat,
		                      o_del, e_del, o_ins, e_ins, xtra[i], 0);
		int32_t *o = &out[7*i];
		o[0] = r.score; o[1] = r.te; o[2] = r.qe; o[3] = r.score2; o[4] = r.te2; o[5] = r.tb; o[6] = r.qb;
	}
}

/* score only (cigar==NULL) or with cigar: out_cigar cap words per problem */
void ref_ksw_global2(int64_t n, const int32_t *qlen, const int64_t *qoff, const uint8_t *q,
                     const int32_t *tlen, const int64_t *toff, const uint8_t *t,
                     const int8_t *mat, int o_del, int e_del, int o_ins, int e_ins, const int32_t *w,
                     int32_t *out_score, int cap_cigar, uint32_t *out_cigar, int32_t *out_ncigar)
{
	int64_t i;
	for (i = 0; i < n; ++i) {
		int nc = 0; uint32_t *cg = 0;
		out_score[i] = ksw_global2(qlen[i], q + qoff[i], tlen[i], t + toff[i], 5, mat, o_del, e_del, o_ins, e_ins,
		                           w[i], out_cigar ? &nc : 0, out_cigar ? &cg : 0);
		if (out_cigar) {
			out_ncigar[i] = nc;
			if (nc <= cap_cigar) memcpy(out_cigar + i * cap_cigar, cg, nc * 4);
			free(cg);
		}
	}
}

/* ---------------------------------------------------------------- full mem_process_seqs -> SAM */

/* names: NUL-separated list of n names; quals may be NULL.  SAM records are concatenated into out
 * (cap bytes). pes0 may be NULL.  Returns bytes written or -(needed). */
int64_t ref_process_seqs(void *p, const void *opt_, int64_t n_processed, int n, const char *names,
                         const uint8_t *seqs, const int64_t *off, const uint8_t *quals,
                         const void *pes0, int64_t cap, char *out)
{
	bwaidx_t *idx = (bwaidx_t*)p; const mem_opt_t *opt = (const mem_opt_t*)opt_;
	bseq1_t *bs = calloc(n, sizeof(bseq1_t));
	const char *nm = names; int64_t tot = 0; int i;
	for (i = 0; i < n; ++i) {
		int l = (int)(off[i+1] - off[i]);
		bs[i].name = strdup(nm); nm += strlen(nm) + 1;
		bs[i].seq = malloc(l + 1); memcpy(bs[i].seq, seqs + off[i], l); bs[i].seq[l] = 0;
		if (quals) { bs[i].qual = malloc(l + 1); memcpy(bs[i].qual, quals + off[i], l); bs[i].qual[l] = 0; }
		bs[i].l_seq = l; bs[i].id = i;
	}
	mem_process_seqs(opt, idx->bwt, idx->bns, idx->pac, n_processed, n, bs, (const mem_pestat_t*)pes0);
	for (i = 0; i < n; ++i) tot += strlen(bs[i].sam);
	if (tot <= cap) {
		char *o = out;
		for (i = 0; i < n; ++i) { size_t l = strlen(bs[i].sam); memcpy(o, bs[i].sam, l); o += l; }
	}
	for (i = 0; i < n; ++i) { free(bs[i].name); free(bs[i].seq); free(bs[i].qual); free(bs[i].sam); }
	free(bs);
	return tot <= cap ? tot : -tot;
}

/* the same with FASTQ comments (NUL-separated list, an empty string = no comment; may be NULL) and a read group id
 * (bwa_rg_id, bwa.c:44; may be NULL) */
extern char bwa_rg_id[256];
static double ref_last_s = 0;
/* wall seconds of the last mem_process_seqs call itself (what its own "[M::mem_process_seqs] ... real sec" line reports) */
double ref_last_process_seqs_seconds(void) { return ref_last_s; }
int64_t ref_process_seqs2(void *p, const void *opt_, int64_t n_processed, int n, const char *names, const char *comments,
                          const uint8_t *seqs, const int64_t *off, const uint8_t *quals, const char *rg,
                          const void *pes0, int64_t cap, char *out)
{
	bwaidx_t *idx = (bwaidx_t*)p; const mem_opt_t *opt = (const mem_opt_t*)opt_;
	bseq1_t *bs = calloc(n, sizeof(bseq1_t));
	const char *nm = names, *cm = comments; int64_t tot = 0; int i;
	for (i = 0; i < n; ++i) {
		int l = (int)(off[i+1] - off[i]);
		bs[i].name = strdup(nm); nm += strlen(nm) + 1;
		if (cm) { if (*cm) bs[i].comment = strdup(cm); cm += strlen(cm) + 1; }
		bs[i].seq = malloc(l + 1); memcpy(bs[i].seq, seqs + off[i], l); bs[i].seq[l] = 0;
		if (quals) { bs[i].qual = malloc(l + 1); memcpy(bs[i].qual, quals + off[i], l); bs[i].qual[l] = 0; }
		bs[i].l_seq = l; bs[i].id = i;
	}
	memset(bwa_rg_id, 0, 256);
	if (rg) strncpy(bwa_rg_id, rg, 255);
	{
		struct timespec t0, t1;
		clock_gettime(CLOCK_MONOTONIC, &t0);
		mem_process_seqs(opt, idx->bwt, idx->bns, idx->pac, n_processed, n, bs, (const mem_pestat_t*)pes0);
		clock_gettime(CLOCK_MONOTONIC, &t1);
		ref_last_s = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
	}
	memset(bwa_rg_id, 0, 256);
	for (i = 0; i < n; ++i) tot += strlen(bs[i].sam);
	if (tot <= cap) {
		char *o = out;
		for (i = 0; i < n; ++i) { size_t l = strlen(bs[i].sam); memcpy(o, bs[i].sam, l); o += l; }
	}
	for (i = 0; i < n; ++i) { free(bs[i].name); free(bs[i].comment); free(bs[i].seq); free(bs[i].qual); free(bs[i].sam); }
	free(bs);
	return tot <= cap ? tot : -tot;
}

/* ---------------------------------------------------------------- paired-end: pestat and the mate-rescue loop */

extern int mem_matesw(const mem_opt_t *opt, const bntseq_t *bns, const uint8_t *pac, const mem_pestat_t pes[4],
                      const mem_alnreg_t *a, int l_ms, const uint8_t *ms, mem_alnreg_v *ma);

static mem_alnreg_v *regs_unflatten(int n_reads, const void *regs, const int32_t *cnt)
{
	mem_alnreg_v *v = calloc(n_reads > 0 ? n_reads : 1, sizeof(mem_alnreg_v));
	const mem_alnreg_t *r = (const mem_alnreg_t*)regs; int i;
	for (i = 0; i < n_reads; ++i) {
		v[i].n = v[i].m = cnt[i];
		v[i].a = malloc((cnt[i] > 0 ? cnt[i] : 1) * sizeof(mem_alnreg_t));
		memcpy(v[i].a, r, cnt[i] * sizeof(mem_alnreg_t)); r += cnt[i];
	}
	return v;
}

/* mem_pestat (bwamem_pair.c:72) over flat region lists; pes_out = 4 x mem_pestat_t */
void ref_pestat(void *p, const void *opt_, int n_reads, const void *regs, const int32_t *cnt, void *pes_out)
{
	bwaidx_t *idx = (bwaidx_t*)p; int i;
	mem_alnreg_v *v = regs_unflatten(n_reads, regs, cnt);
	mem_pestat((const mem_opt_t*)opt_, idx->bns->l_pac, n_reads, v, (mem_pestat_t*)pes_out);
	for (i = 0; i < n_reads; ++i) free(v[i].a);
	free(v);
}

/* The rescue block of mem_sam_pe (bwamem_pair.c:291-302), verbatim control flow around the reference's
 * mem_matesw, for every pair.  seqs: nt4 codes.  Returns total regions or -(needed); *n_sw = sum of mem_matesw returns. */
int64_t ref_mate_rescue(void *p, const void *opt_, const void *pes_, int n_reads, const uint8_t *seqs, const int64_t *off,
                        const void *regs, const int32_t *cnt, int64_t cap, void *out_regs, int32_t *out_n, int64_t *n_sw)
{
	bwaidx_t *idx = (bwaidx_t*)p; const mem_opt_t *opt = (const mem_opt_t*)opt_;
	const mem_pestat_t *pes = (const mem_pestat_t*)pes_;
	mem_alnreg_v *v = regs_unflatten(n_reads, regs, cnt);
	int64_t tot = 0, nsw = 0; int pr, i, j;
	for (pr = 0; pr + 1 < n_reads; pr += 2) {
		mem_alnreg_v *a = &v[pr], b[2];
		kv_init(b[0]); kv_init(b[1]);
		for (i = 0; i < 2; ++i)
			for (j = 0; j < a[i].n; ++j)
				if (a[i].a[j].score >= a[i].a[0].score - opt->pen_unpaired)
					kv_push(mem_alnreg_t, b[i], a[i].a[j]);
		for (i = 0; i < 2; ++i)
			for (j = 0; j < b[i].n && j < opt->max_matesw; ++j)
				nsw += mem_matesw(opt, idx->bns, idx->pac, pes, &b[i].a[j], (int)(off[pr + !i + 1] - off[pr + !i]),
				                  seqs + off[pr + !i], &a[!i]);
		free(b[0].a); free(b[1].a);
	}
	for (i = 0; i < n_reads; ++i) tot += v[i].n;
	if (tot <= cap) {
		mem_alnreg_t *o = (mem_alnreg_t*)out_regs;
		for (i = 0; i < n_reads; ++i) { memcpy(o, v[i].a, v[i].n * sizeof(mem_alnreg_t)); o += v[i].n; }
	}
	for (i = 0; i < n_reads; ++i) { out_n[i] = v[i].n; free(v[i].a); }
	free(v);
	if (n_sw) *n_sw = nsw;
	return tot <= cap ? tot : -tot;
}

void ref_set_verbose(int v) { bwa_verbose = v; }

/* ---------------------------------------------------------------- index from in-memory arrays
 * Wraps caller-owned arrays (e.g. the ones produced by the product's own index builder, which the tests
 * prove byte-identical to bwa_idx_build's) into the reference's bwaidx_t so that the CPU baseline does not
 * have to re-run `bwa index` on large synthetic genomes.  Arrays must outlive the handle. */
void *ref_index_from_arrays(const uint32_t *bwt_words, uint64_t bwt_size, uint64_t primary, const uint64_t *L2,
                            uint64_t seq_len, const uint64_t *sa, uint64_t n_sa, int sa_intv, const uint8_t *pac,
                            int64_t l_pac, int n_seqs, const int64_t *ann_off, const int32_t *ann_len)
{
	bwaidx_t *idx = calloc(1, sizeof(bwaidx_t));
	bwt_t *bwt = calloc(1, sizeof(bwt_t));
	bntseq_t *bns = calloc(1, sizeof(bntseq_t));
	int i;
	bwt->primary = primary; bwt->seq_len = seq_len; bwt->bwt_size = bwt_size;
	for (i = 0; i < 5; ++i) bwt->L2[i] = L2[i];
	bwt->bwt = (uint32_t*)bwt_words; bwt->sa = (bwtint_t*)sa; bwt->n_sa = n_sa; bwt->sa_intv = sa_intv;
	bwt_gen_cnt_table(bwt);
	bns->l_pac = l_pac; bns->n_seqs = n_seqs; bns->seed = 11;
	bns->anns = calloc(n_seqs, sizeof(bntann1_t));
	for (i = 0; i < n_seqs; ++i) {
		char nm[32];
		snprintf(nm, sizeof(nm), "chr%d", i + 1);
		bns->anns[i].offset = ann_off[i]; bns->anns[i].len = ann_len[i];
		bns->anns[i].name = strdup(nm); bns->anns[i].anno = strdup("");
	}
	idx->bwt = bwt; idx->bns = bns; idx->pac = (uint8_t*)pac;
	return idx;
}

void ref_index_release_wrapped(void *p)
{
	bwaidx_t *idx = (bwaidx_t*)p;
	int i;
	for (i = 0; i < idx->bns->n_seqs; ++i) { free(idx->bns->anns[i].name); free(idx->bns->anns[i].anno); }
	free(idx->bns->anns); free(idx->bns); free(idx->bwt); free(idx);
}

/* ---------------------------------------------------------------- mem_reg2aln (bwamem.c:1099-1169) for every region of a batch.
 * aln: per region 8 x int64: pos, rid, is_rev, n_cigar, NM, score, sub, l_md (MD length without the terminating 0);
 * cigar / md: concatenated over the regions in order.  Returns the number of regions, or -1 when a pool is too small
 * (totals[0..1] = words / bytes needed). */
int64_t ref_reg2aln(void *p, const void *opt_, int n_reads, const uint8_t *seqs, const int64_t *off, const void *regs_,
                    const int32_t *cnt, int64_t *aln, int64_t cap_cigar, uint32_t *cigar, int64_t cap_md, char *md, int64_t *totals)
{
	bwaidx_t *idx = (bwaidx_t*)p;
	const mem_opt_t *opt = (const mem_opt_t*)opt_;
	const mem_alnreg_t *regs = (const mem_alnreg_t*)regs_;
	int64_t g = 0, nc = 0, nm = 0;
	int r, k, ovf = 0;
	for (r = 0; r < n_reads; ++r) {
		int l = (int)(off[r + 1] - off[r]);
		for (k = 0; k < cnt[r]; ++k, ++g) {
			mem_aln_t a = mem_reg2aln(opt, idx->bns, idx->pac, l, (const char*)(seqs + off[r]), &regs[g]);
			int l_md = a.cigar ? (int)strlen((char*)(a.cigar + a.n_cigar)) : 0;
			int64_t *o = aln + 8 * g;
			o[0] = a.pos; o[1] = a.rid; o[2] = a.is_rev; o[3] = a.n_cigar; o[4] = a.NM; o[5] = a.score; o[6] = a.sub; o[7] = l_md;
			if (nc + a.n_cigar > cap_cigar || nm + l_md > cap_md) ovf = 1;
			if (!ovf) {
				if (a.n_cigar) memcpy(cigar + nc, a.cigar, (size_t)a.n_cigar * 4);
				if (l_md) memcpy(md + nm, (char*)(a.cigar + a.n_cigar), (size_t)l_md);
			}
			nc += a.n_cigar; nm += l_md;
			free(a.cigar);
		}
	}
	totals[0] = nc; totals[1] = nm;
	return ovf ? -1 : g;
}

/* (8) worker2, single-end (bwamem.c:1195-1205): mem_mark_primary_se [+ mem_reorder_primary5 with MEM_F_PRIMARY5] on each
 * read's regions -- written back in place -- mem_approx_mapq_se of every region (mapq[]), then mem_reg2sam; the SAM
 * text of the reads is concatenated into sam.  seqs are nt4 codes (what worker1 leaves in bseq1_t.seq); reads are named
 * r<global index>, no qualities.  Returns the number of SAM bytes (nothing is written beyond cap_sam). */
extern int mem_mark_primary_se(const mem_opt_t *opt, int n, mem_alnreg_t *a, int64_t id);
extern void mem_reorder_primary5(int T, mem_alnreg_v *a);
extern int mem_approx_mapq_se(const mem_opt_t *opt, const mem_alnreg_t *a);
extern void mem_reg2sam(const mem_opt_t *opt, const bntseq_t *bns, const uint8_t *pac, bseq1_t *s, mem_alnreg_v *a, int extra_flag, const mem_aln_t *m);

int64_t ref_reg2sam_se(void *p, const void *opt_, int n_reads, int64_t n_processed, const uint8_t *seqs, const int64_t *off,
                       void *regs_, const int32_t *cnt, int32_t *mapq, int64_t cap_sam, char *sam)
{
	bwaidx_t *idx = (bwaidx_t*)p;
	const mem_opt_t *opt = (const mem_opt_t*)opt_;
	mem_alnreg_t *regs = (mem_alnreg_t*)regs_;
	int64_t g = 0, ns = 0;
	int r, k;
	for (r = 0; r < n_reads; ++r) {
		int l = (int)(off[r + 1] - off[r]);
		mem_alnreg_v av;
		bseq1_t s;
		char name[32];
		size_t ls;
		av.n = av.m = cnt[r];
		av.a = (mem_alnreg_t*)malloc((size_t)(cnt[r] > 0 ? cnt[r] : 1) * sizeof(mem_alnreg_t));
		memcpy(av.a, regs + g, (size_t)cnt[r] * sizeof(mem_alnreg_t));
		mem_mark_primary_se(opt, (int)av.n, av.a, n_processed + r);
		if (opt->flag & MEM_F_PRIMARY5) mem_reorder_primary5(opt->T, &av);
		memcpy(regs + g, av.a, (size_t)cnt[r] * sizeof(mem_alnreg_t));
		for (k = 0; k < cnt[r]; ++k) mapq[g + k] = mem_approx_mapq_se(opt, &av.a[k]);
		memset(&s, 0, sizeof(s));
		snprintf(name, sizeof(name), "r%lld", (long long)(n_processed + r));
		s.name = name; s.l_seq = l;
		s.seq = (char*)malloc((size_t)l + 1);
		memcpy(s.seq, seqs + off[r], (size_t)l);
		mem_reg2sam(opt, idx->bns, idx->pac, &s, &av, 0, 0);
		ls = strlen(s.sam);
		if (ns + (int64_t)ls <= cap_sam) memcpy(sam + ns, s.sam, ls);
		ns += (int64_t)ls;
		free(s.sam); free(s.seq); free(av.a);
		g += cnt[r];
	}
	return ns;
}
