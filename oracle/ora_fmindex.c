/* TEST INFRASTRUCTURE ONLY -- never linked into or called by the product library.
 *
 * Plain-C restatement of the FM-index side of the hot path, written from the algorithm (not copied): occurrence
 * counts over the reference's BWT layout, bi-directional extension, SMEM search and the three seeding passes.
 * Every function names the reference code it follows (file:line in sunwookim028/bwa-mem-gpu = bwa 0.7.17).
 * Pinned against the committed golden vectors (tests/golden/*.npz, generated from the unmodified reference) and,
 * where oracle/_ref exists, against the reference itself: tests/test_oracle.py.
 *
 * Index layout (bwtindex.c:149-172): the BWT is an array of 32-bit words; every 128 rows there is a block of 16
 * words = four 64-bit counts of A,C,G,T before the block, then 128 two-bit symbols, 16 per word, first symbol in
 * the two most significant bits.  Row numbers exclude the sentinel: rows >= primary are shifted down by one. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
	const uint32_t *bwt;
	uint64_t primary, seq_len, L2[5];
	const uint64_t *sa; uint64_t n_sa; int sa_intv;
} ora_index_t;

typedef struct { uint64_t x0, x1, x2, info; } ora_intv_t;   /* == bwtintv_t, bwt.h:62 */

void *ora_index_new(const uint32_t *bwt, uint64_t primary, const uint64_t *L2, uint64_t seq_len,
                    const uint64_t *sa, uint64_t n_sa, int sa_intv)
{
	ora_index_t *ix = calloc(1, sizeof(*ix));
	ix->bwt = bwt; ix->primary = primary; ix->seq_len = seq_len; ix->sa = sa; ix->n_sa = n_sa; ix->sa_intv = sa_intv;
	memcpy(ix->L2, L2, sizeof(ix->L2));
	return ix;
}
void ora_index_free(void *p) { free(p); }

static inline int bwt_symbol(const ora_index_t *ix, uint64_t row)            /* bwt_B0, bwt.h:46-51 */
{
	const uint32_t *blk = ix->bwt + (row >> 7) * 16 + 8;
	return (int)(blk[(row & 127) >> 4] >> ((~row & 15) << 1)) & 3;
}

/* number of symbols A,C,G,T in rows [0, k] of the BWT column; k = (uint64_t)-1 gives zeros (bwt_occ4, bwt.c:169-187) */
void ora_occ4(const void *p, uint64_t k, uint64_t cnt[4])
{
	const ora_index_t *ix = p;
	int c;
	if (k == (uint64_t)-1) { memset(cnt, 0, 4 * sizeof(uint64_t)); return; }
	if (k >= ix->primary) --k;                                               /* the sentinel row is not stored */
	{
		const uint64_t *chk = (const uint64_t *)(ix->bwt + (k >> 7) * 16);
		uint64_t r, first = k & ~(uint64_t)127;
		for (c = 0; c < 4; ++c) cnt[c] = chk[c];
		for (r = first; r <= k; ++r) ++cnt[bwt_symbol(ix, r)];
	}
}

/* bwt_2occ4 (bwt.c:189-220) is two bwt_occ4 calls that share a block when they can; results are the same */
void ora_2occ4(const void *p, uint64_t k, uint64_t l, uint64_t ck[4], uint64_t cl[4])
{
	ora_occ4(p, k, ck);
	ora_occ4(p, l, cl);
}

/* all four children of a bi-interval (bwt_extend, bwt.c:262-275).  back = 1 prepends a base (uses x0 as the BWT
 * range and shifts x1), back = 0 appends the complement (roles swapped). */
void ora_extend(const void *p, const ora_intv_t *ik, ora_intv_t ok[4], int back)
{
	const ora_index_t *ix = p;
	const uint64_t a = back ? ik->x0 : ik->x1, b = back ? ik->x1 : ik->x0;
	uint64_t tk[4], tl[4], nb[4];
	int c;
	ora_2occ4(p, a - 1, a - 1 + ik->x2, tk, tl);
	/* the other coordinate: children are laid out T, G, C, A after an optional sentinel */
	nb[3] = b + (a <= ix->primary && a + ik->x2 - 1 >= ix->primary);
	nb[2] = nb[3] + (tl[3] - tk[3]);
	nb[1] = nb[2] + (tl[2] - tk[2]);
	nb[0] = nb[1] + (tl[1] - tk[1]);
	for (c = 0; c < 4; ++c) {
		const uint64_t na = ix->L2[c] + 1 + tk[c];
		ok[c].x2 = tl[c] - tk[c];
		ok[c].x0 = back ? na : nb[c];
		ok[c].x1 = back ? nb[c] : na;
		ok[c].info = 0;
	}
}

/* suffix-array value of row k by walking LF to a sampled row (bwt_sa / bwt_invPsi, bwt.c:70-96) */
uint64_t ora_sa(const void *p, uint64_t k)
{
	const ora_index_t *ix = p;
	const uint64_t mask = (uint64_t)ix->sa_intv - 1;
	uint64_t steps = 0;
	while (k & mask) {
		++steps;
		if (k == ix->primary) k = 0;                                         /* the sentinel maps to row 0 */
		else {
			const uint64_t row = k > ix->primary ? k - 1 : k;
			const int c = bwt_symbol(ix, row);
			uint64_t cnt[4];
			ora_occ4(p, k, cnt);
			k = ix->L2[c] + cnt[c];
		}
	}
	return steps + ix->sa[k / ix->sa_intv];
}

static void set_base_intv(const ora_index_t *ix, int c, ora_intv_t *v)       /* bwt_set_intv, bwt.h:98 */
{
	v->x0 = ix->L2[c] + 1; v->x1 = ix->L2[3 - c] + 1; v->x2 = ix->L2[c + 1] - ix->L2[c]; v->info = 0;
}

/* Super-maximal exact matches covering position x (bwt_smem1a, bwt.c:289-351).  out receives at most cap
 * intervals ordered by start; info = start << 32 | end.  Returns the number written (or -1 on overflow);
 * *next = the position after the longest forward match. */
int ora_smem1a(const void *p, int len, const uint8_t *q, int x, int min_intv, uint64_t max_intv,
               int cap, ora_intv_t *out, int *next)
{
	const ora_index_t *ix = p;
	ora_intv_t *fw = malloc(sizeof(ora_intv_t) * (size_t)(len + 2)), *nx = malloc(sizeof(ora_intv_t) * (size_t)(len + 2));
	ora_intv_t cur, kids[4];
	int n_fw = 0, n_out = 0, i, j, ovf = 0;
	if (q[x] > 3) { *next = x + 1; free(fw); free(nx); return 0; }
	if (min_intv < 1) min_intv = 1;
	set_base_intv(ix, q[x], &cur);
	cur.info = (uint64_t)(x + 1);
	/* forward: remember the interval each time its size is about to change */
	for (i = x + 1; i < len; ++i) {
		if (cur.x2 < max_intv) { fw[n_fw++] = cur; break; }
		if (q[i] > 3) { fw[n_fw++] = cur; break; }
		ora_extend(p, &cur, kids, 0);
		{
			const ora_intv_t *k = &kids[3 - q[i]];
			if (k->x2 != cur.x2) {
				fw[n_fw++] = cur;
				if (k->x2 < (uint64_t)min_intv) break;
			}
			cur = *k; cur.info = (uint64_t)(i + 1);
		}
	}
	if (i == len) fw[n_fw++] = cur;
	for (j = 0; j < n_fw / 2; ++j) { ora_intv_t t = fw[j]; fw[j] = fw[n_fw - 1 - j]; fw[n_fw - 1 - j] = t; }   /* longest first */
	*next = (int)fw[0].info;
	/* backward: extend every remembered interval in lock step; an interval that cannot be extended is a SMEM
	 * unless a longer one from the same sweep still can */
	for (i = x - 1; i >= -1 && n_fw > 0; --i) {
		const int c = (i >= 0 && q[i] < 4) ? q[i] : -1;
		int n_nx = 0;
		for (j = 0; j < n_fw; ++j) {
			const ora_intv_t *pv = &fw[j];
			int dead = 1;
			/* the reference tests the size of the LAST interval it touched (`ik`) against max_intv here */
			if (c >= 0 && cur.x2 >= max_intv) {
				ora_extend(p, pv, kids, 1);
				dead = kids[c].x2 < (uint64_t)min_intv;
			}
			if (dead) {
				if (n_nx == 0 && (n_out == 0 || (uint64_t)(i + 1) < out[n_out - 1].info >> 32)) {
					cur = *pv; cur.info |= (uint64_t)(i + 1) << 32;
					if (n_out < cap) out[n_out] = cur; else ovf = 1;
					if (n_out < cap) ++n_out;
				}
			} else if (n_nx == 0 || kids[c].x2 != nx[n_nx - 1].x2) {
				kids[c].info = pv->info;
				nx[n_nx++] = kids[c];
			}
		}
		{ ora_intv_t *t = fw; fw = nx; nx = t; n_fw = n_nx; }
	}
	for (j = 0; j < n_out / 2; ++j) { ora_intv_t t = out[j]; out[j] = out[n_out - 1 - j]; out[n_out - 1 - j] = t; }
	free(fw); free(nx);
	return ovf ? -1 : n_out;
}

/* forward-only seed of at least min_len + 1 bases with fewer than max_intv hits (bwt_seed_strategy1, bwt.c:358-379) */
int ora_seed_strategy1(const void *p, int len, const uint8_t *q, int x, int min_len, int max_intv, ora_intv_t *mem)
{
	const ora_index_t *ix = p;
	ora_intv_t cur, kids[4];
	int i;
	memset(mem, 0, sizeof(*mem));
	if (q[x] > 3) return x + 1;
	set_base_intv(ix, q[x], &cur);
	for (i = x + 1; i < len; ++i) {
		if (q[i] > 3) return i + 1;
		ora_extend(p, &cur, kids, 0);
		cur = kids[3 - q[i]];
		if (cur.x2 < (uint64_t)max_intv && i - x >= min_len) {
			*mem = cur; mem->info = (uint64_t)x << 32 | (uint64_t)(i + 1);
			return i + 1;
		}
	}
	return len;
}

static int by_info(const void *a, const void *b)
{
	const uint64_t x = ((const ora_intv_t *)a)->info, y = ((const ora_intv_t *)b)->info;
	return x < y ? -1 : x > y;
}

/* The three seeding passes of mem_collect_intv (bwamem.c:137-185): all SMEMs, re-seeding of long SMEMs with few
 * hits from their middle, and the forward-only sweep; result sorted by info.  seq: nt4 codes.  Returns the number
 * of intervals (or -1 if cap is too small). */
int ora_collect_intv(const void *p, int len, const uint8_t *seq, int min_seed_len, float split_factor, int split_width,
                     int max_mem_intv, int cap, ora_intv_t *out)
{
	ora_intv_t *tmp = malloc(sizeof(ora_intv_t) * (size_t)(len + 2));
	const int split_len = (int)(min_seed_len * split_factor + .499);
	int n = 0, x = 0, k, m, next, old_n, ovf = 0;
	while (x < len) {                                                          /* pass 1 */
		if (seq[x] > 3) { ++x; continue; }
		m = ora_smem1a(p, len, seq, x, 1, 0, len + 2, tmp, &next);
		for (k = 0; k < m; ++k) {
			const int slen = (int)(uint32_t)tmp[k].info - (int)(tmp[k].info >> 32);
			if (slen >= min_seed_len) { if (n < cap) out[n++] = tmp[k]; else ovf = 1; }
		}
		x = next;
	}
	old_n = n;
	for (k = 0; k < old_n; ++k) {                                              /* pass 2 */
		const int start = (int)(out[k].info >> 32), end = (int)(uint32_t)out[k].info;
		int t;
		if (end - start < split_len || out[k].x2 > (uint64_t)split_width) continue;
		m = ora_smem1a(p, len, seq, (start + end) >> 1, (int)out[k].x2 + 1, 0, len + 2, tmp, &next);
		for (t = 0; t < m; ++t) {
			const int slen = (int)(uint32_t)tmp[t].info - (int)(tmp[t].info >> 32);
			if (slen >= min_seed_len) { if (n < cap) out[n++] = tmp[t]; else ovf = 1; }
		}
	}
	if (max_mem_intv > 0) {                                                    /* pass 3 */
		x = 0;
		while (x < len) {
			ora_intv_t mem;
			if (seq[x] > 3) { ++x; continue; }
			x = ora_seed_strategy1(p, len, seq, x, min_seed_len, max_mem_intv, &mem);
			if (mem.x2 > 0) { if (n < cap) out[n++] = mem; else ovf = 1; }
		}
	}
	qsort(out, (size_t)n, sizeof(ora_intv_t), by_info);
	free(tmp);
	return ovf ? -1 : n;
}
