/* TEST INFRASTRUCTURE ONLY -- never linked into or called by the product library.
 *
 * Plain-C restatement of the seed-extension dynamic programme of the hot path: ksw_extend2 (ksw.c:380-479) -- affine
 * gaps with separate insertion / deletion penalties, a band that is clamped to what the scores allow (ksw.c:399-407),
 * zero-trimmed row ranges (ksw.c:466-469), the z-drop stop (ksw.c:458-464) and the to-end score (ksw.c:450-453).
 * Written from the recurrences with its own data layout (two plain arrays instead of the reference's packed struct
 * and query profile); pinned against tests/golden/ksw_extend2.npz and, where available, oracle/_ref
 * (tests/test_oracle.py).
 *
 * State per query column j while row i is processed:  Hd[j] = H(i-1, j-1) and Ev[j] = E(i, j).
 * A row first reads them, then overwrites Hd[j] with H(i, j-1) and Ev[j] with E(i+1, j).  Entries outside the row's
 * column range keep whatever an earlier row left there -- later rows do read them when the range widens again, which
 * is part of the observable behaviour. */
#include <stdint.h>
#include <stdlib.h>

static int imax(int a, int b) { return a > b ? a : b; }
static int imin(int a, int b) { return a < b ? a : b; }

/* out[6] = score, qle, tle, gtle, gscore, max_off (the return value and the five out-parameters of ksw_extend2) */
void ora_ksw_extend2(int qlen, const uint8_t *query, int tlen, const uint8_t *target, const int8_t *mat /* 5x5 */,
                     int o_del, int e_del, int o_ins, int e_ins, int w, int end_bonus, int zdrop, int h0, int32_t out[6])
{
	int *Hd = calloc((size_t)qlen + 2, sizeof(int)), *Ev = calloc((size_t)qlen + 2, sizeof(int));
	const int open_del = o_del + e_del, open_ins = o_ins + e_ins;
	int best = h0, best_i = -1, best_j = -1, end_i = -1, end_score = -1, max_off = 0;
	int lo = 0, hi = qlen, i, j, top = 0;

	/* row "-1": only insertions into the query from the seed score (ksw.c:395-397) */
	Hd[0] = h0;
	if (qlen >= 1) Hd[1] = h0 > open_ins ? h0 - open_ins : 0;
	for (j = 2; j <= qlen && Hd[j - 1] > e_ins; ++j) Hd[j] = Hd[j - 1] - e_ins;

	/* the band cannot usefully exceed the longest gap the best possible score pays for (ksw.c:399-407) */
	for (i = 0; i < 25; ++i) top = imax(top, mat[i]);
	w = imin(w, imax(1, (int)((double)(qlen * top + end_bonus - o_ins) / e_ins + 1.)));
	w = imin(w, imax(1, (int)((double)(qlen * top + end_bonus - o_del) / e_del + 1.)));

	for (i = 0; i < tlen; ++i) {
		const int8_t *srow = mat + 5 * target[i];
		int left, F = 0, row_best = 0, row_best_j = -1;
		lo = imax(lo, i - w);
		hi = imin(imin(hi, i + w + 1), qlen);
		left = lo == 0 ? imax(0, h0 - (o_del + e_del * (i + 1))) : 0;          /* H(i, lo-1): only deletions from the seed */
		for (j = lo; j < hi; ++j) {
			const int diag = Hd[j], E = Ev[j];
			/* a zero diagonal is a dead path: it may not be revived by a match (ksw.c:433) */
			const int M = diag ? diag + srow[query[j]] : 0;
			const int H = imax(imax(M, E), F);
			Hd[j] = left;                                                          /* becomes H(i, j-1) for row i+1 */
			left = H;
			if (H >= row_best) { row_best = H; row_best_j = j; }                   /* ties: the last column (ksw.c:437-438) */
			/* gaps are opened from M, not from H: a deletion may not follow an insertion directly */
			Ev[j] = imax(E - e_del, imax(0, M - open_del));
			F = imax(F - e_ins, imax(0, M - open_ins));
		}
		Hd[hi] = left; Ev[hi] = 0;
		j = lo < hi ? hi : lo;                                                     /* where the reference's loop variable stops */
		if (j == qlen) {                                                           /* the query end was reached in this row */
			if (!(end_score > left)) end_i = i;
			end_score = imax(end_score, left);
		}
		if (row_best == 0) break;
		if (row_best > best) {
			best = row_best; best_i = i; best_j = row_best_j;
			max_off = imax(max_off, abs(row_best_j - i));
		} else if (zdrop > 0) {
			const int di = i - best_i, dj = row_best_j - best_j;
			const int gap_cost = di > dj ? (di - dj) * e_del : (dj - di) * e_ins;
			if (best - row_best - gap_cost > zdrop) break;
		}
		/* next row: drop the all-zero columns at both ends of [lo, hi] (ksw.c:466-469) */
		for (j = lo; j < hi && Hd[j] == 0 && Ev[j] == 0; ++j) {}
		lo = j;
		for (j = hi; j >= lo && Hd[j] == 0 && Ev[j] == 0; --j) {}
		hi = imin(j + 2, qlen);
	}
	out[0] = best; out[1] = best_j + 1; out[2] = best_i + 1; out[3] = end_i + 1; out[4] = end_score; out[5] = max_off;
	free(Hd); free(Ev);
}

void ora_ksw_extend2_batch(int64_t n, const int32_t *qlen, const int64_t *qoff, const uint8_t *q, const int32_t *tlen,
                           const int64_t *toff, const uint8_t *t, const int8_t *mat, int o_del, int e_del, int o_ins,
                           int e_ins, const int32_t *w, int end_bonus, int zdrop, const int32_t *h0, int32_t *out)
{
	int64_t i;
	for (i = 0; i < n; ++i)
		ora_ksw_extend2(qlen[i], q + qoff[i], tlen[i], t + toff[i], mat, o_del, e_del, o_ins, e_ins, w[i], end_bonus, zdrop,
		                h0[i], out + 6 * i);
}
