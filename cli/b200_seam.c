/* b200_seam.c -- the batch loop of `bwa mem` over libbwamem_b200.so: the three-step pipeline upstream bwa runs through
 * kt_pipeline (read a batch | align it | write its SAM records; the fork left the call commented out at fastmap.c:305,307
 * and calls its GPU driver superBatchMain instead, fastmap.c:311).  Step 1 is the library's mem_process_seqs
 * (bwamem.h:69): worker1, mem_pestat, mate rescue, worker2 and the SAM text on the GPU(s).  Built twice by
 * scripts/build_bwa.sh: against the library (-DB200_SEAM_GPU: `bwa`) and against the reference's own bwamem.c
 * (`bwa_cpu`, the CPU program the fork's CLI used to be), so that the two can be compared on the same command line.
 *
 * Environment: B200_DEVICES=<n> aligns every batch on n GPUs (contiguous blocks of whole pairs, index replicated). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "b200_seam.h"

void kt_pipeline(int n_threads, void *(*func)(void*, int, void*), void *shared_data, int n_steps);
extern int bwa_verbose;

#ifdef B200_SEAM_GPU
void b200_mem_process_seqs(const mem_opt_t *opt, const bwt_t *bwt, const bntseq_t *bns, const uint8_t *pac,
                           int64_t n_processed, int n, bseq1_t *seqs, const mem_pestat_t *pes0);
void b200_process_seqs_use_devices(int n_devices);
void b200_process_seqs_set_rg_id(const char *rg_id);
extern char bwa_rg_id[256];
#define ALIGN_BATCH b200_mem_process_seqs
#else
#define ALIGN_BATCH mem_process_seqs
#endif

typedef struct { int n_seqs; bseq1_t *seqs; } batch_t;

static void *step(void *shared, int which, void *in)
{
	ktp_aux_t *aux = (ktp_aux_t*)shared;
	batch_t *b = (batch_t*)in;
	int i;
	if (which == 0) {                                   /* read */
		int64_t size = 0;
		b = (batch_t*)calloc(1, sizeof(batch_t));
		b->seqs = bseq_read(aux->actual_chunk_size, &b->n_seqs, aux->ks, aux->ks2);
		if (b->seqs == 0) { free(b); return 0; }
		if (!aux->copy_comment)
			for (i = 0; i < b->n_seqs; ++i) { free(b->seqs[i].comment); b->seqs[i].comment = 0; }
		for (i = 0; i < b->n_seqs; ++i) size += b->seqs[i].l_seq;
		if (bwa_verbose >= 3) fprintf(stderr, "[M::%s] read %d sequences (%ld bp)...\n", __func__, b->n_seqs, (long)size);
		return b;
	}
	if (which == 1) {                                   /* align */
		const mem_opt_t *opt = aux->opt;
		const bwaidx_t *idx = aux->idx;
		if (opt->flag & MEM_F_SMARTPE) {                /* -p: interleaved input that may hold unpaired reads */
			bseq1_t *sep[2];
			int n_sep[2];
			mem_opt_t tmp_opt = *opt;
			bseq_classify(b->n_seqs, b->seqs, n_sep, sep);
			if (bwa_verbose >= 3) fprintf(stderr, "[M::%s] %d single-end sequences; %d paired-end sequences\n", __func__, n_sep[0], n_sep[1]);
			if (n_sep[0]) {
				tmp_opt.flag &= ~MEM_F_PE;
				ALIGN_BATCH(&tmp_opt, idx->bwt, idx->bns, idx->pac, aux->n_processed, n_sep[0], sep[0], 0);
				for (i = 0; i < n_sep[0]; ++i) b->seqs[sep[0][i].id].sam = sep[0][i].sam;
			}
			if (n_sep[1]) {
				tmp_opt.flag |= MEM_F_PE;
				ALIGN_BATCH(&tmp_opt, idx->bwt, idx->bns, idx->pac, aux->n_processed + n_sep[0], n_sep[1], sep[1], aux->pes0);
				for (i = 0; i < n_sep[1]; ++i) b->seqs[sep[1][i].id].sam = sep[1][i].sam;
			}
			free(sep[0]); free(sep[1]);
		} else ALIGN_BATCH(opt, idx->bwt, idx->bns, idx->pac, aux->n_processed, b->n_seqs, b->seqs, aux->pes0);
		aux->n_processed += b->n_seqs;
		return b;
	}
	for (i = 0; i < b->n_seqs; ++i) {                   /* write */
		if (b->seqs[i].sam) err_fputs(b->seqs[i].sam, stdout);
		free(b->seqs[i].name); free(b->seqs[i].comment); free(b->seqs[i].seq); free(b->seqs[i].qual); free(b->seqs[i].sam);
	}
	free(b->seqs); free(b);
	return 0;
}

void b200_batch_main(ktp_aux_t *aux)
{
#ifdef B200_SEAM_GPU
	const char *nd = getenv("B200_DEVICES");
	if (nd && atoi(nd) > 1) b200_process_seqs_use_devices(atoi(nd));
	b200_process_seqs_set_rg_id(bwa_rg_id);             /* -R (bwa_set_rg, bwa.c:559) */
#endif
	kt_pipeline(2, step, aux, 3);                       /* reading overlaps aligning, as upstream (no_mt_io would make it 1) */
}
