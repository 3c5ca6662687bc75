/* b200_seam.h -- what the reference's fastmap.c needs in place of "cuda/superbatch_process.h" when `bwa mem` is built
 * against libbwamem_b200.so: the ktp_aux_t main_mem fills (fastmap.c:59-311; declared by the fork in
 * cuda/superbatch_process.h:18-27, upstream in fastmap.c itself) and the batch driver that replaces superBatchMain
 * (cuda/superbatch_process.cpp:133), the fork's broken GPU driver.  No CUDA header is needed by the host program. */
#ifndef B200_SEAM_H_
#define B200_SEAM_H_

#include <locale.h>
#include <zlib.h>
#include "bwa.h"
#include "bwamem.h"
#include "kvec.h"
#include "utils.h"
#include "bntseq.h"
#include "kseq.h"
KSEQ_DECLARE(gzFile)

typedef struct {
	kseq_t *ks, *ks2;
	mem_opt_t *opt;
	mem_pestat_t *pes0;
	int64_t n_processed;
	int copy_comment, actual_chunk_size;
	bwaidx_t *idx;
	void *kmerHashTab;            /* the fork's 12-mer table: unused here (the library builds its own k-mer interval table) */
} ktp_aux_t;

#ifdef __cplusplus
extern "C" {
#endif
/* upstream's kt_pipeline(..., process, &aux, 3) (commented out at fastmap.c:305,307): read a batch of actual_chunk_size
 * bases with bseq_read, align it with mem_process_seqs, write the SAM records, until the input ends */
void b200_batch_main(ktp_aux_t *aux);
#ifdef __cplusplus
}
#endif
#endif
