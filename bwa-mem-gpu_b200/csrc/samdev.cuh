// SAM text on the device: mem_aln2sam (bwamem.c:832-956) with add_cigar (:819-830), the unmapped record of mem_reg2sam
// (:1043-1047) and the XA / XB strings of mem_gen_alt (bwamem_extra.c:143-161), from the arrays worker2 leaves on the device
// (b200_sel_t / b200_aln_t per region, CIGAR and MD pools).  The reference sketches the same step in
// cuda/bwamem_GPU.cu:3323 (SAMGEN_aln2sam_finegrain_kernel) and :3412 (SAMGEN_concatenate_kernel); here it is the two-pass
// shape of csrc/sam_format.cpp (the host formatter, kept as the checker): one templated formatter over a "sink", run
// once with a counting sink (bytes per read), then an exclusive scan places the reads, then run again with a writing sink
// straight into the final text.  One read per lane; the base and quality strings, which are most of a record, are
// written 4 characters per store when the destination is aligned.
#pragma once
#include "ctx.h"

struct SamDevParams {
    mem_opt_t opt;
    // the resident reads (nt4 codes after the pack kernel) and their text fields
    const u8 *seq; const i64 *off;              // batch-relative read r: seq[off[r] .. off[r+1])
    const u8 *qual;                             // same offsets, or null ('*')
    const char *names; const i64 *name_off;     // read r: names[name_off[r] .. name_off[r+1]) (not 0-terminated)
    const char *comments; const i64 *comment_off;   // optional (null), same shape; an empty slice = no comment
    const char *ann_names; const i32 *ann_name_off; // contig names, n_seqs + 1 offsets
    const char *ann_annos; const i32 *ann_anno_off; // optional: bntann1_t.anno for XR (MEM_F_REF_HDR)
    const char *rg; int rg_len;                 // bwa_rg_id or null
    // the chunk: reads r_lo .. r_hi of the batch, regions roff[r] - g0 .. of the chunk arrays
    int r_lo, n_reads; const i64 *roff; i64 g0;
    const b200_sel_t *sel; const b200_aln_t *aln; const u32 *cig; const char *md; i64 cig_base, md_base;
    // paired-end: per read of the batch the chunk-relative region whose alignment is the MATE's record h (bwamem_pair.c:366,
    // :398-407), -1 = mate unmapped (mem_reg2aln with ar = 0), null array = single-end
    const i32 *mate_reg;
    const i32 *read_flag;                       // paired-end: per read the extra flag of its UNMAPPED record (0x41 / 0x81 | proper pair); mapped
                                                // records carry theirs in b200_aln_t.flag; null = 0
    // output
    i32 *len;                                   // per read of the chunk (pass 1)
    const i64 *pos;                             // per read of the chunk: byte offset of its first record (pass 2)
    char *out;
    int *overflow;
};

int launch_sam_count(const SamDevParams &P, int n_sms, cudaStream_t st);
int launch_sam_write(const SamDevParams &P, int n_sms, cudaStream_t st);
