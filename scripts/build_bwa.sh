#!/bin/bash
# Build the reference's `bwa` program against libbwamem_b200.so ("keep the bwa mem CLI"), and the same
# program against the reference's own CPU bwamem.c for comparison.
#   scripts/build_bwa.sh [reference dir, default /root/reference]  ->  cli/_build/bwa, cli/_build/bwa_cpu
# The reference's sources are compiled where they lie; only fastmap.c is COPIED into the git-ignored build directory and
# patched there (three lines at the seam, fastmap.c:35,309,311); nothing of the reference enters the repository.
set -euo pipefail
REF=${1:-/root/reference}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/cli/_build
mkdir -p "$OUT/obj"
CC=${CC:-gcc}
CFLAGS="-g -Wall -Wno-unused-function -Wno-unused-but-set-variable -O2 -DHAVE_PTHREAD -DUSE_MALLOC_WRAPPERS"     # Makefile:5-8
# the seam: include our header instead of the CUDA driver's, drop the 12-mer table load, call the batch loop
sed -e 's|#include "cuda/superbatch_process.h"|#include "b200_seam.h"|' \
    -e 's|^\(\s*\)aux.kmerHashTab = loadKMerIndex("kmers_index/hashTable");|\1aux.kmerHashTab = 0;|' \
    -e 's|^\(\s*\)superBatchMain(&aux);|\1b200_batch_main(\&aux);|' "$REF/fastmap.c" > "$OUT/fastmap_b200.c"
grep -q 'b200_batch_main(&aux);' "$OUT/fastmap_b200.c" && grep -q '"b200_seam.h"' "$OUT/fastmap_b200.c" || { echo "seam not found in $REF/fastmap.c" >&2; exit 1; }
# every C file of the program except the CUDA driver (Makefile:9-16 LOBJS / AOBJS without the cuda/ objects and loadKMerIndex)
SRC="utils kthread kstring ksw bwt bntseq bwa bwamem bwamem_pair bwamem_extra malloc_wrap QSufSort bwt_gen rope rle is bwtindex \
     bwashm bwase bwaseqio bwtgap bwtaln bamlite bwape kopen pemerge maxk bwtsw2_core bwtsw2_main bwtsw2_aux bwt_lite bwtsw2_chain bwtsw2_pair main"
OBJS=""
for f in $SRC; do
    if [ ! -f "$OUT/obj/$f.o" ] || [ "$REF/$f.c" -nt "$OUT/obj/$f.o" ]; then $CC -c $CFLAGS -I"$REF" "$REF/$f.c" -o "$OUT/obj/$f.o"; fi
    OBJS="$OBJS $OUT/obj/$f.o"
done
$CC -c $CFLAGS -I"$ROOT/cli" -I"$REF" "$OUT/fastmap_b200.c" -o "$OUT/obj/fastmap_b200.o"
$CC -c $CFLAGS -DB200_SEAM_GPU -I"$ROOT/cli" -I"$REF" "$ROOT/cli/b200_seam.c" -o "$OUT/obj/seam_gpu.o"
$CC -c $CFLAGS -I"$ROOT/cli" -I"$REF" "$ROOT/cli/b200_seam.c" -o "$OUT/obj/seam_cpu.o"
LIBDIR=$ROOT/bwa-mem-gpu_b200
# `bwa`: batches aligned by the library (its own copy of every device-side function; -rpath so the program finds it in-tree)
g++ -o "$OUT/bwa" $OBJS "$OUT/obj/fastmap_b200.o" "$OUT/obj/seam_gpu.o" -L"$LIBDIR" -lbwamem_b200 -Wl,-rpath,'$ORIGIN/../../bwa-mem-gpu_b200' \
    -L/usr/local/cuda/lib64 -Wl,-rpath,/usr/local/cuda/lib64 -lcudart -lm -lz -lpthread -lrt -ldl
# `bwa_cpu`: the same program with the reference's own mem_process_seqs (what upstream bwa is)
$CC -o "$OUT/bwa_cpu" $OBJS "$OUT/obj/fastmap_b200.o" "$OUT/obj/seam_cpu.o" -lm -lz -lpthread -lrt
echo "built $OUT/bwa and $OUT/bwa_cpu"
