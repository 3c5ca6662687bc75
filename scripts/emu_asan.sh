#!/bin/bash
# AddressSanitizer pass over the kernels: the SIMT-emulated build (the product's kernel sources compiled by g++, tests/emu)
# rebuilt with -fsanitize=address, and a set of parity tests run on it.  Every "global memory" access of a kernel is then a
# checked host access (device buffers are heap blocks in the emulator).  compute-sanitizer is closed on this GPU pool, so
# this is the memcheck the kernels get.  usage: scripts/emu_asan.sh [pytest -k expression]
set -euo pipefail
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/tests/emu/_build_asan
mkdir -p "$OUT"
CSRC=$ROOT/bwa-mem-gpu_b200/csrc
FLAGS="-O1 -g -fPIC -std=c++17 -DB200_EMU -I$ROOT/tests/emu -I$CSRC -fsanitize=address -fno-omit-frame-pointer -ffp-contract=off -w"
SRC=$(sed -n 's/^SRC *= *//p' "$CSRC/Makefile")
objs=""
for f in $SRC; do
  ( g++ $FLAGS -x c++ -c "$CSRC/$f" -o "$OUT/${f%.cu}.o" ) &
  objs="$objs $OUT/${f%.cu}.o"
done
for f in index_build process_seqs sam_format fastq; do ( g++ $FLAGS -c "$CSRC/$f.cpp" -o "$OUT/$f.o" ) & objs="$objs $OUT/$f.o"; done
( g++ $FLAGS -c "$ROOT/tests/emu/simt_emu.cpp" -o "$OUT/simt_emu.o" ) & objs="$objs $OUT/simt_emu.o"
wait
g++ -shared -fsanitize=address -o "$OUT/libbwamem_b200_emu.so" $objs -lpthread -ldl -lz
echo "built $OUT/libbwamem_b200_emu.so"
K=${1:-"golden or reads_to_sam or pairs_to_sam_two or reg2aln or collect_intv or flt_chained"}
cd "$ROOT"
B200_EMU_SO="$OUT/libbwamem_b200_emu.so" LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0:halt_on_error=1:abort_on_error=0 \
  python -m pytest tests/test_golden.py tests/test_stages.py tests/test_reads2sam.py -x -q -m "not gpu" -k "$K"
