#!/bin/bash
# Round-2 candidates (never run on a GPU yet): how P1 reads dictionary entries (L1 bypass).
#   bash tests/build_variants5.sh && python tests/sweep_variants.py --scale 1.0 --runs default gather_cg gather_noalloc p2_mb3 p2_mb2 p2_smem p2_smem_mb3 p1_v8_t1024 bits_6_8 bits_8_6 default
cd "$(dirname "$0")/.."
OUT=hyrise-1_b200/csrc/variants; mkdir -p $OUT; rm -f $OUT/*.so $OUT/*.ptxas
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -shared -I include"
build() { name=$1; shift
  /usr/local/cuda/bin/nvcc $FLAGS "$@" -Xptxas -v hyrise-1_b200/csrc/hj_api.cu -o $OUT/libhj_$name.so -ldl -lpthread -lrt > $OUT/$name.ptxas 2>&1 \
    && echo "built $name" || echo "FAILED $name (see $OUT/$name.ptxas)"; }
build gather_cg -DHJ_TUNE_P1_GATHER=1 &
build gather_noalloc -DHJ_TUNE_P1_GATHER=2 &
# pass 2 with more registers per thread (3 / 2 CTAs of 256 x 16 per SM instead of 4): at 64 registers the compiler
# rematerialises the shared-memory base addresses inside every one of the 16 unrolled items (S2R SR_CgaCtaId + LEAs)
build p2_mb3 -DHJ_TUNE_P2_MINBLOCKS=3 &
build p2_mb2 -DHJ_TUNE_P2_MINBLOCKS=2 &
wait
# pass 1 with 8 tile streams of 1024 rows per SM (was 4 x 2048; pass 2's items are then only ~half full: read P1's time only)
build p1_v8_t1024 -DHJ_TUNE_P1_VIRTUAL32=8 -DHJ_TUNE_P1_TILE32=1024 -DHJ_TUNE_P1_SMALL_VT=1 &
# the 14 radix bits split 6 + 8 / 8 + 6 instead of 7 + 7 (fragment length in pass 1 vs run length in pass 2)
build bits_6_8 -DHJ_TUNE_BITS2_DELTA=1 &
build bits_8_6 -DHJ_TUNE_BITS2_DELTA=-1 &
wait
# pass 2's small shared arrays in one struct: 11 % fewer SASS instructions in an issue-bound kernel (8 bytes of spill)
build p2_smem -DHJ_TUNE_P2_SMEM=1 &
build p2_smem_mb3 -DHJ_TUNE_P2_SMEM=1 -DHJ_TUNE_P2_MINBLOCKS=3 &
wait
