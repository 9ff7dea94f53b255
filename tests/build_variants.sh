#!/bin/bash
# Builds libhj with -D overrides of the HJ_TUNE_* constants (hj_api.cu) into hyrise-1_b200/csrc/variants/ for
# tests/sweep_variants.py. The .so files are git-ignored; they travel to the GPU box with the snapshot.
cd "$(dirname "$0")/.."
OUT=hyrise-1_b200/csrc/variants; mkdir -p $OUT
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -shared -I include"
build() { name=$1; shift
  /usr/local/cuda/bin/nvcc $FLAGS "$@" -Xptxas -v hyrise-1_b200/csrc/hj_api.cu -o $OUT/libhj_$name.so -ldl -lpthread -lrt > $OUT/$name.ptxas 2>&1 \
    && echo "built $name" || echo "FAILED $name (see $OUT/$name.ptxas)"; }
build p2_512x8_mb3 -DHJ_TUNE_P2_THREADS=512 -DHJ_TUNE_P2_ITEMS32=8 -DHJ_TUNE_P2_ITEMS64=4 -DHJ_TUNE_P2_MINBLOCKS=3 &
build p2_256x8_mb6 -DHJ_TUNE_P2_THREADS=256 -DHJ_TUNE_P2_ITEMS32=8 -DHJ_TUNE_P2_ITEMS64=4 -DHJ_TUNE_P2_MINBLOCKS=6 &
build p2_512x8_mb2 -DHJ_TUNE_P2_THREADS=512 -DHJ_TUNE_P2_ITEMS32=8 -DHJ_TUNE_P2_ITEMS64=4 -DHJ_TUNE_P2_MINBLOCKS=2 &
build probe_v2 -DHJ_TUNE_PROBE_V=2 &
wait
build probe_v8 -DHJ_TUNE_PROBE_V=8 &
build p1_v4_t2048 -DHJ_TUNE_ITEMS1_32=8 -DHJ_TUNE_P1_VIRTUAL=4 &
build p1_prefetch -DHJ_TUNE_P1_PREFETCH=1 &
wait
build p1_v2_t2048 -DHJ_TUNE_ITEMS1_32=8 -DHJ_TUNE_P1_VIRTUAL=2 &
build p1_v4_t2048_pf -DHJ_TUNE_ITEMS1_32=8 -DHJ_TUNE_P1_VIRTUAL=4 -DHJ_TUNE_P1_PREFETCH=1 &
build old_find_steps -DHJ_TUNE_BUCKET_FIND=0 -DHJ_TUNE_P2_STEPS=0 &
build p2_512x8_mb3_pf -DHJ_TUNE_P2_THREADS=512 -DHJ_TUNE_P2_ITEMS32=8 -DHJ_TUNE_P2_ITEMS64=4 -DHJ_TUNE_P2_MINBLOCKS=3 -DHJ_TUNE_P1_PREFETCH=1 &
wait
