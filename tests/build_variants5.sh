#!/bin/bash
# Round-2 candidates (never run on a GPU yet): how P1 reads dictionary entries (L1 bypass).
#   bash tests/build_variants5.sh && python tests/sweep_variants.py --scale 1.0 --runs default gather_cg gather_noalloc default
cd "$(dirname "$0")/.."
OUT=hyrise-1_b200/csrc/variants; mkdir -p $OUT; rm -f $OUT/*.so $OUT/*.ptxas
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -shared -I include"
build() { name=$1; shift
  /usr/local/cuda/bin/nvcc $FLAGS "$@" -Xptxas -v hyrise-1_b200/csrc/hj_api.cu -o $OUT/libhj_$name.so -ldl -lpthread -lrt > $OUT/$name.ptxas 2>&1 \
    && echo "built $name" || echo "FAILED $name (see $OUT/$name.ptxas)"; }
build gather_cg -DHJ_TUNE_P1_GATHER=1 &
build gather_noalloc -DHJ_TUNE_P1_GATHER=2 &
wait
