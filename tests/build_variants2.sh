#!/bin/bash
# Second tuning round (pass 2 with 2048-row slabs): owner-search steps, owner-table granularity, item fill.
cd "$(dirname "$0")/.."
OUT=hyrise-1_b200/csrc/variants; mkdir -p $OUT; rm -f $OUT/*.so $OUT/*.ptxas
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -shared -I include"
build() { name=$1; shift
  /usr/local/cuda/bin/nvcc $FLAGS "$@" -Xptxas -v hyrise-1_b200/csrc/hj_api.cu -o $OUT/libhj_$name.so -ldl -lpthread -lrt > $OUT/$name.ptxas 2>&1 \
    && echo "built $name" || echo "FAILED $name (see $OUT/$name.ptxas)"; }
build s1 -DHJ_TUNE_P2_STEPS=1 &
build s2 -DHJ_TUNE_P2_STEPS=2 &
build s3 -DHJ_TUNE_P2_STEPS=3 &
build own4 -DHJ_TUNE_P2_OWN_SHIFT=4 &
wait
build own4_s1 -DHJ_TUNE_P2_OWN_SHIFT=4 -DHJ_TUNE_P2_STEPS=1 &
build own3 -DHJ_TUNE_P2_OWN_SHIFT=3 &
build fill16 -DHJ_TUNE_P2_FILL16=16 &
build fill12 -DHJ_TUNE_P2_FILL16=12 &
wait
build fill24 -DHJ_TUNE_P2_FILL16=24 &
build fill28_s2 -DHJ_TUNE_P2_FILL16=28 -DHJ_TUNE_P2_STEPS=2 &
build own4_s2_fill28 -DHJ_TUNE_P2_OWN_SHIFT=4 -DHJ_TUNE_P2_STEPS=2 -DHJ_TUNE_P2_FILL16=28 &
wait
