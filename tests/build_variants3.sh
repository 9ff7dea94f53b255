#!/bin/bash
cd "$(dirname "$0")/.."
OUT=hyrise-1_b200/csrc/variants; mkdir -p $OUT; rm -f $OUT/*.so $OUT/*.ptxas
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -shared -I include"
build() { name=$1; shift
  /usr/local/cuda/bin/nvcc $FLAGS "$@" -Xptxas -v hyrise-1_b200/csrc/hj_api.cu -o $OUT/libhj_$name.so -ldl -lpthread -lrt > $OUT/$name.ptxas 2>&1 \
    && echo "built $name" || echo "FAILED $name (see $OUT/$name.ptxas)"; }
build own3_s1 -DHJ_TUNE_P2_OWN_SHIFT=3 -DHJ_TUNE_P2_STEPS=1 &
build own2 -DHJ_TUNE_P2_OWN_SHIFT=2 &
build own2_s1 -DHJ_TUNE_P2_OWN_SHIFT=2 -DHJ_TUNE_P2_STEPS=1 &
build own4_s2 -DHJ_TUNE_P2_OWN_SHIFT=4 -DHJ_TUNE_P2_STEPS=2 &
wait
build mcs12 -DHJ_TUNE_MAX_CHUNK_SHIFT=12 &
build mcs16 -DHJ_TUNE_MAX_CHUNK_SHIFT=16 &
build jt512_v8 -DHJ_TUNE_JOIN_THREADS=512 -DHJ_TUNE_PROBE_V=8 &
build jt512_v4 -DHJ_TUNE_JOIN_THREADS=512 &
wait
build jt768_v4 -DHJ_TUNE_JOIN_THREADS=768 &
build outcs -DHJ_TUNE_OUT_CS=1 &
wait
