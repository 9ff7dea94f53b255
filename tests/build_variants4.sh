#!/bin/bash
cd "$(dirname "$0")/.."
OUT=hyrise-1_b200/csrc/variants; mkdir -p $OUT; rm -f $OUT/*.so $OUT/*.ptxas
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -shared -I include"
P2="-DHJ_TUNE_P2_OWN_SHIFT=3 -DHJ_TUNE_P2_STEPS=1"
build() { name=$1; shift
  /usr/local/cuda/bin/nvcc $FLAGS "$@" -Xptxas -v hyrise-1_b200/csrc/hj_api.cu -o $OUT/libhj_$name.so -ldl -lpthread -lrt > $OUT/$name.ptxas 2>&1 \
    && echo "built $name" || echo "FAILED $name (see $OUT/$name.ptxas)"; }
build p2best $P2 &
build dictpf $P2 -DHJ_TUNE_P1_DICT_PREFETCH=1 &
build ilp $P2 -DHJ_TUNE_FIND_ILP=1 &
build dictpf_ilp $P2 -DHJ_TUNE_P1_DICT_PREFETCH=1 -DHJ_TUNE_FIND_ILP=1 &
wait
