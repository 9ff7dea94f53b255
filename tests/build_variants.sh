#!/bin/bash
# Builds variant libraries of libhj for tests/sweep_variants.py: every argument is name=FLAGS, e.g.
#   bash tests/build_variants.sh p1_old="-DHJ_TUNE_P1_PIPE=0" p1p_v2="-DHJ_TUNE_P1_VIRTUAL32=2"
# -> hyrise-1_b200/csrc/variants/libhj_<name>.so (+ <name>.ptxas with the register / spill report). Four builds at a time.
# The HJ_TUNE_* macros are documented where they are defined (hj_api.cu, hj_fused.cuh, hj_kernels.cuh).
cd "$(dirname "$0")/.."
OUT=hyrise-1_b200/csrc/variants; mkdir -p $OUT
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -shared -I include"
n=0
for spec in "$@"; do
  name=${spec%%=*}; flags=${spec#*=}
  ( /usr/local/cuda/bin/nvcc $FLAGS $flags -Xptxas -v hyrise-1_b200/csrc/hj_api.cu -o $OUT/libhj_$name.so -ldl -lpthread -lrt > $OUT/$name.ptxas 2>&1 \
      && echo "built $name ($flags)" || echo "FAILED $name (see $OUT/$name.ptxas)" ) &
  n=$((n + 1)); [ $((n % 4)) -eq 0 ] && wait
done
wait
