#!/bin/bash
# Round 2, first GPU call (one B200): the verification debt (tests/gpu_round2_first.sh), then the prepared variant sweep and
# the gather microbenchmark. Every step has its own timeout; logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/r2_box.txt 2>&1
nproc >> gpurun_out/r2_box.txt; free -g >> gpurun_out/r2_box.txt
bash tests/gpu_round2_first.sh
( time timeout 90 tests/micro/gather_bench > gpurun_out/r2_gather_bench.txt 2>&1 ) 2>> gpurun_out/round2_first.log
echo "gather_bench rc=$?" | tee -a gpurun_out/round2_first.log; cat gpurun_out/r2_gather_bench.txt
( time timeout 400 python tests/sweep_variants.py --scale 1.0 --runs default gather_cg gather_noalloc p2_mb3 p2_mb2 p2_smem p2_smem_mb3 p1_v8_t1024 bits_6_8 bits_8_6 default \
    --out gpurun_out/r2_sweep_variants5.json > gpurun_out/r2_sweep_variants5.out 2>&1 ) 2>> gpurun_out/round2_first.log
echo "sweep rc=$?" | tee -a gpurun_out/round2_first.log
grep '^{' gpurun_out/r2_sweep_variants5.out | cut -c1-400
