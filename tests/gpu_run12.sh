#!/bin/bash
# Tuning sweep on one B200 (C2 at full size, every variant library of tests/build_variants.sh), C3 at full size with the
# default library, then the parity file that exercises the join kernels hardest.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/run12.log; : > $L
( time timeout 300 python tests/sweep_variants.py --scale 1.0 --joins 4 \
    --runs default old_find_steps p1_prefetch p1_v4_t2048 p1_v4_t2048_pf p1_v2_t2048 p2_512x8_mb2 p2_512x8_mb3 p2_256x8_mb6 \
           p2_512x8_mb3_pf probe_v2 probe_v8 default:13 default:15 default \
    --out gpurun_out/sweep_main.json > gpurun_out/sweep_main.out 2>&1 ) 2>> $L; echo "sweep main rc=$?" | tee -a $L
grep '^{' gpurun_out/sweep_main.out | tee -a $L; grep -v '^{' gpurun_out/sweep_main.out | tail -5 | tee -a $L
( time timeout 150 python tests/sweep_variants.py --workload c3 --scale 1.0 --joins 3 --runs default old_find_steps \
    --out gpurun_out/sweep_c3_full.json > gpurun_out/sweep_c3_full.out 2>&1 ) 2>> $L; echo "c3 full rc=$?" | tee -a $L
grep '^{' gpurun_out/sweep_c3_full.out | tee -a $L; grep -v '^{' gpurun_out/sweep_c3_full.out | tail -5 | tee -a $L
( time timeout 240 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --durations=8 > gpurun_out/pytest_parity.log 2>&1 ) 2>> $L; echo "pytest parity rc=$?" | tee -a $L
tail -15 gpurun_out/pytest_parity.log | tee -a $L
