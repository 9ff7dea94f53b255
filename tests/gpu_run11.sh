#!/bin/bash
# Re-entry check of HEAD on one B200 + tuning sweep: (1) variant sweep at C2 full size, (2) ncu --set full with source
# of the three hot kernels of one C2 join, (3) quick device-only bench line, (4) the whole `-m gpu` suite.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/run11.log; : > $L
nvidia-smi -L | tee -a $L; nproc | tee -a $L
( time timeout 300 python tests/sweep_variants.py --scale 1.0 --joins 4 \
    --runs default p1_prefetch p2_512x8_mb2 p2_512x8_mb3 p2_256x8_mb6 probe_v2 probe_v8 default:13 default:15 p1_v4_t2048 \
    --out gpurun_out/sweep_main.json > gpurun_out/sweep_main.out 2>&1 ) 2>> $L; echo "sweep main rc=$?" | tee -a $L
cat gpurun_out/sweep_main.out | grep '^{' | tee -a $L
( time timeout 300 ncu --set full --clock-control none --import-source on -k regex:'k_part1|k_scatter2f|k_build_probe' -s 5 -c 5 \
    -o gpurun_out/ncu_full_c2_r11 -f python tests/profile_join.py --scale 1.0 > gpurun_out/ncu_full.out 2>&1 ) 2>> $L
echo "ncu full rc=$?" | tee -a $L; tail -3 gpurun_out/ncu_full.out | tee -a $L
( time timeout 120 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err ) 2>> $L
echo "bench quick rc=$?" | tee -a $L; cat gpurun_out/bench_quick.json | tee -a $L
( time timeout 150 python tests/sweep_variants.py --workload c3 --scale 1.0 --joins 3 --runs default \
    --out gpurun_out/sweep_c3_full.json > gpurun_out/sweep_c3_full.out 2>&1 ) 2>> $L; echo "c3 full rc=$?" | tee -a $L
grep '^{' gpurun_out/sweep_c3_full.out | tee -a $L; tail -2 gpurun_out/sweep_c3_full.out >> $L
( time timeout 420 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1 ) 2>> $L; echo "pytest rc=$?" | tee -a $L
tail -5 gpurun_out/pytest_gpu.log | tee -a $L
