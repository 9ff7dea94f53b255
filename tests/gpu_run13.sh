#!/bin/bash
# Evidence run for the round's final state on one B200: default bench (CPU baseline + e2e), reference arm, ncu launch
# list of the bench command, ncu --set full of the hot kernels of one C2 join, C3 at full size, then parity subsets.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/run13.log; : > $L
( time timeout 300 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2>> $L; echo "bench rc=$?" | tee -a $L
cat gpurun_out/bench_default.json | tee -a $L
( time timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err ) 2>> $L; echo "ref rc=$?" | tee -a $L
( time timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 200 --csv --log-file gpurun_out/ncu_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launches.out 2>&1 ) 2>> $L; echo "ncu launches rc=$?" | tee -a $L
( time timeout 240 ncu --set full --clock-control none --import-source on -k regex:'k_part1|k_scatter2f|k_build_probe' -s 5 -c 5 \
    -o gpurun_out/ncu_full_c2_r13 -f python tests/profile_join.py --scale 1.0 > gpurun_out/ncu_full.out 2>&1 ) 2>> $L
echo "ncu full rc=$?" | tee -a $L; tail -3 gpurun_out/ncu_full.out | tee -a $L
( time timeout 100 python tests/sweep_variants.py --workload c3 --scale 1.0 --joins 3 --runs default \
    --out gpurun_out/sweep_c3_full.json > gpurun_out/sweep_c3_full.out 2>&1 ) 2>> $L; echo "c3 full rc=$?" | tee -a $L
grep '^{' gpurun_out/sweep_c3_full.out | tee -a $L
( time timeout 220 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --durations=5 \
    -k "forced_radix or random_small or edge_sizes or null_probe or medium_1m or irregular or bound_output or (two_pass and int32 and 14) or attr_widths or device_resident or reference_case or all_duplicate or skewed" \
    > gpurun_out/pytest_parity_subset.log 2>&1 ) 2>> $L; echo "pytest parity subset rc=$?" | tee -a $L
tail -4 gpurun_out/pytest_parity_subset.log | tee -a $L
( time timeout 200 python -m pytest tests/test_gpu_configs.py -m gpu -q -x > gpurun_out/pytest_configs.log 2>&1 ) 2>> $L; echo "pytest configs rc=$?" | tee -a $L
tail -3 gpurun_out/pytest_configs.log | tee -a $L
