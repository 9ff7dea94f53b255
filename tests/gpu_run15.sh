#!/bin/bash
# Final state of the round on one B200: device-only bench line, ncu launch list of that command, ncu --set full of the hot
# kernels of one C2 join, the parity subset that exercises the changed kernels, C3 at full size.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/run15.log; : > $L
( time timeout 60 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_final_quick.json 2> gpurun_out/bench_final_quick.err ) 2>> $L; echo "bench rc=$?" | tee -a $L
cat gpurun_out/bench_final_quick.json | tee -a $L
( time timeout 90 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 200 --csv --log-file gpurun_out/ncu_launches_final.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launches.out 2>&1 ) 2>> $L; echo "ncu launches rc=$?" | tee -a $L
( time timeout 120 ncu --set full --clock-control none --import-source on -k regex:'k_part1|k_scatter2f|k_build_probe' -s 5 -c 5 \
    -o gpurun_out/ncu_full_c2_final -f python tests/profile_join.py --scale 1.0 > gpurun_out/ncu_full.out 2>&1 ) 2>> $L
echo "ncu full rc=$?" | tee -a $L; tail -3 gpurun_out/ncu_full.out | tee -a $L
( time timeout 150 python -m pytest tests/test_gpu_parity.py -m gpu -q -x \
    -k "forced_radix or random_small or edge_sizes or null_probe or medium_1m or irregular or bound_output or (two_pass and int32 and 14) or attr_widths or device_resident or reference_case or all_duplicate or skewed or int64_and_mixed or sentinel" \
    > gpurun_out/pytest_parity_final.log 2>&1 ) 2>> $L; echo "pytest parity subset rc=$?" | tee -a $L
tail -4 gpurun_out/pytest_parity_final.log | tee -a $L
( time timeout 40 python tests/sweep_variants.py --workload c3 --scale 1.0 --joins 3 --runs default \
    --out gpurun_out/sweep_c3_final.json > gpurun_out/sweep_c3_final.out 2>&1 ) 2>> $L; echo "c3 full rc=$?" | tee -a $L
grep '^{' gpurun_out/sweep_c3_final.out | tee -a $L
