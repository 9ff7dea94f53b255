#!/bin/bash
# final evidence run (1 GPU): default bench, reference arm, ncu launch list of the bench command, ncu --set full of one join
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/run7.log; : > $L
( time timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2>> $L; echo "bench rc=$?" | tee -a $L
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err ) 2>> $L; echo "ref rc=$?" | tee -a $L
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 200 --csv --log-file gpurun_out/ncu_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launches.out 2>&1; echo "ncu launches rc=$?" | tee -a $L
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_ -s 12 -c 12 -o gpurun_out/ncu_full_c2 -f \
  python tests/profile_join.py --scale 1.0 > gpurun_out/ncu_full.out 2>&1; echo "ncu full rc=$?" | tee -a $L
tail -2 gpurun_out/ncu_full.out
cat $L
