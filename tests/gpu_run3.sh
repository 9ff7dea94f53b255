#!/bin/bash
# Round-1 state capture on the GPU box: smoke, full GPU suite, both bench arms, ncu launch list + one full capture.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/run3.log
: > $L
nvidia-smi -L >> $L; nproc >> $L; free -g | head -2 >> $L
ls /usr/include/boost /usr/include/tbb >> $L 2>&1
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $L
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $L
tail -5 gpurun_out/pytest_gpu.log | tee -a $L
( time timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2>> $L; echo "bench rc=$?" | tee -a $L
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err ) 2>> $L; echo "bench ref rc=$?" | tee -a $L
# ncu launch list of the bench command (times under ncu are never bench values)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 200 --csv --log-file gpurun_out/ncu_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launches.out 2>&1; echo "ncu launches rc=$?" | tee -a $L
# full capture of one join at full scale (second join of profile_join.py; 14 launches per join)
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_ -s 14 -c 14 -o gpurun_out/ncu_full_c2 -f \
  python tests/profile_join.py --scale 1.0 > gpurun_out/ncu_full.out 2>&1; echo "ncu full rc=$?" | tee -a $L
cat gpurun_out/bench_default.json | head -c 3000
tail -3 gpurun_out/bench_default.err
cat gpurun_out/bench_reference.json | head -c 1500
