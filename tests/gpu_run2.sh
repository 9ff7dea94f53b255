#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee gpurun_out/run2.log
timeout 2400 bash tests/run_gpu_suite.sh > gpurun_out/suite_stdout.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/run2.log
for sc in 0.1 1.0; do
  timeout 900 python bench.py --scale $sc --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_scale_$sc.json 2> gpurun_out/bench_scale_$sc.err
  echo "bench scale $sc rc=$?" | tee -a gpurun_out/run2.log
done
tail -3 gpurun_out/smoke.log
tail -25 gpurun_out/suite/summary.txt
tail -3 gpurun_out/bench_scale_1.0.err
