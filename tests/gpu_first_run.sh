#!/bin/bash
# First-contact script for a GPU box: environment probe, smoke, per-test parity suite, short benches.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
  echo "== env"; nvidia-smi -L; nproc; free -g | head -2
  ls -d /usr/include/boost /usr/include/tbb 2>&1 | head -2
  echo "== smoke"
  timeout 600 python __graft_entry__.py smoke
  echo "smoke rc=$?"
} > gpurun_out/first_run.log 2>&1
timeout 2400 bash tests/run_gpu_suite.sh > gpurun_out/suite_stdout.log 2>&1
echo "suite rc=$?" >> gpurun_out/first_run.log
for sc in 0.01 0.1; do
  timeout 600 python bench.py --scale $sc --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_scale_$sc.json 2> gpurun_out/bench_scale_$sc.err
  echo "bench scale $sc rc=$?" >> gpurun_out/first_run.log
done
tail -5 gpurun_out/first_run.log
tail -30 gpurun_out/suite/summary.txt
