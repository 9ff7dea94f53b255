#!/bin/bash
# Quick iteration loop on the GPU box: a fast parity subset, then per-phase timings of C2 (and optionally C3).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/iter.log
: > $L
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "${1:-random_small or forced_radix or int64 or medium_1m or skew or overflow or edge_sizes}" > gpurun_out/iter_pytest.log 2>&1
echo "pytest rc=$? $(tail -1 gpurun_out/iter_pytest.log)" | tee -a $L
grep -E "^E  |Error" gpurun_out/iter_pytest.log | head -10 | tee -a $L
timeout 600 python tests/profile_join.py --scale 1.0 --joins 4 2>&1 | tail -3 | tee -a $L
if [ -n "$2" ]; then timeout 600 python tests/profile_join.py --workload c3 --scale $2 --joins 3 2>&1 | tail -2 | tee -a $L; fi
