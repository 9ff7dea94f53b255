#!/bin/bash
# Runs every GPU test function in its own process (a CUDA fault is sticky for a process) and writes logs + a summary
# under gpurun_out/. Usage on the GPU box:  bash tests/run_gpu_suite.sh [pytest -k filter]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/suite
SUMMARY=gpurun_out/suite/summary.txt
: > "$SUMMARY"
nvidia-smi -L >> "$SUMMARY" 2>&1
nproc >> "$SUMMARY"
TESTS=$(python -m pytest tests/test_gpu_parity.py -m gpu --collect-only -q ${1:+-k "$1"} 2>/dev/null | grep '::' | sed 's/\[.*//' | sort -u)
fail=0
for t in $TESTS; do
  name=$(echo "$t" | sed 's/.*:://')
  timeout 900 python -m pytest "$t" -m gpu -q -x --no-header -p no:cacheprovider > "gpurun_out/suite/$name.log" 2>&1
  rc=$?
  line=$(tail -1 "gpurun_out/suite/$name.log")
  echo "$name rc=$rc $line" | tee -a "$SUMMARY"
  if [ $rc -ne 0 ]; then fail=1; grep -E "Error|error|assert|differ" "gpurun_out/suite/$name.log" | head -8 | tee -a "$SUMMARY"; fi
done
echo "suite_failed=$fail" | tee -a "$SUMMARY"
exit $fail
