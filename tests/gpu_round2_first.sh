#!/bin/bash
# First GPU call of the next round (one B200, ~15 min of box time): everything that was written after round 1's GPU minutes
# ran out, in order of how much depends on it. Each step has its own timeout and log under gpurun_out/; a failing step does
# not stop the later ones.
#   gpurun --timeout 1500 -- 'bash tests/gpu_round2_first.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/round2_first.log; : > $L
step() {  # step <name> <timeout s> <command...>
  local name=$1 t=$2; shift 2
  ( time timeout "$t" "$@" > "gpurun_out/r2_$name.out" 2>&1 ) 2>> $L
  echo "$name rc=$?" | tee -a $L
  tail -4 "gpurun_out/r2_$name.out" | tee -a $L
}
export HJ_RUN_UNVALIDATED=1
# 1. the gated device paths: scan -> join on the device, FOR/RLE scans, float / double scans, value-id (string) scans,
#    hj_gather_column
step scan_gated 150 python -m pytest tests/test_gpu_scan.py -m gpu -q -x -k "without_leaving or for_and_rle or float or string or value_id or gather or integer_column"
step aggregate_gated 600 python -m pytest tests/test_gpu_aggregate.py -m gpu -q
# 2. the C++ operators: TableScan -> hj_scan in front of JoinHash, float + string scans
step shim_gated 150 python -m pytest tests/test_shim.py -m gpu -q
# 3. C1 on the reference's own dbgen keys (SF1), through the C ABI against the oracle
step c1_dbgen 120 python -m pytest tests/test_gpu_configs.py -m gpu -q -k c1_tpch_sf1_dbgen
unset HJ_RUN_UNVALIDATED
# 4. hj_scan against its algorithmic bytes on the C2 probe column (1 B rows, 4-byte attribute vectors)
step scan_timing 90 python tests/profile_scan.py --workload c2 --scale 1.0
step pipeline_timing 90 python tests/profile_pipeline.py --workload c2 --scale 0.25 --selectivity 0.5
# 5. bench legs that have not run: C1 with dbgen keys; and the headline once more as a reference point for the session
step bench_c1 90 python bench.py --workload c1 --steps 5 --warmup 3 --no-cpu
step bench_c2 60 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu
