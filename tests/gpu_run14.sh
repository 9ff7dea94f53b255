#!/bin/bash
# Pass-2 tuning sweep with 2048-row slabs (tests/build_variants2.sh) on one B200.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/run14.log; : > $L
( time timeout 200 python tests/sweep_variants.py --scale 1.0 --joins 4 \
    --runs default p2best dictpf ilp dictpf_ilp default \
    --out gpurun_out/sweep_r4.json > gpurun_out/sweep_r4.out 2>&1 ) 2>> $L; echo "sweep rc=$?" | tee -a $L
grep '^{' gpurun_out/sweep_r4.out | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['variant'], r.get('ok'), r.get('materialize'), r.get('scatter2'), r.get('build_probe'), r.get('total'), r.get('error', ''))" | tee -a $L
