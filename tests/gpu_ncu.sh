#!/bin/bash
# ncu --set full of selected kernels of one C2 join. usage: gpu_ncu.sh <kernel regex> <scale> <skip> <count> <tag>
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:$1" -s ${3:-0} -c ${4:-4} -o gpurun_out/ncu_${5:-iter} -f \
  python tests/profile_join.py --scale ${2:-1.0} --joins 2 > gpurun_out/ncu_${5:-iter}.out 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_${5:-iter}.out
