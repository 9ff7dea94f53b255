#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for sr in 2000000000 500000000 250000000; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $NG --steps 4 --warmup 3 --no-cpu --no-e2e --slice-rows $sr > gpurun_out/bench_s$sr.json 2> gpurun_out/bench_s$sr.err; echo "slice-rows $sr rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_s$sr.json").read().strip().splitlines()[-1])
    r=d.get("roofline") or {}
    print("value=%.4g"%d["value"], "ms=%.2f"%d["ms_per_step"], {k:round(v,2) for k,v in r.get("phase_ms",{}).items()})
except Exception as e:
    print("no bench line:", e)
PY
done
