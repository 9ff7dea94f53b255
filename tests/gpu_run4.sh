#!/bin/bash
# bench both arms at N=1 + 2-GPU bench (if 2 GPUs visible)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
L=gpurun_out/run4.log; : > $L
NG=$(nvidia-smi -L | wc -l)
( time timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err ) 2>> $L; echo "bench n1 rc=$?" | tee -a $L
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err ) 2>> $L; echo "ref rc=$?" | tee -a $L
if [ "$NG" -ge 2 ]; then
  ( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err ) 2>> $L; echo "bench n2 rc=$?" | tee -a $L
  tail -3 gpurun_out/bench_n2.err
fi
tail -3 gpurun_out/bench_n1.err
python - <<'PY'
import json
for f in ("bench_n1","bench_n2","bench_ref"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    r=d.get("roofline") or {}
    print(f, "value=%.3g"%d["value"], "ms=%.2f"%d["ms_per_step"], "e2e=", (d.get("e2e") or {}).get("value"), "cpu=", (d.get("cpu_baseline") or {}).get("value"),
          "kernel=", r.get("kernel"), "frac=", r.get("frac"), "whole=", (r.get("whole_join") or {}).get("frac"), "phases=", {k: round(v,2) for k,v in (r.get("phase_ms") or {}).items()}, "nvlink=", r.get("nvlink"))
PY
