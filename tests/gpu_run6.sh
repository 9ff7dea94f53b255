#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded.py -m gpu -x -q 2>&1 | tail -5
NG=$(nvidia-smi -L | wc -l)
for n in 2 4 8; do
  if [ "$NG" -ge $n ]; then
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $n --steps 5 --warmup 3 --no-cpu ${BENCH_EXTRA} > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench n$n rc=$?"
    tail -2 gpurun_out/bench_n$n.err
  fi
done
python - <<'PY'
import json
for f in ("bench_n2","bench_n4","bench_n8"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    except Exception as e:
        continue
    r=d.get("roofline") or {}
    print(f, "value=%.3g"%d["value"], "ms=%.2f"%d["ms_per_step"], "e2e=", (d.get("e2e") or {}).get("value"),
          "phases=", {k: round(v,2) for k,v in (r.get("phase_ms") or {}).items()}, "nvlink=", r.get("nvlink"))
PY
