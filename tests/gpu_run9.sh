#!/bin/bash
# N-GPU: sharded parity (direct, staged, sliced) + one weak-scaling bench line at N = all visible GPUs
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
if [ -z "$SKIP_TESTS" ]; then
timeout 900 python -m pytest tests/test_sharded.py -m gpu -x -q > gpurun_out/run9_pytest.log 2>&1
echo "pytest rc=$? $(tail -1 gpurun_out/run9_pytest.log)"
grep -E "MISMATCH|^E  |Error|error" gpurun_out/run9_pytest.log | head -20
fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $NG --steps 5 --warmup 3 --no-cpu ${BENCH_EXTRA} > gpurun_out/bench_n$NG.json 2> gpurun_out/bench_n$NG.err; echo "bench n$NG rc=$?"
tail -3 gpurun_out/bench_n$NG.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n$NG.json").read().strip().splitlines()[-1])
    r=d.get("roofline") or {}
    print("value=%.4g"%d["value"], "ms=%.2f"%d["ms_per_step"], "e2e=", (d.get("e2e") or {}).get("value"))
    print({k:round(v,2) for k,v in r.get("phase_ms",{}).items()})
    print(r.get("nvlink"))
except Exception as e:
    print("no bench line:", e)
PY
