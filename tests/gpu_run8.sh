#!/bin/bash
# streamed-join parity + bench (value, e2e) on one GPU
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bound_output or random_small or two_pass" > gpurun_out/run8_pytest.log 2>&1
echo "pytest rc=$? $(tail -1 gpurun_out/run8_pytest.log)"
grep -E "^E  |Error" gpurun_out/run8_pytest.log | head -10
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/run8_bench.json 2> gpurun_out/run8_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/run8_bench.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/run8_bench.json').read().strip().splitlines()[-1])
print('value', l['value']/1e9, 'ms', l['ms_per_step'], 'e2e', l['e2e'], 'roofline', {k:l['roofline'][k] for k in ('kernel','achieved','frac','kernel_ms')}, 'whole', l['roofline']['whole_join'])
print('cpu', l['cpu_baseline'])
PY
