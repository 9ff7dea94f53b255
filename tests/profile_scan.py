#!/usr/bin/env python3
"""Times hj_scan on the probe column of a bench workload that is resident in HBM (NOT yet run on a GPU: written after the
round's GPU minutes were spent). For every selectivity: wall time around the call (hj_scan ends with a
host read-back of the match count, so the call is synchronous), rows/s, and the algorithmic HBM bytes of DESIGN.md §4:
w_in bytes per row read + 2 x 1 bit per row (mask written by S1, read by S3) + 8 bytes per match written.

    python tests/profile_scan.py --workload c2 --scale 1.0
    ncu --set full -k regex:k_scan python tests/profile_scan.py --scale 0.1 --repeats 1

The literal is a fraction of the largest key, so `key < literal` matches about that share of the (uniform) rows."""
import argparse
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

import bench

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c2")
ap.add_argument("--scale", type=float, default=0.1)
ap.add_argument("--repeats", type=int, default=5)
ap.add_argument("--selectivities", default="0.0,0.01,0.1,0.5,1.0")
ap.add_argument("--peak-gbs", type=float, default=None, help="HBM peak to report against (default: MEASURED_PEAKS.json)")
args = ap.parse_args()

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
dev = torch.device("cuda", 0)
data = bench.gen_device_workload(args.workload, args.scale, dev, seed=0xB202)
probe = bench.relation_from_chunks(hj, data["probe"], abi.LOC_DEVICE)
n_rows = int(probe.n_rows)
# chunks are (chunk_id, dictionary | values, attribute vector | None); bytes read per row: the attribute vector of a
# dictionary chunk, the values of a value chunk
_, first_keys, first_attr = data["probe"][0]
w_in = (first_attr if first_attr is not None else first_keys).element_size()
key_max = max(int(keys.max()) for _, keys, _ in data["probe"] if len(keys))
peak = args.peak_gbs
if peak is None:
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            mp = json.load(f)
        peak = float(mp.get("hbm_gbs") or 0) or None
    except (OSError, ValueError):
        peak = None
torch.cuda.synchronize()
j = hj.HashJoin(device=0)
out = []
for sel in [float(x) for x in args.selectivities.split(",")]:
    literal = int(key_max * sel) + (1 if sel >= 1.0 else 0)
    times, info = [], None
    for _ in range(args.repeats + 1):  # the first call allocates the result buffers and is not counted
        t0 = time.perf_counter()
        info = j.scan(probe, abi.PRED_LESS_THAN, literal)
        times.append(time.perf_counter() - t0)
    best = min(times[1:])
    model_bytes = n_rows * w_in + 2 * n_rows / 8 + info.n_rows * 8
    rec = {"selectivity_asked": sel, "literal": literal, "rows": n_rows, "matches": int(info.n_rows), "ms": best * 1e3,
           "rows_per_s": n_rows / best, "model_gb": model_bytes / 1e9, "gbs": model_bytes / best / 1e9,
           "frac_of_peak": (model_bytes / best / 1e9 / peak) if peak else None}
    out.append(rec)
    print(json.dumps(rec), flush=True)
j.close()
