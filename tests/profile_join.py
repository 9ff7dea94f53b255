#!/usr/bin/env python3
"""One warm-up join + one join of a bench workload, for ncu: `ncu -k regex:k_ -s <launches of warm-up> ... python tests/profile_join.py`.
Prints the library's per-phase timings of the second join (never a bench number when run under a profiler)."""
import argparse
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

import bench

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c2")
ap.add_argument("--scale", type=float, default=0.1)
ap.add_argument("--mode", default="Inner")
ap.add_argument("--joins", type=int, default=2)
ap.add_argument("--flags", type=lambda x: int(x, 0), default=0, help="extra HJ_FLAG_* bits")
ap.add_argument("--sharded", action="store_true",
                help="hj_join_sharded with ONE rank: the exchange kernels store into the rank's own receive regions, the local join "
                     "runs over them — the multi-GPU code path, profilable with ncu on one GPU")
args = ap.parse_args()
hj = importlib.import_module("hyrise-1_b200")
dev = torch.device("cuda", 0)
data = bench.gen_device_workload(args.workload, args.scale, dev, seed=0xB202)
b = bench.relation_from_chunks(hj, data["build"], hj.abi.LOC_DEVICE)
p = bench.relation_from_chunks(hj, data["probe"], hj.abi.LOC_DEVICE)
torch.cuda.synchronize()
j = hj.HashJoin(device=0, flags=hj.abi.FLAG_TIMINGS | args.flags)
if args.sharded:
    j.comm_init(hj.comm_unique_id(), 1, 0)
for i in range(args.joins):
    info = j.join_sharded(b, p, args.mode) if args.sharded else j.join(b, p, args.mode)
    t = j.timings()
    print(f"join {i}: rows={info.n_rows} bits={info.radix_bits} overflow_parts={info.n_overflow_partitions} "
          f"passes={info.n_probe_passes} " + " ".join(f"{k}={v:.3f}" for k, v in t.items() if isinstance(v, float)),
          f"launches={t['n_kernel_launches']}", flush=True)
j.close()
