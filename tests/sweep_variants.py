#!/usr/bin/env python3
"""Times builds of libhj with different HJ_TUNE_* constants (tests/build_variants.sh) side by side on one GPU.

One process, one copy of the workload in HBM; every variant library is dlopen'ed in turn, joins the same relations
`--joins` times and reports the library's per-phase CUDA-event timings (best join) plus a result check: row count and
the sum of the emitted probe-side positions against the value torch computes from the keys. Development tool, not a
bench: the numbers it prints go into profiles/ only as tuning evidence.

    python tests/sweep_variants.py --scale 1.0 [--runs default probe_v2 default:13]
"""
import argparse
import ctypes as C
import glob
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

import bench

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c2")
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--joins", type=int, default=4)
ap.add_argument("--runs", nargs="*", default=None,
                help="variant[:radix_bits] in the order to run, e.g. default p2_512x8_mb3 default:13 (default: every variant)")
ap.add_argument("--out", default=None)
args = ap.parse_args()

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
dev = torch.device("cuda", 0)
data = bench.gen_device_workload(args.workload, args.scale, dev, seed=0xB202)
b = bench.relation_from_chunks(hj, data["build"], abi.LOC_DEVICE)
p = bench.relation_from_chunks(hj, data["probe"], abi.LOC_DEVICE)
torch.cuda.synchronize()


def flat_keys(chunks):
    return torch.cat([(a if at is None else a[at.long() & (0xFFFFFFFF if at.dtype == torch.int32 else 0xFFFF if at.dtype == torch.int16 else 0xFF)])
                      for _, a, at in chunks]).to(torch.int64)


# expected row count and probe-position checksum (every probe row appears once per matching build row)
bk, pk = flat_keys(data["build"]), flat_keys(data["probe"])
cb = torch.bincount(bk, minlength=int(max(bk.max(), pk.max())) + 1)
mpp = cb[pk]
want_rows = int(mpp.sum())
want_sum = int((torch.arange(len(pk), device=dev) * mpp).sum())
del bk, pk, cb, mpp
torch.cuda.empty_cache()

libs = {"default": hj.LIB_PATH}
for path in sorted(glob.glob(os.path.join(ROOT, "hyrise-1_b200", "csrc", "variants", "libhj_*.so"))):
    libs[os.path.basename(path)[len("libhj_"):-3]] = path
runs = [(r.split(":")[0], int(r.split(":")[1]) if ":" in r else -1) for r in (args.runs or list(libs))]

results = []
for name, bits in runs:
    path = libs.get(name)
    if True:
        rec = {"variant": name, "radix_bits_opt": bits}
        try:
            if path is None:
                raise FileNotFoundError(f"no library for variant {name}")
            hj._lib, hj.LIB_PATH = None, path
            j = hj.HashJoin(device=0, flags=abi.FLAG_TIMINGS, radix_bits=bits)
            best = None
            for _ in range(args.joins):
                info = j.join(b, p, "Inner")
                t = j.timings()
                if best is None or t["total"] < best["total"]:
                    best = t
            rec.update(bits=int(info.radix_bits), **{k: round(v, 3) for k, v in best.items() if isinstance(v, float)})
            _, pp, n = j.result_device()
            rec.update(rows=int(n))
            ok = n == want_rows
            if ok and n:
                class _W:  # the device PosList as a torch tensor (cuda array interface; int32 view: ids < 2^31)
                    __cuda_array_interface__ = {"shape": (n, 2), "typestr": "<i4", "data": (pp, False), "version": 2}
                rid = torch.as_tensor(_W(), device=dev)
                got = 0
                for lo in range(0, n, 1 << 28):
                    part = rid[lo:lo + (1 << 28)]
                    got += int((part[:, 0].to(torch.int64) * bench.CHUNK + part[:, 1].to(torch.int64)).sum())
                ok = got == want_sum
                del rid, part
            rec.update(ok=bool(ok))
            j.close()
            torch.cuda.empty_cache()
        except Exception as e:  # a variant that fails must not stop the sweep
            rec.update(ok=False, error=repr(e)[:300])
        results.append(rec)
        print(json.dumps(rec), flush=True)
if args.out:
    with open(args.out, "w") as f:
        json.dump(results, f, indent=1)
