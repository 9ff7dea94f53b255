#!/usr/bin/env python3
"""scan -> join -> translate -> gather -> aggregate on resident columns, timed per stage (NOT yet run on a GPU).

The plan `SELECT COUNT(*), SUM(p.payload) FROM build b JOIN probe p ON b.key = p.key WHERE p.key < :literal` on a bench
workload: the probe key column is filtered by hj_scan, its PosLists go into hj_join as REFERENCE chunks without leaving
HBM, the join's probe side is translated to rows of the probe table (hj_translate_device), a payload column is read
through it (hj_gather_column) and aggregated (hj_aggregate_column). Only the aggregate comes back. Every call of the
library ends synchronised, so wall time around a call is its device time + launch overhead. Result checked with torch.

    python tests/profile_pipeline.py --workload c2 --scale 0.1 --selectivity 0.5
"""
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
ap.add_argument("--selectivity", type=float, default=0.5)
ap.add_argument("--repeats", type=int, default=3)
args = ap.parse_args()

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
dev = torch.device("cuda", 0)
data = bench.gen_device_workload(args.workload, args.scale, dev, seed=0xB202)
build = bench.relation_from_chunks(hj, data["build"], abi.LOC_DEVICE)
probe = bench.relation_from_chunks(hj, data["probe"], abi.LOC_DEVICE)
key_type = abi.KEY_INT32 if data["dtype"] == "int32" else abi.KEY_INT64


def decoded(chunk):
    _, keys, attr = chunk
    if attr is None:
        return keys
    mask = {torch.int32: 0xFFFFFFFF, torch.int16: 0xFFFF, torch.uint8: 0xFF}[attr.dtype]
    return keys[attr.long() & mask]


# a payload column with the probe table's chunking: payload = row number mod 1000 (int64 values, no NULLs)
payload_chunks, keep, row0 = hj.RelationBuilder(abi.LOC_DEVICE), [], 0
for cid, keys, attr in data["probe"]:
    n_chunk_rows = len(attr) if attr is not None else len(keys)
    t = (torch.arange(row0, row0 + n_chunk_rows, device=dev, dtype=torch.int64) % 1000).contiguous()
    keep.append(t)
    payload_chunks.add_value(cid, t)
    row0 += n_chunk_rows
payload = payload_chunks.build()
key_max = max(int(keys.max()) for _, keys, _ in data["probe"] if len(keys))
literal = int(key_max * args.selectivity)

# expected: probe rows with key < literal, weighted by the number of build rows with that key
bk = torch.cat([decoded(c) for c in data["build"]]).long()
pk = torch.cat([decoded(c) for c in data["probe"]]).long()
per_key = torch.bincount(bk, minlength=int(max(int(bk.max()), key_max)) + 1)
mult = per_key[pk] * (pk < literal)
want_rows = int(mult.sum())
want_sum = int(((torch.arange(len(pk), device=dev) % 1000) * mult).sum())
n_in = len(bk) + len(pk)
del bk, pk, per_key, mult
torch.cuda.empty_cache()


def device_view(ptr, n):
    class W:
        __cuda_array_interface__ = {"shape": (max(n, 1), 2), "typestr": "<i4", "data": (ptr, False), "version": 2}
    return torch.as_tensor(W(), device=dev)[:n]


j = hj.HashJoin(device=0)
best = None
for rep in range(args.repeats + 1):  # the first round allocates every buffer and is not counted
    torch.cuda.synchronize()
    t = [time.perf_counter()]
    j.scan(probe, abi.PRED_LESS_THAN, literal, slot=0)
    t.append(time.perf_counter())
    ptr, _, n_scan, n_chunks = j.scan_result_device(0)
    begin = j.scan_chunk_begin(0)  # host metadata of the reference table's chunks; the PosLists stay in HBM
    view = device_view(ptr, n_scan)
    rb = hj.RelationBuilder(abi.LOC_DEVICE)
    for c in range(n_chunks):
        if begin[c + 1] > begin[c]:
            rb.add_reference(c, view[int(begin[c]):int(begin[c + 1])], probe, key_type)
    filtered = rb.build()
    torch.cuda.synchronize()
    t.append(time.perf_counter())
    info = j.join(build, filtered, abi.MODE_INNER)
    t.append(time.perf_counter())
    _, p_ptr, n = j.result_device()
    j.translate_device(p_ptr, n, filtered)
    t.append(time.perf_counter())
    values_ptr, nulls_ptr, kt = j.gather_column(p_ptr, n, payload)
    t.append(time.perf_counter())
    agg = j.aggregate_column(values_ptr, nulls_ptr, n, kt)
    t.append(time.perf_counter())
    assert n == want_rows and agg["count"] == want_rows and agg["sum"] == want_sum, (n, want_rows, agg, want_sum)
    ms = {name: (b - a) * 1e3 for name, a, b in zip(("scan", "host_glue", "join", "translate", "gather", "aggregate"), t, t[1:])}
    ms["device_total"] = sum(v for k, v in ms.items() if k != "host_glue")
    if rep and (best is None or ms["device_total"] < best["device_total"]):
        best = ms
print(json.dumps({"workload": args.workload, "scale": args.scale, "literal": literal, "rows_scanned": row0, "rows_after_scan": n_scan,
                  "result_rows": want_rows, "ms": best, "input_tuples_per_s": n_in / (best["device_total"] / 1e3)}))
j.close()
