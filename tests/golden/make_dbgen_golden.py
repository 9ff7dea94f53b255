#!/usr/bin/env python3
"""Writes tests/golden/dbgen_keys_golden.json: row counts, sums and CRC32s of the TPC-H key columns the reference's own
dbgen produces (oracle/_ref/libref_dbgen.so, built by `make -C oracle ref` where /root/reference exists) at scale
factors 0.01, 0.1 and 1. Run from the repo root in the container that has the reference checkout."""
import json
import os
import sys
import zlib

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dbgen_binding as db

out = {}
for sf in (0.01, 0.1, 1.0):
    k = db.tpch_keys(sf)
    out[str(sf)] = {name: {"n": int(len(a)), "sum": int(a.astype("int64").sum()), "crc32": zlib.crc32(a.tobytes())}
                    for name, a in k.items()}
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "dbgen_keys_golden.json"), "w") as f:
    json.dump(out, f, indent=1)
print(json.dumps({sf: {n: v["n"] for n, v in d.items()} for sf, d in out.items()}))
