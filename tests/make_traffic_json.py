#!/usr/bin/env python3
"""profiles/traffic_c2_full.json from an `ncu --set full` raw export (ncu -i x.ncu-rep --page raw --csv) of ONE C2 join:
dram__bytes_read.sum + dram__bytes_write.sum summed per kernel over the join's launches (build side + probe side = one
"launch set", the unit bench.py's per-phase algorithmic bytes use), stamped with the hash of the kernel sources it was
captured from — bench.py reports `roofline.traffic` only while that hash still matches (a stale capture reads null).

    python tests/make_traffic_json.py gpurun_out/x_ncu1.raw.csv [--out profiles/traffic_c2_full.json]"""
import argparse
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

ap = argparse.ArgumentParser()
ap.add_argument("raw_csv")
ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "traffic_c2_full.json"))
args = ap.parse_args()
rows = list(csv.reader(open(args.raw_csv)))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {n: i for i, n in enumerate(hdr)}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
out = {"source": f"{os.path.basename(args.raw_csv)}: ncu --set full of one C2 join (100M x 1B); dram__bytes_read.sum + dram__bytes_write.sum "
                 "summed over the kernel's launches in the join",
       "csrc_sha256": bench.csrc_hash(), "launches": {}}
for r in data:
    name = re.match(r"(?:void )?(?:hj::)?(\w+)", r[col["Kernel Name"]]).group(1)
    b = 0.0
    for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        b += float(r[col[m]].replace(",", "")) * scale[units[col[m]]]
    out[name] = out.get(name, 0) + int(b)
    out["launches"][name] = out["launches"].get(name, 0) + 1
with open(args.out, "w") as f:
    json.dump(out, f, indent=1)
print(json.dumps(out))
