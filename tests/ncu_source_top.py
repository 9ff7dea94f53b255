#!/usr/bin/env python3
"""Top source lines by warp-stall samples from `ncu --page source --csv --print-source sass,cuda` exports (gz or plain):
   python tests/ncu_source_top.py profiles/x.source.csv.gz [kernel substring] [n]"""
import csv, gzip, sys, collections
path = sys.argv[1]; want = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 12
f = gzip.open(path, "rt") if path.endswith(".gz") else open(path)
fn = fpath = None; hdr = None; seen = set()
agg = collections.defaultdict(lambda: [0, 0, ""])  # (function, file, line) -> [samples, instructions, text]
for r in csv.reader(f):
    if not r: continue
    if r[0] == "File Path": fpath = r[1]; hdr = None; continue
    if r[0] == "Function Name": fn = r[1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or not want in (fn or ""): continue
    if not r[0].strip().isdigit(): continue   # SASS rows have no line number: the CUDA-line rows carry the per-line sums
    try:
        s = int(r[hdr.index("# Samples")] or 0); n = int(r[hdr.index("Instructions Executed")] or 0)
    except ValueError:
        continue
    k = (fn.split("(")[0][-60:], fpath.split("/")[-1], r[0])
    agg[k][0] += s; agg[k][1] += n; agg[k][2] = r[1][:110]
byfn = collections.defaultdict(list)
for k, v in agg.items(): byfn[k[0]].append((v[0], v[1], k[1], k[2], v[2]))
for fnname, rows in byfn.items():
    tot = sum(x[0] for x in rows) or 1; toti = sum(x[1] for x in rows) or 1
    print(f"== {fnname}: {tot} samples, {toti} warp instructions")
    for s, n, fl, ln, txt in sorted(rows, reverse=True)[:top]:
        print(f"  {100*s/tot:5.1f}% stalls {100*n/toti:5.1f}% instr  {fl}:{ln}  {txt.strip()}")
