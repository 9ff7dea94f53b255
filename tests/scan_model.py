"""Test-side model of the tables and operator plumbing around a single-column TableScan (tests only).

Replays the cases of tests/golden/scan_fixtures.json (the reference's src/test/operators/table_scan_test.cpp, restated by
tests/golden/make_scan_golden.py) through ANY scan function with the signature of oracle.table_scan_oracle.table_scan —
the CPU oracle itself, or the GPU library behind the C ABI (gpu_scan below) — and evaluates the expectations the way the
reference's ASSERT_COLUMN_EQ does: the values of another column of the returned rows, as a multiset.

What the model does around the scan is what TableScan::_on_execute does around scan_chunk (table_scan.cpp:84-171): one
output chunk per input chunk that has matches; for a data table the matches ARE the output PosList, for a reference
table every distinct input PosList is filtered by the match offsets (:112-131)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

INVALID = 0xFFFFFFFF
PREDICATES = {"Equals": 0, "NotEquals": 1, "LessThan": 2, "LessThanEquals": 3, "GreaterThan": 4, "GreaterThanEquals": 5,
              "Between": 6, "In": 7, "Like": 8, "NotLike": 9, "IsNull": 10, "IsNotNull": 11}


TYPE_NAMES = {"int": np.int32, "long": np.int64, "float": np.float32, "double": np.float64, "string": str}  # load_table.cpp column types


def load_fixtures():
    with open(os.path.join(ROOT, "tests", "golden", "scan_fixtures.json"), encoding="utf-8") as f:
        return json.load(f)


def dict_encode(values, nulls):
    """dictionary_encoder.hpp:58-130 + fixed_size_byte_aligned_compressor.cpp:21-30 (same as workloads.dict_encode)."""
    valid = nulls == 0
    dictionary = np.unique(values[valid]).astype(values.dtype)
    attr = np.searchsorted(dictionary, values).astype(np.uint32)
    attr[~valid] = len(dictionary)
    m = len(dictionary) + 1
    width = np.uint8 if m <= 0xFF else (np.uint16 if m <= 0xFFFF else np.uint32)
    return {"kind": "dict", "dictionary": dictionary, "attr": attr.astype(width), "null_value_id": len(dictionary)}


def for_encode(values, nulls):
    """frame_of_reference_encoder.hpp:30-90: 2048-row blocks, NULL stored as 0, offset = value - block minimum."""
    v = values.copy()
    v[nulls != 0] = 0
    n = len(v)
    minima = np.array([v[b:b + 2048].min() for b in range(0, n, 2048)], dtype=values.dtype)
    off = v.astype(np.int64) - np.repeat(minima.astype(np.int64), 2048)[:n]
    top = int(off.max(initial=0))
    width = np.uint8 if top < 256 else np.uint16 if top < 65536 else np.uint32
    return {"kind": "for", "minima": minima, "offsets": off.astype(width), "nulls": nulls.copy() if nulls.any() else None}


def rle_encode(values, nulls):
    """run_length_encoder.hpp:20-60: a run ends where the value or the NULL-ness changes; inclusive end positions."""
    n = len(values)
    change = np.ones(n, dtype=bool)
    change[1:] = (nulls[1:] != nulls[:-1]) | ((nulls[1:] == 0) & (values[1:] != values[:-1]))
    starts = np.nonzero(change)[0]
    ends = np.append(starts[1:] - 1, n - 1).astype(np.uint32) if n else np.zeros(0, np.uint32)
    return {"kind": "rle", "n_rows": n, "run_values": np.ascontiguousarray(values[starts]), "end_positions": ends,
            "run_nulls": np.ascontiguousarray(nulls[starts])}


ENCODERS = {"dict": dict_encode, "for": for_encode, "rle": rle_encode}


class DataTable:
    """load_table(name, chunk_size) + ChunkEncoder::encode_chunks on the listed chunks with EncodingType `style`."""

    def __init__(self, columns, chunk_size, encode, dtype=np.int32, style="dict", dtypes=None):
        self.style = style
        dtypes = dtypes or {}
        self.names = list(columns)
        n = len(next(iter(columns.values())))
        self.bounds = [(s, min(s + chunk_size, n)) for s in range(0, n, chunk_size)] or [(0, 0)]
        self.values, self.nulls = {}, {}
        for name, col in columns.items():
            self.values[name] = np.array([0 if v is None else v for v in col], dtype=dtypes.get(name, dtype))
            self.nulls[name] = np.array([1 if v is None else 0 for v in col], dtype=np.uint8)
        all_chunks = list(range(len(self.bounds)))
        self.encoded = set(all_chunks if encode == "all" else [] if encode == "none" else encode)

    def chunk(self, name, c):
        lo, hi = self.bounds[c]
        v, nl = np.ascontiguousarray(self.values[name][lo:hi]), np.ascontiguousarray(self.nulls[name][lo:hi])
        if c in self.encoded:
            return ENCODERS[self.style](v, nl)
        return {"kind": "value", "values": v, "nulls": nl if nl.any() else None}

    def column_chunks(self, name):
        return [self.chunk(name, c) for c in range(len(self.bounds))]

    def value_at(self, name, chunk_id, offset):
        i = self.bounds[chunk_id][0] + offset
        return None if self.nulls[name][i] else self.values[name][i].item()


def pos_array(rows):
    return np.array([[INVALID, INVALID] if r is None else r for r in rows], dtype=np.uint32).reshape(-1, 2)


def run_case(case, tables, scan_fn, encode_override=None, style="dict"):
    """Runs one fixture case through scan_fn; returns ("rows", n) / ("values", sorted list) to compare with the expectation."""
    if "generated" in case:  # get_table_op_with_n_dict_entries: values 0..n in one chunk
        n = case["generated"]["n"]
        table = DataTable({"a": list(range(n + 1))}, n + 1, case["encode"] if encode_override is None else encode_override,
                          style=style)
    else:
        t = tables[case["table"]]
        dtypes = {n: TYPE_NAMES[ty.replace("_null", "")] for n, ty in zip(t["columns"], t["types"])} if "types" in t else None
        table = DataTable(t["data"], case["chunk_size"], case["encode"] if encode_override is None else encode_override,
                          style=style, dtypes=dtypes)
    n_chunks = len(table.bounds)
    excluded = set(case.get("excluded_chunks", []))
    # current relation: list of chunks; a chunk is None (data chunk c of `table`) or {column: pos_list} (reference chunk)
    ref = case.get("reference")
    if ref is None:
        rel = [("data", c) for c in range(n_chunks) if c not in excluded]
    elif "pos_list" in ref:
        pl = pos_array(ref["pos_list"])
        rel = [("ref", {name: pl for name in table.names})]
    elif ref.get("identity"):
        pl = pos_array([[c, o] for c, (lo, hi) in enumerate(table.bounds) for o in range(hi - lo)])
        rel = [("ref", {name: pl for name in table.names})]
    else:
        rel = [("ref", {name: pos_array(rows) for name, rows in ref["pos_lists"].items()})]

    for step in case["steps"]:
        col, pred = step["column"], PREDICATES[step["predicate"]]
        value = step.get("value")
        chunks, ids = [], []
        for i, (kind, payload) in enumerate(rel):
            ids.append(payload if kind == "data" else i)
            if kind == "data":
                chunks.append(table.chunk(col, payload))
            else:
                chunks.append({"kind": "reference", "pos_list": payload[col], "referenced": table.column_chunks(col)})
        pos, begin = scan_fn(chunks, ids, pred, value)
        out = []
        for i, (kind, payload) in enumerate(rel):
            m = pos[begin[i]:begin[i + 1]]
            if len(m) == 0:
                continue  # table_scan.cpp:97: chunks without matches are dropped
            assert (m[:, 0] == ids[i]).all(), "a match carries the wrong chunk id"
            assert (np.diff(m[:, 1].astype(np.int64)) > 0).all(), "matches are not in ascending chunk offset"
            if kind == "data":
                shared = m.copy()
                out.append(("ref", {name: shared for name in table.names}))
            else:
                filtered = {}
                for name, pl in payload.items():  # one filtered list per DISTINCT input PosList (:112-131)
                    filtered[name] = filtered.get(id(pl), pl[m[:, 1]])
                    filtered[id(pl)] = filtered[name]
                out.append(("ref", {name: filtered[name] for name in table.names}))
        rel = out

    n_rows = sum(len(next(iter(p.values()))) if k == "ref" else table.bounds[p][1] - table.bounds[p][0] for k, p in rel)
    if "expect_rows" in case:
        return "rows", n_rows
    name = case["expect_column"]
    got = []
    for kind, payload in rel:
        for c, o in payload[name]:
            got.append(None if (c == INVALID and o == INVALID) else table.value_at(name, int(c), int(o)))
    return "values", sorted(got, key=lambda v: (v is None, v))


def expected_of(case):
    if "expect_rows" in case:
        return "rows", case["expect_rows"]
    return "values", sorted(case["expect"], key=lambda v: (v is None, v))


def encodings_for(case):
    """The reference runs every test for Unencoded / Dictionary / RunLength / FrameOfReference (:211-214) with the case's
    own chunk selection; here: the case's selection as dictionary chunks, plus all-value and all-dictionary."""
    own = case["encode"]
    outs = [own]
    for e in ("none", "all"):
        if e != own:
            outs.append(e)
    return outs
