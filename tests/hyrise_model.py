"""Test-side model of the slice of Hyrise the JoinHash tests touch (TEST INFRASTRUCTURE).

Mirrors, in small, the reference's storage layer (Table/Chunk/ValueColumn/DictionaryColumn/ReferenceColumn,
src/lib/storage), `load_table` (src/lib/utils/load_table.cpp:14-48), the dictionary encoder
(storage/dictionary_column/dictionary_encoder.hpp:58-130 + fixed_size_byte_aligned_compressor.cpp:21-30),
TableScan's output conventions (operators/table_scan.cpp:78-160) and the host half of JoinHash
(operators/join_hash.cpp:49-87 side selection, :533-613 + :829-855 output assembly), so that the reference's own
test cases (tests/golden/join_fixtures.txt) can be replayed against any *join engine*:
    engine(build: abi.Relation, probe: abi.Relation, mode: int) -> (build_side (n,2) uint32 | None, probe_side (n,2) uint32)
The engine is either the CPU oracle (tests/oracle_binding.py) or libhj on the GPU (hyrise-1_b200.HashJoin).
"""
import os as _os
import sys as _sys

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
if _ROOT not in _sys.path:
    _sys.path.insert(0, _ROOT)
import importlib
import math
import os

import numpy as np

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NP_TYPES = {"int": np.int32, "long": np.int64, "float": np.float32, "double": np.float64, "string": object}
NULL_ROW = (abi.INVALID_CHUNK_ID, abi.INVALID_CHUNK_OFFSET)


# ---------------------------------------------------------------------------------------------- storage
class ValueColumn:
    def __init__(self, values, nulls=None):
        self.values = values  # np array (object dtype for strings)
        self.nulls = nulls    # np.uint8 array or None (non-nullable)

    def __len__(self):
        return len(self.values)

    def get(self, i):
        if self.nulls is not None and self.nulls[i]:
            return None
        return self.values[i]


class DictionaryColumn:
    def __init__(self, dictionary, attr, null_value_id):
        self.dictionary, self.attr, self.null_value_id = dictionary, attr, null_value_id

    def __len__(self):
        return len(self.attr)

    def get(self, i):
        v = int(self.attr[i])
        return None if v == self.null_value_id else self.dictionary[v]


class ReferenceColumn:
    def __init__(self, table, column_id, pos_list):
        assert table.type == "Data"
        self.table, self.column_id, self.pos_list = table, column_id, pos_list  # pos_list: (n,2) uint32, shared

    def __len__(self):
        return len(self.pos_list)

    def get(self, i):
        c, o = int(self.pos_list[i, 0]), int(self.pos_list[i, 1])
        if o == abi.INVALID_CHUNK_OFFSET:
            return None
        return self.table.chunks[c][self.column_id].get(o)


class Table:
    def __init__(self, names, types, nullable, ttype="Data"):
        self.names, self.types, self.nullable, self.type = list(names), list(types), list(nullable), ttype
        self.chunks = []  # list of list of columns

    def row_count(self):
        return sum(len(c[0]) for c in self.chunks)

    def rows(self):
        out = []
        for chunk in self.chunks:
            for i in range(len(chunk[0])):
                out.append(tuple(col.get(i) for col in chunk))
        return out


def encode_dictionary(col, np_type):
    """dictionary_encoder.hpp:58-130: sorted unique non-NULL values; NULL -> null_value_id == dict size; narrowest of
    u8/u16/u32 that holds dict.size()+1 (fixed_size_byte_aligned_compressor.cpp:21-30)."""
    n = len(col)
    valid = np.ones(n, bool) if col.nulls is None else (col.nulls == 0)
    if np_type is object:
        dictionary = np.array(sorted(set(col.values[valid])), dtype=object)
        lookup = {v: i for i, v in enumerate(dictionary)}
        attr = np.array([lookup[v] if ok else len(dictionary) for v, ok in zip(col.values, valid)], dtype=np.uint32)
    else:
        dictionary = np.unique(col.values[valid]).astype(np_type)
        attr = np.searchsorted(dictionary, col.values).astype(np.uint32)
        attr[~valid] = len(dictionary)
    max_value = len(dictionary) + 1
    width = np.uint8 if max_value <= 0xFF else (np.uint16 if max_value <= 0xFFFF else np.uint32)
    return DictionaryColumn(dictionary, attr.astype(width), len(dictionary))


# ---------------------------------------------------------------------------------------------- fixtures
class Fixtures:
    def __init__(self, path=os.path.join(GOLDEN, "join_fixtures.txt")):
        self.tables, self.cases = {}, []
        with open(path) as f:
            lines = f.read().split("\n")
        i = 0
        while i < len(lines):
            line = lines[i]
            if line.startswith("@table "):
                name, body = line.split()[1], []
                i += 1
                while lines[i] != "@end":
                    body.append(lines[i])
                    i += 1
                self.tables[name] = body
            elif line.startswith("@case "):
                _, name, cite = line.split()
                ops = []
                i += 1
                while lines[i] != "@end":
                    ops.append(lines[i].split())
                    i += 1
                self.cases.append((name, cite, ops))
            i += 1

    def load_table(self, name, chunk_size):
        """utils/load_table.cpp:14-48."""
        body = self.tables[name]
        names = body[0].split("|")
        types, nullable = [], []
        for t in body[1].split("|"):
            parts = t.split("_")
            types.append(parts[0])
            nullable.append(len(parts) > 1 and parts[1] == "null")
        table = Table(names, types, nullable)
        rows = [l.split("|") for l in body[2:] if l != ""]
        for start in range(0, len(rows), chunk_size):
            part = rows[start:start + chunk_size]
            chunk = []
            for c, t in enumerate(types):
                np_type = NP_TYPES[t]
                raw = [r[c] for r in part]
                nulls = np.array([1 if (nullable[c] and v == "null") else 0 for v in raw], dtype=np.uint8)
                if np_type is object:
                    vals = np.array([("" if nl else v) for v, nl in zip(raw, nulls)], dtype=object)
                else:
                    conv = float if t in ("float", "double") else int
                    vals = np.array([(0 if nl else conv(v)) for v, nl in zip(raw, nulls)], dtype=np_type)
                chunk.append(ValueColumn(vals, nulls if nullable[c] else None))
            table.chunks.append(chunk)
        return table


def encode_chunks(table, which):
    ids = range(len(table.chunks)) if which == "all" else [int(x) for x in which.split(",")]
    for cid in ids:
        if cid < len(table.chunks):
            table.chunks[cid] = [encode_dictionary(col, NP_TYPES[t]) for col, t in zip(table.chunks[cid], table.types)]


# ---------------------------------------------------------------------------------------------- operators
_OPS = {">=": lambda a, b: a >= b, ">": lambda a, b: a > b, "=": lambda a, b: a == b, "<=": lambda a, b: a <= b,
        "<": lambda a, b: a < b}


def table_scan(table, column_id, op, value):
    """operators/table_scan.cpp:78-160: one output chunk per input chunk with matches; output ReferenceColumns point at
    the Data table (pos lists of reference inputs are resolved); NULL never matches."""
    out = Table(table.names, table.types, table.nullable, "References")
    pred = _OPS[op]
    for cid, chunk in enumerate(table.chunks):
        col = chunk[column_id]
        matches = [i for i in range(len(col)) if col.get(i) is not None and pred(col.get(i), value)]
        if not matches:
            continue
        if table.type == "References":
            cache = {}
            cols = []
            for c in chunk:
                key = id(c.pos_list)
                if key not in cache:
                    cache[key] = np.ascontiguousarray(c.pos_list[matches])
                cols.append(ReferenceColumn(c.table, c.column_id, cache[key]))
        else:
            pl = np.array([(cid, i) for i in matches], dtype=np.uint32).reshape(-1, 2)
            cols = [ReferenceColumn(table, c, pl) for c in range(len(chunk))]
        out.chunks.append(cols)
    return out


def flatten_key_column(table, column_id, location=abi.LOC_HOST):
    """What the shim does before calling the library: one hj_segment per chunk of the key column."""
    key_np = NP_TYPES[table.types[column_id]]
    if key_np is object:
        raise NotImplementedError(f"key type {table.types[column_id]} is outside the scope so far (SURVEY §8f N3)")
    rb = hj.RelationBuilder(location)
    for cid, chunk in enumerate(table.chunks):
        col = chunk[column_id]
        if isinstance(col, ValueColumn):
            rb.add_value(cid, np.ascontiguousarray(col.values), col.nulls)
        elif isinstance(col, DictionaryColumn):
            rb.add_dict(cid, np.ascontiguousarray(col.dictionary), col.attr, col.null_value_id)
        else:  # ReferenceColumn: gather the keys through the PosList (reference_column_iterable.hpp:59-90)
            n = len(col)
            vals = np.zeros(n, dtype=key_np)
            nulls = np.zeros(n, dtype=np.uint8)
            for i in range(n):
                v = col.get(i)
                if v is None:
                    nulls[i] = 1  # value stays T{}
                else:
                    vals[i] = v
            rb.add_value(cid, vals, nulls if nulls.any() else None, kind=abi.SEG_GATHERED)
    return rb.build()


def _output_columns(in_table, pos_list):
    """write_output_columns, join_hash.cpp:564-613."""
    cols = []
    if in_table.type == "References":
        if in_table.chunks:
            cache = {}
            for c in range(len(in_table.names)):
                in_lists = tuple(id(ch[c].pos_list) for ch in in_table.chunks)
                if in_lists not in cache:
                    new = np.empty_like(pos_list)
                    for i in range(len(pos_list)):
                        cid, off = int(pos_list[i, 0]), int(pos_list[i, 1])
                        if off == abi.INVALID_CHUNK_OFFSET:
                            new[i] = NULL_ROW
                        else:
                            new[i] = in_table.chunks[cid][c].pos_list[off]
                    cache[in_lists] = new
                ref = in_table.chunks[0][c]
                cols.append(ReferenceColumn(ref.table, ref.column_id, cache[in_lists]))
        else:
            dummy = Table(in_table.names, in_table.types, in_table.nullable)  # Table::create_dummy_table
            cols = [ReferenceColumn(dummy, c, pos_list) for c in range(len(in_table.names))]
    else:
        cols = [ReferenceColumn(in_table, c, pos_list) for c in range(len(in_table.names))]
    return cols


def join_hash(left, right, mode, column_ids, engine):
    """JoinHash::_on_execute (join_hash.cpp:49-87) + output assembly (:690-708, :829-855) around `engine`."""
    mode_id = abi.MODES[mode] if isinstance(mode, str) else mode
    swapped = mode_id in (abi.MODE_LEFT, abi.MODE_ANTI, abi.MODE_SEMI)
    if not swapped and left.row_count() > right.row_count():
        swapped = True
    if swapped:
        build_t, probe_t, build_c, probe_c = right, left, column_ids[1], column_ids[0]
    else:
        build_t, probe_t, build_c, probe_c = left, right, column_ids[0], column_ids[1]
    build_rel = flatten_key_column(build_t, build_c)
    probe_rel = flatten_key_column(probe_t, probe_c)
    build_side, probe_side = engine(build_rel, probe_rel, mode_id)
    semi_anti = mode_id in (abi.MODE_SEMI, abi.MODE_ANTI)
    if semi_anti:
        names, types, nullable = probe_t.names, probe_t.types, probe_t.nullable
    elif swapped:
        names, types = probe_t.names + build_t.names, probe_t.types + build_t.types
        nullable = probe_t.nullable + build_t.nullable
    else:
        names, types = build_t.names + probe_t.names, build_t.types + probe_t.types
        nullable = build_t.nullable + probe_t.nullable
    out = Table(names, types, nullable, "References")
    if len(probe_side) == 0:
        return out
    probe_cols = _output_columns(probe_t, probe_side)
    if semi_anti:
        out.chunks.append(probe_cols)
    else:
        build_cols = _output_columns(build_t, build_side)
        out.chunks.append(probe_cols + build_cols if swapped else build_cols + probe_cols)
    return out


def nested_loop_join(left, right, mode, column_ids):
    """Semantic stand-in for the reference's JoinNestedLoop in MixNestedLoopAndHash (join_equi_test.cpp:360)."""
    assert mode in ("Inner", "Left")
    lrows, rrows = [], []
    lpos = [(c, i) for c, ch in enumerate(left.chunks) for i in range(len(ch[0]))]
    rpos = [(c, i) for c, ch in enumerate(right.chunks) for i in range(len(ch[0]))]
    for lc, li in lpos:
        lv = left.chunks[lc][column_ids[0]].get(li)
        hit = False
        for rc, ri in rpos:
            rv = right.chunks[rc][column_ids[1]].get(ri)
            if lv is not None and rv is not None and lv == rv:
                lrows.append((lc, li))
                rrows.append((rc, ri))
                hit = True
        if not hit and mode == "Left":
            lrows.append((lc, li))
            rrows.append(NULL_ROW)
    out = Table(left.names + right.names, left.types + right.types, left.nullable + right.nullable, "References")
    if lrows:
        lp = np.array(lrows, dtype=np.uint32).reshape(-1, 2)
        rp = np.array(rrows, dtype=np.uint32).reshape(-1, 2)
        out.chunks.append(_output_columns(left, lp) + _output_columns(right, rp))
    return out


# ---------------------------------------------------------------------------------------------- comparison
def _norm(v):
    if v is None:
        return (1, 0)
    if isinstance(v, (float, np.floating)):
        return (0, round(float(v), 3))
    if isinstance(v, (int, np.integer)):
        return (0, int(v))
    return (0, str(v))


def tables_equal_unordered(actual, expected, eps=1e-4):
    """EXPECT_TABLE_EQ_UNORDERED, src/test/testing_assert.cpp:118-254: same column count/types, same multiset of rows,
    floats compared with an absolute epsilon."""
    if len(actual.types) != len(expected.types):
        return False, f"column count {len(actual.types)} != {len(expected.types)}"
    if actual.types != expected.types:
        return False, f"column types {actual.types} != {expected.types}"
    a = sorted(actual.rows(), key=lambda r: tuple(_norm(v) for v in r))
    e = sorted(expected.rows(), key=lambda r: tuple(_norm(v) for v in r))
    if len(a) != len(e):
        return False, f"row count {len(a)} != {len(e)}\nactual={a}\nexpected={e}"
    for ra, re_ in zip(a, e):
        for va, ve in zip(ra, re_):
            if (va is None) != (ve is None):
                return False, f"row {ra} != {re_}"
            if va is None:
                continue
            if isinstance(ve, (float, np.floating)):
                if not math.isclose(float(va), float(ve), abs_tol=eps, rel_tol=1e-6):
                    return False, f"row {ra} != {re_}"
            elif va != ve:
                return False, f"row {ra} != {re_}"
    return True, ""


def case_in_scope(ops, numeric_floats=False):
    """Integer-key cases; with numeric_floats also the int x float / int x double cases (GPU library only: the oracle
    restates the integer path)."""
    scopes = [op[1] for op in ops if op[0] == "scope"]
    return not scopes or (numeric_floats and all(s in ("float", "double") for s in scopes))


def run_case(fx, ops, engine):
    """Replays one @case; returns (ok, message)."""
    env = {}
    for op in ops:
        kind = op[0]
        if kind == "scope":
            continue
        if kind == "base":
            t = fx.load_table(op[2], int(op[3]))
            if len(op) > 4 and op[4] == "dict":
                encode_chunks(t, op[5])
            env[op[1]] = t
        elif kind == "scan":
            env[op[1]] = table_scan(env[op[2]], int(op[3]), op[4], float(op[5]) if "." in op[5] else int(op[5]))
        elif kind == "join":
            env[op[1]] = join_hash(env[op[2]], env[op[3]], op[4], (int(op[5]), int(op[6])), engine)
        elif kind == "nlj":
            env[op[1]] = nested_loop_join(env[op[2]], env[op[3]], op[4], (int(op[5]), int(op[6])))
        elif kind == "expect":
            expected = fx.load_table(op[2], 1)
            return tables_equal_unordered(env[op[1]], expected)
        else:
            raise ValueError(op)
    raise ValueError("case without expect")
