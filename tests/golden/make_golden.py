#!/usr/bin/env python3
"""Generates the committed golden fixtures from the reference checkout (run in the build container only).

    python tests/golden/make_golden.py            # needs /root/reference; the GPU box never runs this

Outputs (committed, read by the tests at run time — nothing under /root/reference is read by any test):
  tests/golden/join_fixtures.txt  every .tbl table the reference's JoinHash tests load, plus one `@case` block per
                                  reference test that exercises JoinHash (src/test/operators/join_equi_test.cpp,
                                  join_null_test.cpp, join_semi_anti_test.cpp), written in a small line DSL that both
                                  the Python harness (tests/hyrise_model.py) and the C++ shim test runner parse.
  tests/golden/murmur_kat.json    known answers of the reference's own src/lib/utils/murmur_hash.cpp (compiled into
                                  oracle/_ref/libref_murmur.so by oracle/Makefile) for int32/int64 keys, seed 17.

Case DSL (one op per line, names are local to the case):
  base NAME TABLE CHUNK_SIZE [dict all | dict 0,1,..]   load_table(TABLE, CHUNK_SIZE) (+ ChunkEncoder on those chunks)
  scan NAME INPUT COLUMN OP VALUE                        TableScan(INPUT, ColumnID, PredicateCondition, value)
  join NAME LEFT RIGHT MODE LCOL RCOL                    the join under test (JoinHash)
  nlj  NAME LEFT RIGHT MODE LCOL RCOL                    JoinNestedLoop of the reference test (semantic join in the harness)
  expect NAME TABLE                                      EXPECT_TABLE_EQ_UNORDERED(NAME.get_output(), load_table(TABLE))
  scope KIND                                             key types outside v1 scope (float/double/string)
"""
import ctypes
import json
import os
import random
import re
import struct
import subprocess
import sys

REF = "/root/reference"
TABLES = os.path.join(REF, "src/test/tables")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

# join_test.hpp:28-94 table wrappers: name -> (tbl path relative to src/test/tables, chunk size)
W = {
    "a": ("int_float", 2), "b": ("int_float2", 2), "c": ("int_string", 4), "d": ("string_int", 3),
    "e": ("int_int", 4), "f": ("int_int2", 4), "g": ("int_int3", 4), "h": ("int_int4", 4),
    "i": ("int5", 1), "j": ("int3", 1), "k": ("int4", 1), "l": ("int", 1),
    "m": ("aggregateoperator/groupby_int_1gb_0agg/input_null", 20),
    "n": ("aggregateoperator/groupby_int_1gb_1agg/input_null", 20),
    "o": ("float_zero_precision", 1), "p": ("double_zero_precision", 1), "q": ("string_numbers", 1),
    "a_null": ("int_float_with_null", 2),  # join_null_test.cpp:36
    "semi_a": ("joinoperators/semi_left", 2), "semi_b": ("joinoperators/semi_right", 2),  # join_semi_anti_test.cpp:28-31
}
DICT = {  # join_test.hpp:48-66, join_null_test.cpp:41-45
    "a_dict": ("a", "all"), "b_dict": ("b", "all"), "c_dict": ("a", "0"), "m_dict": ("m", "0"), "n_dict": ("n", "0"),
    "a_null_dict": ("a_null", "all"),
}

J = "joinoperators/"


def base(name):
    if name in DICT:
        src, which = DICT[name]
        t, cs = W[src]
        return f"base {name} {t} {cs} dict {which}"
    t, cs = W[name]
    return f"base {name} {t} {cs}"


def simple(test, line, left, right, mode, lcol, rcol, expect, scope=None):
    ops = [base(left)] + ([base(right)] if right != left else [])
    ops.append(f"join j {left} {right} {mode} {lcol} {rcol}")
    ops.append(f"expect j {expect}")
    if scope:
        ops.insert(0, f"scope {scope}")
    return (test, line, ops)


def scanned(test, line, specs, join_ops, expect):
    """specs: list of (scan name, base wrapper, column, op, value)."""
    ops, seen = [], set()
    for _, wrapper, *_ in specs:
        if wrapper not in seen:
            ops.append(base(wrapper))
            seen.add(wrapper)
    for sname, wrapper, col, op, val in specs:
        ops.append(f"scan {sname} {wrapper} {col} {op} {val}")
    ops += join_ops
    ops.append(f"expect {join_ops[-1].split()[1]} {expect}")
    return (test, line, ops)


E, N, S = "join_equi_test.cpp", "join_null_test.cpp", "join_semi_anti_test.cpp"
CASES = [
    simple("JoinEquiTest.LeftJoin", f"{E}:45", "a", "b", "Left", 0, 0, J + "int_left_join"),
    simple("JoinEquiTest.InnerJoinIntFloat.1", f"{E}:51", "a", "o", "Inner", 0, 0, J + "int_float_inner", "float"),
    simple("JoinEquiTest.InnerJoinIntFloat.2", f"{E}:51", "o", "a", "Inner", 0, 0, J + "float_int_inner", "float"),
    simple("JoinEquiTest.InnerJoinIntDouble.1", f"{E}:80", "a", "p", "Inner", 0, 0, J + "int_double_inner", "double"),
    simple("JoinEquiTest.InnerJoinIntDouble.2", f"{E}:80", "p", "a", "Inner", 0, 0, J + "double_int_inner", "double"),
    simple("JoinEquiTest.InnerJoinIntString", f"{E}:96", "a", "q", "Inner", 0, 0, J + "int_string_inner", "string"),
    simple("JoinEquiTest.LeftJoinOnString", f"{E}:106", "c", "d", "Left", 1, 0, J + "string_left_join", "string"),
    simple("JoinEquiTest.RightJoin", f"{E}:112", "a", "b", "Right", 0, 0, J + "int_right_join"),
    simple("JoinEquiTest.InnerJoin", f"{E}:128", "a", "b", "Inner", 0, 0, J + "int_inner_join"),
    simple("JoinEquiTest.InnerJoinOnString", f"{E}:134", "c", "d", "Inner", 1, 0, J + "string_inner_join", "string"),
    scanned("JoinEquiTest.InnerRefJoin", f"{E}:140", [("sa", "a", 0, ">=", 0), ("sb", "b", 0, ">=", 0)],
            ["join j sa sb Inner 0 0"], J + "int_inner_join"),
    simple("JoinEquiTest.InnerValueDictJoin", f"{E}:154", "a", "b_dict", "Inner", 0, 0, J + "int_inner_join"),
    simple("JoinEquiTest.InnerDictValueJoin", f"{E}:160", "a_dict", "b", "Inner", 0, 0, J + "int_inner_join"),
    scanned("JoinEquiTest.InnerValueDictRefJoin", f"{E}:166", [("sa", "a", 0, ">=", 0), ("sb", "b_dict", 0, ">=", 0)],
            ["join j sa sb Inner 0 0"], J + "int_inner_join"),
    scanned("JoinEquiTest.InnerDictValueRefJoin", f"{E}:180", [("sa", "a_dict", 0, ">=", 0), ("sb", "b", 0, ">=", 0)],
            ["join j sa sb Inner 0 0"], J + "int_inner_join"),
    scanned("JoinEquiTest.InnerRefJoinFiltered", f"{E}:194", [("sa", "a", 0, ">", 1000), ("sb", "b", 0, ">=", 0)],
            ["join j sa sb Inner 0 0"], J + "int_inner_join_filtered"),
    simple("JoinEquiTest.InnerDictJoin", f"{E}:206", "a_dict", "b_dict", "Inner", 0, 0, J + "int_inner_join"),
    scanned("JoinEquiTest.InnerRefDictJoin", f"{E}:212", [("sa", "a_dict", 0, ">=", 0), ("sb", "b_dict", 0, ">=", 0)],
            ["join j sa sb Inner 0 0"], J + "int_inner_join"),
    scanned("JoinEquiTest.InnerRefDictJoinFiltered", f"{E}:226",
            [("sa", "a_dict", 0, ">", 1000), ("sb", "b_dict", 0, ">=", 0)],
            ["join j sa sb Inner 0 0"], J + "int_inner_join_filtered"),
    simple("JoinEquiTest.InnerJoinBig", f"{E}:239", "c", "d", "Inner", 0, 1, J + "int_string_inner_join"),
    scanned("JoinEquiTest.InnerRefJoinFilteredBig", f"{E}:245", [("sc", "c", 0, ">=", 0), ("sd", "d", 1, ">=", 6)],
            ["join j sc sd Inner 0 1"], J + "int_string_inner_join_filtered"),
    simple("JoinEquiTest.JoinOnMixedValueAndDictionaryColumns", f"{E}:258", "c_dict", "b", "Inner", 0, 0,
           J + "int_inner_join"),
    scanned("JoinEquiTest.JoinOnMixedValueAndReferenceColumns", f"{E}:264", [("sa", "a", 0, ">=", 0)],
            [base("b"), "join j sa b Inner 0 0"], J + "int_inner_join"),
    scanned("JoinEquiTest.MultiJoinOnReferenceLeft", f"{E}:275",
            [("sa", "f", 0, ">=", 0), ("sb", "g", 0, ">=", 0), ("sc", "h", 0, ">=", 0)],
            ["join j1 sa sb Inner 0 0", "join j2 j1 sc Inner 0 0"], J + "int_inner_multijoin_ref_ref_ref_left"),
    scanned("JoinEquiTest.MultiJoinOnReferenceRight", f"{E}:297",
            [("sa", "f", 0, ">=", 0), ("sb", "g", 0, ">=", 0), ("sc", "h", 0, ">=", 0)],
            ["join j1 sa sb Inner 0 0", "join j2 sc j1 Inner 0 0"], J + "int_inner_multijoin_ref_ref_ref_right"),
    scanned("JoinEquiTest.MultiJoinOnReferenceLeftFiltered", f"{E}:319",
            [("sa", "f", 0, ">", 6), ("sb", "g", 0, ">=", 0), ("sc", "h", 0, ">=", 0)],
            ["join j1 sa sb Inner 0 0", "join j2 j1 sc Inner 0 0"],
            J + "int_inner_multijoin_ref_ref_ref_left_filtered"),
    ("JoinEquiTest.MultiJoinOnValue", f"{E}:340",
     [base("f"), base("g"), base("h"), "join j1 f g Inner 0 0", "join j2 j1 h Inner 0 0",
      "expect j2 " + J + "int_inner_multijoin_val_val_val_left"]),
    ("JoinEquiTest.MultiJoinOnRefOuter", f"{E}:350",
     [base("f"), base("g"), base("h"), "join j1 f g Left 0 0", "join j2 j1 h Inner 0 0",
      "expect j2 " + J + "int_inner_multijoin_val_val_val_leftouter"]),
    ("JoinEquiTest.MixNestedLoopAndHash", f"{E}:360",
     [base("f"), base("g"), base("h"), "nlj j1 f g Left 0 0", "join j2 j1 h Inner 0 0",
      "expect j2 " + J + "int_inner_multijoin_nlj_hash"]),
    ("JoinEquiTest.MixHashAndNestedLoop", f"{E}:370",
     [base("f"), base("g"), base("h"), "join j1 f g Left 0 0", "join j2 j1 h Inner 0 0",
      "expect j2 " + J + "int_inner_multijoin_nlj_hash"]),
    scanned("JoinEquiTest.RightJoinRefColumn", f"{E}:380", [("sa", "a", 0, ">=", 0)],
            [base("b"), "join j sa b Right 0 0"], J + "int_right_join"),
    scanned("JoinEquiTest.LeftJoinRefColumn", f"{E}:391", [("sb", "b", 0, ">=", 0)],
            [base("a"), "join j a sb Left 0 0"], J + "int_left_join"),
    scanned("JoinEquiTest.RightJoinEmptyRefColumn", f"{E}:402", [("sa", "a", 0, "=", 0)],
            [base("b"), "join j sa b Right 0 0"], J + "int_join_empty"),
    scanned("JoinEquiTest.LeftJoinEmptyRefColumn", f"{E}:413", [("sb", "b", 0, "=", 0)],
            ["join j b sb Left 0 0"], J + "int_join_empty_left"),
    # ---- join_null_test.cpp
    simple("JoinNullTest.InnerJoinWithNull", f"{N}:55", "a", "a_null", "Inner", 0, 0, J + "int_float_null_inner"),
    simple("JoinNullTest.InnerJoinWithNullDict", f"{N}:61", "a_dict", "a_null_dict", "Inner", 0, 0,
           J + "int_float_null_inner"),
    simple("JoinNullTest.InnerJoinWithNull2", f"{N}:67", "m", "n", "Inner", 0, 0, J + "int_inner_join_null"),
    simple("JoinNullTest.InnerJoinWithNullDict2", f"{N}:73", "m_dict", "n_dict", "Inner", 0, 0,
           J + "int_inner_join_null"),
    scanned("JoinNullTest.InnerJoinWithNullRef2", f"{N}:79", [("sa", "m", 1, ">=", 0), ("sb", "n", 1, ">=", 0)],
            ["join j sa sb Inner 0 0"], J + "int_inner_join_null_ref"),
    simple("JoinNullTest.LeftJoinWithNullAsOuter", f"{N}:92", "a_null", "b", "Left", 0, 0, J + "int_left_join_null"),
    simple("JoinNullTest.LeftJoinWithNullAsOuterDict", f"{N}:98", "a_null_dict", "b_dict", "Left", 0, 0,
           J + "int_left_join_null"),
    simple("JoinNullTest.LeftJoinWithNullAsInner", f"{N}:104", "b", "a_null", "Left", 0, 0,
           J + "int_left_join_null_inner"),
    simple("JoinNullTest.LeftJoinWithNullAsInnerDict", f"{N}:110", "b_dict", "a_null_dict", "Left", 0, 0,
           J + "int_left_join_null_inner"),
    simple("JoinNullTest.RightJoinWithNullAsOuter", f"{N}:116", "b", "a_null", "Right", 0, 0,
           J + "int_right_join_null"),
    simple("JoinNullTest.RightJoinWithNullAsOuterDict", f"{N}:122", "b_dict", "a_null_dict", "Right", 0, 0,
           J + "int_right_join_null"),
    simple("JoinNullTest.RightJoinWithNullAsInner", f"{N}:128", "a_null", "b", "Right", 0, 0,
           J + "int_right_join_null_inner"),
    simple("JoinNullTest.RightJoinWithNullAsInnerDict", f"{N}:134", "a_null_dict", "b_dict", "Right", 0, 0,
           J + "int_right_join_null_inner"),
    # ---- join_semi_anti_test.cpp
    simple("JoinSemiAntiTest.SemiJoin", f"{S}:42", "k", "a", "Semi", 0, 0, "int"),
    scanned("JoinSemiAntiTest.SemiJoinRefColumns", f"{S}:47", [("sa", "k", 0, ">=", 0), ("sb", "a", 0, ">=", 0)],
            ["join j sa sb Semi 0 0"], "int"),
    simple("JoinSemiAntiTest.SemiJoinBig", f"{S}:58", "semi_a", "semi_b", "Semi", 0, 0, J + "semi_result"),
    simple("JoinSemiAntiTest.AntiJoin", f"{S}:64", "k", "a", "Anti", 0, 0, J + "anti_int4"),
    scanned("JoinSemiAntiTest.AntiJoinRefColumns", f"{S}:69", [("sa", "k", 0, ">=", 0), ("sb", "a", 0, ">=", 0)],
            ["join j sa sb Anti 0 0"], J + "anti_int4"),
    simple("JoinSemiAntiTest.AntiJoinBig", f"{S}:80", "semi_a", "semi_b", "Anti", 0, 0, J + "anti_result"),
]


def find_line(cite, test):
    """Real line of the TEST in the reference file (the literals above are only hints)."""
    fname = cite.split(":")[0]
    tname = test.split(".")[1]
    pat = re.compile(r"TYPED_TEST\(\w+, %s\)|TEST_F\(\w+, %s\)" % (tname, tname))
    with open(os.path.join(REF, "src/test/operators", fname)) as f:
        for i, l in enumerate(f, 1):
            if pat.search(l):
                return f"{fname}:{i}"
    raise KeyError(test)


def collect_tables():
    names = set()
    for _, _, ops in CASES:
        for op in ops:
            parts = op.split()
            if parts[0] == "base":
                names.add(parts[2])
            elif parts[0] == "expect":
                names.add(parts[2])
    return sorted(names)


def write_fixtures(path):
    out = ["# GENERATED by tests/golden/make_golden.py from the reference's test tables (src/test/tables/**.tbl) and",
           "# JoinHash test cases (src/test/operators/join_{equi,null,semi_anti}_test.cpp). Do not edit by hand.",
           "# .tbl format: utils/load_table.cpp:14-48 — line 1 column names, line 2 types (suffix _null = nullable), rows."]
    for name in collect_tables():
        with open(os.path.join(TABLES, name + ".tbl")) as f:
            body = f.read().rstrip("\n")
        out.append(f"@table {name}")
        out.append(body)
        out.append("@end")
    for test, line, ops in CASES:
        out.append(f"@case {test} src/test/operators/{find_line(line, test)}")
        out += ops
        out.append("@end")
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")
    print(f"wrote {path}: {len(collect_tables())} tables, {len(CASES)} cases")


def write_murmur_kat(path):
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/libref_murmur.so"])
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle/_ref/libref_murmur.so"))
    fn = lib.ref_murmur_hash2
    fn.restype = ctypes.c_uint32
    fn.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32]
    rng = random.Random(0xB200)
    k32 = [0, 1, -1, 2, 17, 2**31 - 1, -2**31, 12345, 6000000] + [rng.randrange(-2**31, 2**31) for _ in range(1200)]
    k64 = [0, 1, -1, 2**31, 2**32, 2**63 - 1, -2**63, 4000000000] + [rng.randrange(-2**63, 2**63) for _ in range(1200)]
    kat = {"seed": 17, "source": "src/lib/utils/murmur_hash.cpp compiled verbatim (oracle/Makefile -> oracle/_ref)",
           "int32": [], "int64": []}
    for k in k32:
        buf = ctypes.create_string_buffer(struct.pack("<i", k), 4)
        kat["int32"].append([k, fn(buf, 4, 17)])
    for k in k64:
        buf = ctypes.create_string_buffer(struct.pack("<q", k), 8)
        kat["int64"].append([k, fn(buf, 8, 17)])
    with open(path, "w") as f:
        json.dump(kat, f)
    print(f"wrote {path}: {len(k32)} int32 + {len(k64)} int64 known answers")


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs the reference checkout at /root/reference")
    write_fixtures(os.path.join(HERE, "join_fixtures.txt"))
    write_murmur_kat(os.path.join(HERE, "murmur_kat.json"))
