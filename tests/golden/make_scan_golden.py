#!/usr/bin/env python3
"""Generates tests/golden/scan_fixtures.json from the reference checkout (run in the build container only; the tests
read the committed JSON, never /root/reference).

The tables are the .tbl files the reference's TableScan tests load (src/test/operators/table_scan_test.cpp); the cases
restate the integer-column expectations of that file: which rows a scan returns, given as the values of ANOTHER column
of the returned rows (the test's ASSERT_COLUMN_EQ, :164-193, an unordered multiset comparison), or as a row count.

Case fields: name, cite (file:line of the expectation), table, chunk_size, encode ("none" | "all" | [chunk ids] — the
test's ChunkEncoder calls; the parity tests run every case with the encodings our scan supports), steps = list of
{column, predicate (PredicateCondition name), value (int | null)} applied one after the other (a TableScan on the
previous TableScan's output), reference (optional: how the first input is wrapped into a reference table), and
expect_column + expect (multiset of values, null = NULL) or expect_rows."""
import json
import os

REF = "/root/reference"
TABLES = os.path.join(REF, "src/test/tables")
HERE = os.path.dirname(os.path.abspath(__file__))
T = "src/test/operators/table_scan_test.cpp"


def load_tbl(name):
    """src/lib/utils/load_table.cpp:14-80: header line of names, line of types (`_null` suffix = nullable), rows with `|`."""
    with open(os.path.join(TABLES, name + ".tbl")) as f:
        lines = [l.rstrip("\n") for l in f if l.strip()]
    names, types = lines[0].split("|"), lines[1].split("|")
    cols = {n: [] for n in names}
    for l in lines[2:]:
        for n, t, v in zip(names, types, l.split("|")):
            if v == "null":
                cols[n].append(None)
            elif t.startswith("int") or t.startswith("long"):
                cols[n].append(int(v))
            elif t.startswith("string"):
                cols[n].append(v)
            else:
                cols[n].append(float(v))
    return {"columns": names, "types": types, "data": cols}


ALL = [100, 102, 104, 106, 108, 110, 112] * 2
cases = []


def per_predicate(name, cite, table, chunk, encode, column, value, expect_column, tests, reference=None, first=None):
    for pred, expect in tests.items():
        steps = ([first] if first else []) + [{"column": column, "predicate": pred, "value": value}]
        c = {"name": f"{name}[{pred}]", "cite": f"{T}:{cite}", "table": table, "chunk_size": chunk, "encode": encode,
             "steps": steps, "expect_column": expect_column, "expect": expect}
        if reference:
            c["reference"] = reference
        cases.append(c)


for table, chunk, encode in (("int_int_shuffled", 7, [0, 1]), ("int_int_shuffled_2", 5, [0, 1])):  # SetUp :27-40
    # ScanOnCompressedColumn :221-254 (a <op> 6, column b of the result)
    per_predicate("ScanOnCompressedColumn", "224-232", table, chunk, encode, "a", 6, "b", {
        "Equals": [106, 106], "NotEquals": [100, 102, 104, 108, 110, 112] * 2, "LessThan": [100, 102, 104] * 2,
        "LessThanEquals": [100, 102, 104, 106] * 2, "GreaterThan": [108, 110, 112] * 2,
        "GreaterThanEquals": [106, 108, 110, 112] * 2, "IsNull": [], "IsNotNull": ALL})
    # ScanOnReferencedCompressedColumn :256-293 (b < 108, then a <op> 4 on the reference table)
    per_predicate("ScanOnReferencedCompressedColumn", "259-267", table, chunk, encode, "a", 4, "b", {
        "Equals": [104, 104], "NotEquals": [100, 102, 106] * 2, "LessThan": [100, 102] * 2,
        "LessThanEquals": [100, 102, 104] * 2, "GreaterThan": [106, 106], "GreaterThanEquals": [104, 106] * 2,
        "IsNull": [], "IsNotNull": [100, 102, 104, 106] * 2},
        first={"column": "b", "predicate": "LessThan", "value": 108})
    # ...ValueGreaterThanMaxDictionaryValue :316-340 (literal 30), ...LessThanMinDictionaryValue :342-366 (literal -10)
    per_predicate("ValueGreaterThanMaxDictionaryValue", "321-327", table, chunk, encode, "a", 30, "b", {
        "Equals": [], "NotEquals": ALL, "LessThan": ALL, "LessThanEquals": ALL, "GreaterThan": [], "GreaterThanEquals": []})
    per_predicate("ValueLessThanMinDictionaryValue", "347-353", table, chunk, encode, "a", -10, "b", {
        "Equals": [], "NotEquals": ALL, "LessThan": [], "LessThanEquals": [], "GreaterThan": ALL, "GreaterThanEquals": ALL})
    # ScanOnCompressedColumnAroundBounds :441-464 (literal 0 = the smallest dictionary entry)
    per_predicate("ScanOnCompressedColumnAroundBounds", "445-452", table, chunk, encode, "a", 0, "b", {
        "Equals": [100, 100], "LessThan": [], "LessThanEquals": [100, 100], "GreaterThan": [102, 104, 106, 108, 110, 112] * 2,
        "GreaterThanEquals": ALL, "NotEquals": [102, 104, 106, 108, 110, 112] * 2, "IsNull": [], "IsNotNull": ALL})

# ScanWeirdPosList :295-314 over get_table_op_filtered() :55-83: one reference chunk over the partly compressed table
per_predicate("ScanWeirdPosList", "297-304", "int_int_shuffled_2", 5, [0, 1], "a", 10, "b", {
    "Equals": [110, 110], "NotEquals": [100, 102, 106, 108, 112], "LessThan": [100, 102, 106, 108],
    "LessThanEquals": [100, 102, 106, 108, 110, 110], "GreaterThan": [112], "GreaterThanEquals": [110, 110, 112],
    "IsNull": [], "IsNotNull": [100, 102, 106, 108, 110, 110, 112]},
    reference={"pos_list": [[2, 0], [1, 1], [1, 3], [0, 2], [2, 2], [0, 0], [0, 4]]})

# ScanForNullValues* :503-600: IS [NOT] NULL on column b, column a of the result
NULL_TESTS = {"IsNull": [12, 123], "IsNotNull": [12345, None, 1234, 12345, 12, 1234]}
per_predicate("ScanForNullValuesOnValueColumn", "507-509", "int_int_w_null_8_rows", 4, "none", "b", None, "a", NULL_TESTS)
per_predicate("ScanForNullValuesOnCompressedColumn", "521-523", "int_int_w_null_8_rows", 4, "all", "b", None, "a", NULL_TESTS)
per_predicate("ScanForNullValuesOnReferencedValueColumn", "557-559", "int_int_w_null_8_rows", 4, "none", "b", None, "a",
              NULL_TESTS, reference={"identity": True})  # to_referencing_table :103-133
per_predicate("ScanForNullValuesOnReferencedCompressedColumn", "571-573", "int_int_w_null_8_rows", 4, "all", "b", None, "a",
              NULL_TESTS, reference={"identity": True})
per_predicate("ScanForNullValuesOnValueColumnWithoutNulls", "533-534", "int_float", 4, "none", "a", None, "a",
              {"IsNull": [], "IsNotNull": [12345, 123, 1234]})
per_predicate("ScanForNullValuesOnReferencedValueColumnWithoutNulls", "545-546", "int_float", 4, "none", "a", None, "a",
              {"IsNull": [], "IsNotNull": [12345, 123, 1234]}, reference={"identity": True})
# ...WithNullRowID... :578-600 over create_referencing_table_w_null_row_id :135-160: columns a and b have DIFFERENT PosLists
for enc, cite in (("none", "584-585"), ("all", "596-597")):
    per_predicate("ScanForNullValuesWithNullRowID" + ("OnReferencedCompressedColumn" if enc == "all" else "OnReferencedValueColumn"),
                  cite, "int_int_w_null_8_rows", 4, enc, "b", None, "a", {"IsNull": [123, 1234], "IsNotNull": [12345, None]},
                  reference={"pos_lists": {"a": [[0, 1], [1, 0], [0, 2], [0, 3]], "b": [None, [0, 0], [1, 2], [0, 1]]}})

# NullSemantics :602-617: a comparison with the NULL literal returns nothing
for pred in ("Equals", "NotEquals", "LessThan", "LessThanEquals", "GreaterThan", "GreaterThanEquals"):
    cases.append({"name": f"NullSemantics[{pred}]", "cite": f"{T}:603-611", "table": "int_float_with_null", "chunk_size": 2,
                  "encode": "none", "steps": [{"column": "a", "predicate": pred, "value": None, "null_literal": True}],
                  "expect_rows": 0})

# SingleScanReturnsCorrectRowCount :216-219 / DoubleScan's first scan :205-214: a >= 1234 on int_float.tbl -> int_float_filtered2.tbl
cases.append({"name": "SingleScanReturnsCorrectRowCount", "cite": f"{T}:216-219", "table": "int_float", "chunk_size": 2,
              "encode": "none", "steps": [{"column": "a", "predicate": "GreaterThanEquals", "value": 1234}],
              "expect_column": "a", "expect": load_tbl("int_float_filtered2")["data"]["a"]})
# EmptyResultScan :207-213
cases.append({"name": "EmptyResultScan", "cite": f"{T}:207-213", "table": "int_float", "chunk_size": 2, "encode": "none",
              "steps": [{"column": "a", "predicate": "GreaterThan", "value": 90000}], "expect_rows": 0})
# ScanOnWideDictionaryColumn :478-494: values 0..n in ONE dictionary chunk (get_table_op_with_n_dict_entries :85-101)
cases.append({"name": "ScanOnWideDictionaryColumn[16bit]", "cite": f"{T}:480-485", "generated": {"n": (1 << 8) + 1},
              "encode": "all", "steps": [{"column": "a", "predicate": "GreaterThan", "value": 200}], "expect_rows": 57})
cases.append({"name": "ScanOnWideDictionaryColumn[32bit]", "cite": f"{T}:488-493", "generated": {"n": (1 << 16) + 1},
              "encode": "all", "steps": [{"column": "a", "predicate": "GreaterThan", "value": 65500}], "expect_rows": 37})
# ScanWithExcludedFirstChunk :619-628 is TableScan::set_excluded_chunk_ids (the operator skips chunk 0): host-side, the
# parity tests model it by not passing that chunk.
cases.append({"name": "ScanWithExcludedFirstChunk", "cite": f"{T}:620-627", "table": "int_int_shuffled_2", "chunk_size": 5,
              "encode": [0, 1], "excluded_chunks": [0], "steps": [{"column": "a", "predicate": "GreaterThanEquals", "value": 0}],
              "expect_column": "b", "expect": [100, 102, 104, 106, 108, 110, 112, 102, 104]})
# refused predicate: Between "Will throw" :230,247-251
cases.append({"name": "ScanOnCompressedColumn[Between]", "cite": f"{T}:230,247-251", "table": "int_int_shuffled", "chunk_size": 7,
              "encode": [0, 1], "steps": [{"column": "a", "predicate": "Between", "value": 6}], "expect_error": True})

tables = {n: load_tbl(n) for n in ("int_int_shuffled", "int_int_shuffled_2", "int_int_w_null_8_rows", "int_float",
                                   "int_float_with_null")}
for t in tables.values():  # only the integer columns are scanned; keep the others out of the fixture
    for n, ty in zip(list(t["columns"]), list(t["types"])):
        if not ty.startswith("int"):
            del t["data"][n]
    t["columns"] = [n for n in t["columns"] if n in t["data"]]
    del t["types"]

# Scans on a float column (kept apart from `cases`: their tables carry per-column types)
typed_tables = {"int_float": load_tbl("int_float")}
float_cases = [
    # DoubleScan :211-221: a >= 1234, then b < 457.9 on the reference table -> int_float_filtered.tbl
    {"name": "DoubleScan", "cite": f"{T}:211-221", "table": "int_float", "chunk_size": 2, "encode": "none",
     "steps": [{"column": "a", "predicate": "GreaterThanEquals", "value": 1234},
               {"column": "b", "predicate": "LessThan", "value": 457.9}],
     "expect_column": "a", "expect": load_tbl("int_float_filtered")["data"]["a"]},
    # ScanWithEmptyInput :466-478: the second scan runs on an empty reference table
    {"name": "ScanWithEmptyInput", "cite": f"{T}:466-478", "table": "int_float", "chunk_size": 2, "encode": "none",
     "steps": [{"column": "a", "predicate": "GreaterThan", "value": 12345},
               {"column": "b", "predicate": "Equals", "value": 456.7}], "expect_rows": 0},
]

# String column scans, src/test/operators/table_scan_string_test.cpp (the comparison predicates; LIKE is out of scope):
# int_string_like.tbl in chunks of 5, every chunk encoded (:48-53), column b <op> literal, column a of the result
TS = "src/test/operators/table_scan_string_test.cpp"
typed_tables["int_string_like"] = load_tbl("int_string_like")
string_cases = [
    {"name": "ScanEquals", "cite": f"{TS}:74-80", "predicate": "Equals", "value": "Reeperbahn", "expected": "int_string_like_equals"},
    {"name": "ScanNotEquals", "cite": f"{TS}:82-89", "predicate": "NotEquals", "value": "Reeperbahn",
     "expected": "int_string_like_not_equals"},
    {"name": "ScanLessThan", "cite": f"{TS}:91-97", "predicate": "LessThan", "value": "Schiff", "expected": "int_string_like_less_than"},
]
string_cases = [{"name": c["name"], "cite": c["cite"], "table": "int_string_like", "chunk_size": 5, "encode": "all",
                 "steps": [{"column": "b", "predicate": c["predicate"], "value": c["value"]}], "expect_column": "a",
                 "expect": load_tbl(c["expected"])["data"]["a"]} for c in string_cases]

with open(os.path.join(HERE, "scan_fixtures.json"), "w") as f:
    json.dump({"source": "generated by tests/golden/make_scan_golden.py from /root/reference/" + T, "tables": tables,
               "cases": cases, "typed_tables": typed_tables, "float_cases": float_cases, "string_cases": string_cases}, f, indent=0,
              ensure_ascii=False)
print(len(cases), "cases,", len(tables), "tables,", len(float_cases), "float cases,", len(string_cases), "string cases")


# The same float / string cases for the C++ operator (hyrise-1_b200/host/shim_test, grammar of join_fixtures.txt):
# opossum::TableScan -> hj_scan / hj_scan_value_ids with HJ_SHIM_DEVICE_SCAN=1
def tbl_text(name):
    with open(os.path.join(TABLES, name + ".tbl"), encoding="utf-8") as f:
        return "@table " + name + "\n" + "".join(l for l in f if l.strip()).rstrip("\n") + "\n@end\n"


OPS = {"Equals": "=", "NotEquals": "!=", "LessThan": "<", "LessThanEquals": "<=", "GreaterThan": ">", "GreaterThanEquals": ">="}
shim_tables = ["int_float", "int_float_filtered", "int_string_like", "int_string_like_equals", "int_string_like_not_equals",
               "int_string_like_less_than"]
with open(os.path.join(HERE, "scan_shim_fixtures.txt"), "w", encoding="utf-8") as f:
    f.write("# GENERATED by tests/golden/make_scan_golden.py from the reference's TableScan tests. Do not edit by hand.\n")
    for t in shim_tables:
        f.write(tbl_text(t))
    f.write(f"@case TableScanTest.DoubleScan {T}:211-221\nbase t int_float 2\nscan s1 t 0 >= 1234\nscan s2 s1 1 < 457.9\n"
            "expect s2 int_float_filtered\n@end\n")
    f.write(f"@case TableScanTest.DoubleScanDict {T}:211-221\nbase t int_float 2 dict all\nscan s1 t 0 >= 1234\n"
            "scan s2 s1 1 < 457.9\nexpect s2 int_float_filtered\n@end\n")
    for name, pred, lit, expected in (("ScanEquals", "Equals", "Reeperbahn", "int_string_like_equals"),
                                      ("ScanNotEquals", "NotEquals", "Reeperbahn", "int_string_like_not_equals"),
                                      ("ScanLessThan", "LessThan", "Schiff", "int_string_like_less_than")):
        f.write(f"@case TableScanStringTest.{name} {TS}\nbase t int_string_like 5 dict all\nscan s t 1 {OPS[pred]} {lit}\n"
                f"expect s {expected}\n@end\n")
