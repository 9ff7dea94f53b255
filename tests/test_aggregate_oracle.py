"""The aggregate oracle (oracle/aggregate_oracle.py) against the reference's own expectations for Aggregates without
group-by columns (tests/golden/aggregate_fixtures.json, from src/test/operators/aggregate_test.cpp) and against a plain
Python restatement of the functors. CPU only."""
import numpy as np
import pytest

import aggregate_model as am
from oracle import aggregate_oracle as ao
from oracle import table_scan_oracle as tso

FIX = am.load_fixtures()
PRED = {"LessThan": tso.LESS_THAN, "GreaterThanEquals": tso.GREATER_THAN_EQUALS}


def rows_after(table, scans):
    """Row numbers that survive the scans (oracle TableScan on one unencoded chunk per scan, like the operator chain)."""
    rows = np.arange(len(next(iter(table["data"].values()))))
    for s in scans:
        v, n = am.column_of(table, s["column"])
        chunk = {"kind": "value", "values": v[rows], "nulls": n[rows] if n.any() else None}
        pos, _ = tso.table_scan([chunk], [0], PRED[s["predicate"]], s["value"])
        rows = rows[pos[:, 1]]
    return rows


@pytest.mark.parametrize("case", FIX["cases"], ids=[c["name"] for c in FIX["cases"]])
@pytest.mark.parametrize("variant", [0, 1], ids=["data", "reference"])
def test_oracle_reproduces_reference_expectation(case, variant):
    table = FIX["tables"][case["table"]]
    rows = rows_after(table, am.scans_of(case)[variant])
    v, n = am.column_of(table, case["column"])
    got = ao.aggregate_column(v[rows], n[rows])[case["function"]]
    assert am.matches_reference(got, case["expect"]), (got, case["expect"], case["cite"])


def functor_restated(values, is_null):
    """aggregate.cpp:141-201 + :203-252 row by row in plain Python (ints are Python ints, floats go through np.float64)."""
    floating = values.dtype.kind == "f"
    cur_sum = cur_min = cur_max = avg_sum = None
    count = 0
    for x, z in zip(values, is_null):
        if z:
            continue
        count += 1
        x = x.item()
        add = np.float64(x) if floating else x
        cur_sum = add if cur_sum is None else cur_sum + add
        avg_sum = np.float64(x) if avg_sum is None else avg_sum + np.float64(x)
        if cur_min is None or x < cur_min:
            cur_min = x
        if cur_max is None or x > cur_max:
            cur_max = x
    if not floating and cur_sum is not None:
        cur_sum = (cur_sum + 2**63) % 2**64 - 2**63   # int64_t accumulator
    return {"count_rows": len(values), "count": count, "sum": None if cur_sum is None else (float(cur_sum) if floating else cur_sum),
            "min": cur_min, "max": cur_max, "avg": None if count == 0 else float(avg_sum) / count}


@pytest.mark.parametrize("dtype", [np.int32, np.int64, np.float32, np.float64])
@pytest.mark.parametrize("null_share", [0.0, 0.3, 1.0])
def test_oracle_equals_the_functors_row_by_row(dtype, null_share):
    rng = np.random.default_rng(5)
    for n in (0, 1, 7, 1000):
        if np.dtype(dtype).kind == "f":
            v = (rng.standard_normal(n) * 1e3).astype(dtype)
        else:
            v = rng.integers(-10**6, 10**6, n).astype(dtype)
        z = (rng.random(n) < null_share).astype(np.uint8)
        assert ao.aggregate_column(v, z) == functor_restated(v, z)


def test_integer_sum_wraps_like_an_int64_accumulator_and_avg_runs_in_double():
    v = np.array([2**62, 2**62, 2**62], dtype=np.int64)
    got = ao.aggregate_column(v)
    assert got["sum"] == (3 * 2**62 + 2**63) % 2**64 - 2**63
    assert got["avg"] == float(2**62)
    assert functor_restated(v, np.zeros(3, np.uint8)) == got


# ---- with group-by columns: oracle groundwork for the device GROUP BY (no device path yet) --------------------
def rows_match(got, want):
    """check_table_equal, unordered (testing_assert.cpp:118-254): same multiset of rows, floats within EPSILON."""
    if len(got) != len(want):
        return False
    key = lambda row: tuple((v is None, 0 if v is None else round(float(v), 3)) for v in row)
    got, want = sorted(got, key=key), sorted(want, key=key)
    return all(len(g) == len(w) and all(am.matches_reference(a, b) for a, b in zip(g, w)) for g, w in zip(got, want))


@pytest.mark.parametrize("case", FIX["grouped_cases"], ids=[c["name"] for c in FIX["grouped_cases"]])
def test_grouped_oracle_reproduces_reference_expectation(case):
    table = FIX["grouped_tables"][case["table"]]
    columns = {name: am.column_of(table, name) for name in table["columns"]}
    got = ao.aggregate_grouped(columns, case["group_by"], [tuple(a) for a in case["aggregates"]])
    assert rows_match(got, case["expected_rows"]), (got, case["expected_rows"], case["cite"])
