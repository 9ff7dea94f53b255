"""Test-side plumbing for the ungrouped-aggregate cases of tests/golden/aggregate_fixtures.json (tests only): builds the
case's column, optionally behind the TableScan the reference's test_output puts in front (aggregate_test.cpp:100-107),
and compares like check_table_equal (testing_assert.cpp:21,103-110)."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TYPE_NAMES = {"int": np.int32, "long": np.int64, "float": np.float32, "double": np.float64}
EPSILON = 0.0001  # src/test/testing_assert.cpp:21


def load_fixtures():
    with open(os.path.join(ROOT, "tests", "golden", "aggregate_fixtures.json")) as f:
        return json.load(f)


def column_of(table, name):
    """(values, is_null) of one column of a fixture table, in the column's type."""
    ty = table["types"][table["columns"].index(name)].replace("_null", "")
    col = table["data"][name]
    values = np.array([0 if v is None else v for v in col], dtype=TYPE_NAMES[ty])
    return values, np.array([v is None for v in col], dtype=np.uint8)


def matches_reference(got, want):
    if want is None or got is None:
        return got is None and want is None
    if isinstance(want, float) or isinstance(got, float):
        return abs(got - want) < max(EPSILON, abs(want * EPSILON))
    return got == want


def scans_of(case):
    """The scans in front of the aggregate: the case's own filter, and the reference-table variant `a >= 0` of test_output."""
    own = [case["filter"]] if "filter" in case else []
    return [own, own + [{"column": "a", "predicate": "GreaterThanEquals", "value": 0}]]
