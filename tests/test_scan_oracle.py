"""The TableScan oracle (oracle/table_scan_oracle.py) against the reference's own expectations: every integer-column case
of src/test/operators/table_scan_test.cpp as restated in tests/golden/scan_fixtures.json, under each chunk encoding the
device scan supports. This pins the oracle the GPU parity tests (tests/test_gpu_scan.py) compare against."""
import numpy as np
import pytest

import scan_model as sm
from oracle import table_scan_oracle as tso

FIX = sm.load_fixtures()
CASES = [(c, e) for c in FIX["cases"] for e in sm.encodings_for(c)]
IDS = [f"{c['name']}-{e if isinstance(e, str) else 'chunks' + ''.join(map(str, e))}" for c, e in CASES]


def oracle_scan(chunks, ids, pred, value):
    return tso.table_scan(chunks, ids, pred, value)


@pytest.mark.parametrize("case,encode", CASES, ids=IDS)
def test_oracle_reproduces_reference_expectation(case, encode):
    if case.get("expect_error"):
        with pytest.raises(ValueError):
            sm.run_case(case, FIX["tables"], oracle_scan, encode)
        return
    assert sm.run_case(case, FIX["tables"], oracle_scan, encode) == sm.expected_of(case), case["cite"]


ENCODED = [(ce, i) for ce, i in zip(CASES, IDS) if ce[1] != "none" and not ce[0].get("expect_error")]


@pytest.mark.parametrize("style", ["for", "rle"])
@pytest.mark.parametrize("case,encode", [ce for ce, _ in ENCODED], ids=[i for _, i in ENCODED])
def test_oracle_reproduces_reference_expectation_for_and_rle(case, encode, style):
    """EncodingType::FrameOfReference / RunLength of the reference's INSTANTIATE_TEST_CASE_P (table_scan_test.cpp:211-214)."""
    assert sm.run_case(case, FIX["tables"], oracle_scan, encode, style) == sm.expected_of(case), case["cite"]


def test_dictionary_value_id_scan_equals_value_scan_on_random_chunks():
    """The value-id formulation (single_column_table_scan_impl.cpp:83-199) and the plain comparison agree, NULLs included."""
    rng = np.random.default_rng(7)
    for trial in range(200):
        n = int(rng.integers(0, 60))
        values = rng.integers(-5, 12, n).astype(np.int32)
        nulls = (rng.random(n) < 0.2).astype(np.uint8)
        d = sm.dict_encode(values, nulls)
        v = {"kind": "value", "values": values, "nulls": nulls}
        for pred in list(range(6)) + [10, 11]:
            for lit in (-7, -5, 0, 3, 11, 20, None):
                assert np.array_equal(tso.scan_chunk(d, pred, lit), tso.scan_chunk(v, pred, lit)), (trial, pred, lit)


def test_dereference_emits_the_input_pos_list_entries():
    t = sm.DataTable(FIX["tables"]["int_int_shuffled_2"]["data"], 5, [0, 1])
    pl = sm.pos_array([[2, 0], [1, 1], None, [0, 2]])
    chunk = {"kind": "reference", "pos_list": pl, "referenced": t.column_chunks("a")}
    pos, begin = tso.table_scan([chunk], [0], tso.IS_NOT_NULL, None, dereference=True)
    assert np.array_equal(pos, pl[[0, 1, 3]]) and list(begin) == [0, 3]
    pos, _ = tso.table_scan([chunk], [0], tso.IS_NULL, None, dereference=True)
    assert np.array_equal(pos, pl[[2]])


# ---- float / double columns ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", FIX["float_cases"], ids=[c["name"] for c in FIX["float_cases"]])
@pytest.mark.parametrize("encode,style", [("none", "dict"), ("all", "dict"), ("all", "rle")])
def test_oracle_reproduces_reference_expectation_on_float_columns(case, encode, style):
    assert sm.run_case(case, FIX["typed_tables"], oracle_scan, encode, style) == sm.expected_of(case), case["cite"]


def float_chunks(rng, n_rows, chunk, dtype, encoding, nullable):
    """Float keys from a small domain incl. -0.0, NaN and +-inf, so that every predicate has matches and ties."""
    domain = np.array([-np.inf, -7.25, -0.0, 0.0, 0.1, 1.0, 457.7, 457.9, 3.0e9, np.inf, np.nan], dtype=dtype)
    values = domain[rng.integers(0, len(domain), n_rows)]
    nulls = (rng.random(n_rows) < 0.1).astype(np.uint8) if nullable else np.zeros(n_rows, np.uint8)
    out = []
    for c, lo in enumerate(range(0, max(n_rows, 1), chunk)):
        v, nl = np.ascontiguousarray(values[lo:lo + chunk]), np.ascontiguousarray(nulls[lo:lo + chunk])
        enc = encoding if encoding != "mixed" else ("dict" if c % 2 else "value")
        if enc == "dict":   # a dictionary holds no NaN in these tests: std::sort on NaN keys is undefined in the reference
            v = np.where(np.isnan(v), dtype(2.5), v).astype(dtype)
            out.append(sm.dict_encode(v, nl))
        else:
            out.append({"kind": "value", "values": v, "nulls": nl if nullable else None})
    return out


FLOAT_LITERALS = (-np.inf, -8, -7.25, 0, 0.1, 457.8, 457.9, 3000000000, 1e30)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("encoding", ["value", "dict", "mixed"])
def test_float_scan_equals_a_row_by_row_comparison_in_the_column_type(dtype, encoding):
    """single_column_table_scan_impl.cpp:36-81: the literal becomes the column type, rows compare with the C++ operator
    of that type — NaN matches only !=, -0.0 == 0.0, 0.1f != 0.1."""
    rng = np.random.default_rng(11)
    ops = {0: lambda a, b: a == b, 1: lambda a, b: a != b, 2: lambda a, b: a < b, 3: lambda a, b: a <= b,
           4: lambda a, b: a > b, 5: lambda a, b: a >= b}
    chunks = float_chunks(rng, 700, 256, dtype, encoding, True)
    for pred, op in ops.items():
        for lit in FLOAT_LITERALS:
            pos, begin = tso.table_scan(chunks, list(range(len(chunks))), pred, lit)
            want = []
            for cid, ch in enumerate(chunks):
                v, n = tso.decode(ch)
                with np.errstate(over="ignore"):
                    right = dtype(lit)
                want += [(cid, i) for i in range(len(v)) if not n[i] and op(v[i], right)]
            assert [tuple(r) for r in pos.tolist()] == want, (pred, lit)


def test_floating_literal_on_an_integer_column_is_truncated_like_numeric_cast():
    """type_cast.hpp:48-57: `a < 457.9` on an int column scans `a < 457`; -0.5 becomes 0; outside the range it throws."""
    chunks = [{"kind": "value", "values": np.array([455, 456, 457, 458, -1, 0], np.int32), "nulls": None}]
    pos, _ = tso.table_scan(chunks, [0], tso.LESS_THAN, 457.9)
    assert pos[:, 1].tolist() == [0, 1, 4, 5]
    pos, _ = tso.table_scan(chunks, [0], tso.EQUALS, -0.5)
    assert pos[:, 1].tolist() == [5]
    with pytest.raises(OverflowError):
        tso.table_scan(chunks, [0], tso.EQUALS, 3.0e9)


# ---- string columns (dictionary chunks are scanned on value ids; the oracle compares the strings) -----------------------
@pytest.mark.parametrize("case", FIX["string_cases"], ids=[c["name"] for c in FIX["string_cases"]])
@pytest.mark.parametrize("encode,style", [("none", "dict"), ("all", "dict"), ("all", "rle")])
def test_oracle_reproduces_reference_expectation_on_string_columns(case, encode, style):
    assert sm.run_case(case, FIX["typed_tables"], oracle_scan, encode, style) == sm.expected_of(case), case["cite"]


def value_id_bounds(chunks, literal):
    """What the caller of hj_scan_value_ids does per dictionary chunk: BaseDictionaryColumn::lower_bound / upper_bound."""
    return np.array([[np.searchsorted(c["dictionary"], literal, side="left"), np.searchsorted(c["dictionary"], literal, side="right")]
                     for c in chunks], dtype=np.uint32).reshape(-1, 2)


def string_chunks(rng, n_rows, chunk, n_words, nullable=True):
    words = np.array(sorted({"".join(rng.choice(list("abcXYZ09 é"), rng.integers(0, 6))) for _ in range(n_words * 3)})[:n_words])
    values = words[rng.integers(0, len(words), n_rows)]
    nulls = (rng.random(n_rows) < 0.1).astype(np.uint8) if nullable else np.zeros(n_rows, np.uint8)
    return [sm.dict_encode(np.ascontiguousarray(values[lo:lo + chunk]), np.ascontiguousarray(nulls[lo:lo + chunk]))
            for lo in range(0, max(n_rows, 1), chunk)], words


def test_value_id_comparison_with_host_side_bounds_equals_comparing_the_strings():
    """single_column_table_scan_impl.cpp:83-133 needs only (lower bound, upper bound, attribute vector): the rows whose
    value id passes the id comparison are the rows whose STRING passes the string comparison."""
    rng = np.random.default_rng(3)
    chunks, words = string_chunks(rng, 3000, 700, 40)
    ops = {0: lambda a, b: a == b, 1: lambda a, b: a != b, 2: lambda a, b: a < b, 3: lambda a, b: a <= b,
           4: lambda a, b: a > b, 5: lambda a, b: a >= b}
    for lit in ["", "a", str(words[7]), str(words[-1]), "zzzz", "Xé"]:
        bounds = value_id_bounds(chunks, lit)
        for pred, op in ops.items():
            pos, _ = tso.table_scan(chunks, list(range(len(chunks))), pred, lit)
            want = []
            for cid, (c, (lb, ub)) in enumerate(zip(chunks, bounds.tolist())):
                attr = c["attr"].astype(np.int64)
                ok = {0: (attr == lb) & (lb != ub), 1: (attr != lb) | (lb == ub), 2: attr < lb, 3: attr < ub, 4: attr >= ub,
                      5: attr >= lb}[pred] & (attr != c["null_value_id"])
                want += [(cid, int(i)) for i in np.flatnonzero(ok)]
                v, n = tso.decode(c)
                assert [int(i) for i in np.flatnonzero(ok)] == [i for i in range(len(v)) if not n[i] and op(str(v[i]), lit)]
            assert [tuple(r) for r in pos.tolist()] == want, (pred, lit)
