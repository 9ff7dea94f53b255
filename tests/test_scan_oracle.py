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
