"""C1/C4 inputs "as dbgen makes them": the reference's own tpch-dbgen compiled in place (oracle/_ref/libref_dbgen.so) against
the committed summary tests/golden/dbgen_keys_golden.json, against the closed forms the GPU config tests use where the
library is absent, and through the CPU oracle (lineitem x orders: every lineitem has exactly one order)."""
import json
import os
import zlib

import numpy as np
import pytest

import dbgen_binding as db
import oracle_binding as ob
import workloads as wl

pytestmark = pytest.mark.skipif(not db.available(), reason="oracle/_ref/libref_dbgen.so not built (needs /root/reference)")
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dbgen_keys_golden.json")


@pytest.fixture(scope="module")
def sf01():
    return db.tpch_keys(0.1)


@pytest.mark.parametrize("sf", ["0.01", "0.1"])
def test_dbgen_keys_match_golden_summary(sf):
    with open(GOLDEN) as f:
        want = json.load(f)[sf]
    got = db.tpch_keys(float(sf))
    for name, w in want.items():
        a = got[name]
        assert len(a) == w["n"], name
        assert int(a.astype(np.int64).sum()) == w["sum"], name
        assert zlib.crc32(a.tobytes()) == w["crc32"], name


def test_golden_has_the_published_tpch_row_counts():
    """TPC-H fixes these cardinalities: SF1 = 150 000 customers, 1 500 000 orders, 6 001 215 lineitems."""
    with open(GOLDEN) as f:
        g = json.load(f)
    assert g["1.0"]["c_custkey"]["n"] == 150_000 and g["1.0"]["o_orderkey"]["n"] == 1_500_000
    assert g["1.0"]["l_orderkey"]["n"] == 6_001_215 and g["0.1"]["l_orderkey"]["n"] == 600_572


def test_closed_forms_used_by_the_gpu_config_tests(sf01):
    """o_orderkey = mk_sparse(i) (build.c:142-146: 8 of every 32 values used); c_custkey = 1..n; o_custkey never divisible by
    3 (CUST_MORTALITY, dss.h:371, build.c:184); 1..7 lineitems per order, generated order by order."""
    i = np.arange(1, len(sf01["o_orderkey"]) + 1, dtype=np.int64)
    assert np.array_equal(sf01["o_orderkey"], ((i >> 3) << 5) | (i & 7))
    assert np.array_equal(sf01["c_custkey"], np.arange(1, len(sf01["c_custkey"]) + 1))
    assert not (sf01["o_custkey"] % 3 == 0).any()
    assert sf01["o_custkey"].min() >= 1 and sf01["o_custkey"].max() <= len(sf01["c_custkey"])
    assert sf01["o_lines"].min() == 1 and sf01["o_lines"].max() == 7
    assert np.array_equal(sf01["l_orderkey"], np.repeat(sf01["o_orderkey"], sf01["o_lines"]))


@pytest.mark.parametrize("sf", [0.1, 1.0])
def test_custkey_columns_alone_equal_the_full_generator(sf):
    """tpch_custkeys (what C4 at SF100 uses: only random stream O_CKEY_SD is run) == the columns mk_cust / mk_order produce."""
    full = db.tpch_keys(sf)
    c, oc = db.tpch_custkeys(sf)
    assert np.array_equal(c, full["c_custkey"]) and np.array_equal(oc, full["o_custkey"])


@pytest.mark.parametrize("encoding", ["value", "dict"])
def test_oracle_joins_dbgen_lineitem_with_orders(sf01, encoding):
    """C1 at SF 0.1 through the oracle: Inner on l_orderkey = o_orderkey, chunk size 100 000 (hyriseBenchmarkTPCH default)."""
    build = wl.make_relation(sf01["o_orderkey"], 100_000, encoding)
    probe = wl.make_relation(sf01["l_orderkey"], 100_000, encoding)
    b, p, info = ob.oracle_join(build, probe, ob.abi.MODE_INNER, n_threads=os.cpu_count() or 1)
    assert info["n_rows"] == len(sf01["l_orderkey"]) == 600_572
    ppos = p[:, 0].astype(np.int64) * 100_000 + p[:, 1]
    assert np.array_equal(np.sort(ppos), np.arange(len(ppos)))  # every lineitem exactly once
    bpos = b[:, 0].astype(np.int64) * 100_000 + b[:, 1]
    assert np.array_equal(sf01["o_orderkey"][bpos], sf01["l_orderkey"][ppos])


def test_oracle_c4_modes_on_dbgen_customer_orders(sf01):
    """C4 shape on real dbgen keys (SF 0.1): orders is the build side, customer probes (join_hash.cpp:58-70); one third of
    the customers never ordered."""
    build = wl.make_relation(sf01["o_custkey"], 100_000, "value")
    probe = wl.make_relation(sf01["c_custkey"], 100_000, "value")
    n_c = len(sf01["c_custkey"])
    has_order = np.isin(sf01["c_custkey"], sf01["o_custkey"])
    _, _, semi = ob.oracle_join(build, probe, ob.abi.MODE_SEMI)
    _, _, anti = ob.oracle_join(build, probe, ob.abi.MODE_ANTI)
    _, _, left = ob.oracle_join(build, probe, ob.abi.MODE_LEFT)
    assert semi["n_rows"] == int(has_order.sum()) and anti["n_rows"] == n_c - int(has_order.sum())
    assert left["n_rows"] == len(sf01["o_custkey"]) + anti["n_rows"]
    assert abs(anti["n_rows"] / n_c - 1 / 3) < 0.01
