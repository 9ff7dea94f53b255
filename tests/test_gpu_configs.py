"""BASELINE.json's configs as GPU parity cases.

C1/C3/C4 shapes run at sizes the oracle finishes in seconds and are compared bit-exactly (sorted multiset of RowID pairs);
C2 and C3 additionally run at (or near) their published sizes, where the result is checked through size-independent
properties computed independently with torch: exact row count, a checksum over the emitted RowIDs, and key equality of
sampled result rows. Everything goes through the C ABI (`-m gpu`)."""
import importlib
import os
import sys

import numpy as np
import pytest

import oracle_binding as ob
import workloads as wl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def join():
    j = hj.HashJoin(device=0)
    yield j
    j.close()


def assert_same(j, build, probe, mode, what):
    mode_id = abi.MODES[mode]
    info = j.join(build, probe, mode_id)
    b, p = j.fetch()
    ob_b, ob_p, oinfo = ob.oracle_join(build, probe, mode_id, n_threads=os.cpu_count() or 1)
    assert info.n_rows == oinfo["n_rows"], f"{what} {mode}: {info.n_rows} rows, oracle {oinfo['n_rows']}"
    got, want = ob.canonical(b, p), ob.canonical(ob_b, ob_p)
    assert np.array_equal(got, want), f"{what} {mode}: sorted result rows differ"
    return info


# ------------------------------------------------------------------------------------------------------------------ C1
def dbgen_orderkeys(n_orders):
    """o_orderkey as dbgen makes it: mk_sparse keeps the low 3 bits of the order index and spreads the rest so that 8 of
    every 32 key values are used (third_party/tpch-dbgen/build.c:142-146, dss.h:144-145; SF1: 1.5 M orders, max 6 000 000)."""
    i = np.arange(1, n_orders + 1, dtype=np.int64)
    return (((i >> 3) << 5) | (i & 7)).astype(np.int32)


@pytest.mark.parametrize("encoding", ["value", "dict"])
def test_c1_tpch_sf1_lineitem_orders(join, encoding):
    """C1: TPC-H SF1 lineitem x orders on l_orderkey = o_orderkey, Inner, int32, chunk size 100 000
    (tpch_db_generator.cpp:219-236 shapes: 1.5 M unique sparse order keys, 1..7 lineitems per order)."""
    rng = np.random.default_rng(0xB201)
    okeys = dbgen_orderkeys(1_500_000)
    lines = rng.integers(1, 8, size=len(okeys))
    lkeys = np.repeat(okeys, lines)  # lineitem is generated order by order
    build = wl.make_relation(okeys, 100_000, encoding)
    probe = wl.make_relation(lkeys, 100_000, encoding)
    info = assert_same(join, build, probe, "Inner", f"C1 {encoding}")
    assert info.n_rows == len(lkeys)  # every lineitem has exactly one order


@pytest.fixture(scope="module")
def dbgen_sf1():
    import dbgen_binding as db
    if not db.available():
        pytest.skip("oracle/_ref/libref_dbgen.so not built")
    return db.tpch_keys(1.0)


@pytest.mark.parametrize("encoding", ["value", "dict"])
def test_c1_tpch_sf1_dbgen_keys(join, dbgen_sf1, encoding):
    """C1 on the key columns the reference's own dbgen generates at SF1 (oracle/_ref/libref_dbgen.so = third_party/tpch-dbgen
    compiled in place and driven like tpch_db_generator.cpp:203-237; it travels to the GPU box as a built file):
    1 500 000 orders x 6 001 215 lineitems, chunk size 100 000, bit-exact vs the oracle."""
    k = dbgen_sf1
    assert len(k["l_orderkey"]) == 6_001_215
    build = wl.make_relation(k["o_orderkey"], 100_000, encoding)
    probe = wl.make_relation(k["l_orderkey"], 100_000, encoding)
    info = assert_same(join, build, probe, "Inner", f"C1 dbgen {encoding}")
    assert info.n_rows == 6_001_215


# ------------------------------------------------------------------------------------------------------------------ C3
@pytest.mark.parametrize("mode", ["Inner", "Left", "Semi", "Anti"])
def test_c3_zipf_int64_scaled(join, mode):
    """C3 shape at 1/100: Zipf(1.0)-skewed int64 build keys (hottest key ~5 % of the rows => build partitions far larger
    than one shared-memory table, heavy duplicate expansion), uniform probe."""
    rng = np.random.default_rng(0xB203)
    n_b, n_p = 2_000_000, 20_000_000 // (1 if mode != "Inner" else 4)
    bk = wl.zipf_keys(rng, n_b, n_b, 1.0, np.int64)
    pk = rng.integers(1, n_b + 1, n_p, dtype=np.int64)
    info = assert_same(join, wl.make_relation(bk, 1 << 20), wl.make_relation(pk, 1 << 20), mode, "C3/100")
    assert info.n_overflow_partitions >= 1


# ------------------------------------------------------------------------------------------------------------------ C4
@pytest.mark.parametrize("encoding", ["value", "dict"])
@pytest.mark.parametrize("mode", ["Left", "Semi", "Anti"])
def test_c4_customer_orders_nulls(join, mode, encoding):
    """C4 shape at 1/10 (TPC-H SF10 sizes): customer (1.5 M unique c_custkey, left input = probe side for Left/Semi/Anti)
    x orders (15 M o_custkey; customers with custkey % 3 == 0 never order, dss.h:371 CUST_MORTALITY), 1 % NULL keys on
    both sides."""
    rng = np.random.default_rng(0xB204)
    n_cust, n_ord = 1_500_000, 15_000_000
    ckeys = np.arange(1, n_cust + 1, dtype=np.int32)
    okeys = rng.integers(1, n_cust + 1, n_ord).astype(np.int32)
    okeys[okeys % 3 == 0] -= 1  # never 0: 3 -> 2, ...
    cn = (rng.random(n_cust) < 0.01).astype(np.uint8)
    on = (rng.random(n_ord) < 0.01).astype(np.uint8)
    build = wl.make_relation(okeys, 1 << 20, encoding, on)   # orders build, customer probes (join_hash.cpp:58-70)
    probe = wl.make_relation(ckeys, 1 << 20, encoding, cn)
    info = assert_same(join, build, probe, mode, f"C4/10 {encoding}")
    non_null_c = int((cn == 0).sum())
    if mode == "Anti":  # NULL customers are NOT returned (quirk Q5); about 1/3 of the others never ordered
        assert 0.30 * non_null_c < info.n_rows < 0.37 * non_null_c
    if mode == "Semi":
        assert 0.63 * non_null_c < info.n_rows < 0.70 * non_null_c


def _pair_digest(b, p):
    """Order-independent digest of a result: (rows, wrapping sum, xor) of a 64-bit mix of every (build RowID, probe RowID) row.
    Equal multisets give equal digests; used where sorting 150 M rows twice would take longer than the join by far."""
    x = p.reshape(-1, 2).astype(np.uint64)
    h = (x[:, 0] << np.uint64(32) | x[:, 1]) * np.uint64(0x9E3779B97F4A7C15)
    if b is not None:
        y = b.reshape(-1, 2).astype(np.uint64)
        h ^= ((y[:, 0] << np.uint64(32) | y[:, 1]) + np.uint64(0x632BE59BD9B4E019)) * np.uint64(0xC2B2AE3D27D4EB4F)
    h ^= h >> np.uint64(29)
    h *= np.uint64(0xBF58476D1CE4E5B9)
    h ^= h >> np.uint64(32)
    return len(x), int(h.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(h)) if len(h) else 0


@pytest.fixture(scope="module")
def c4_sf100():
    """C4 as published: TPC-H SF100 customer (15 M c_custkey) x orders (150 M o_custkey) on the keys the reference's own dbgen
    draws (tests/dbgen_binding.py tpch_custkeys: random stream O_CKEY_SD + CUST_MORTALITY, equal to mk_order's column), 1 % of
    the keys of both sides set NULL by a seeded generator."""
    import dbgen_binding as db
    if not db.available():
        pytest.skip("oracle/_ref/libref_dbgen.so not built")
    ckeys, okeys = db.tpch_custkeys(100.0)
    assert len(ckeys) == 15_000_000 and len(okeys) == 150_000_000
    rng = np.random.default_rng(0xB204)
    cn = (rng.random(len(ckeys)) < 0.01).astype(np.uint8)
    on = (rng.random(len(okeys)) < 0.01).astype(np.uint8)
    return ckeys, okeys, cn, on


@pytest.mark.parametrize("mode,encoding", [("Left", "value"), ("Semi", "value"), ("Anti", "dict")])
def test_c4_sf100_customer_orders_dbgen_keys(join, c4_sf100, mode, encoding):
    """C4 at its published size against the oracle: orders build, customer probes (join_hash.cpp:58-70). Semi / Anti: sorted
    RowID lists bit-exact; Left (150 M pairs): row count + an order-independent 64-bit digest of all (build, probe) RowID rows."""
    ckeys, okeys, cn, on = c4_sf100
    build = wl.make_relation(okeys, 1 << 20, encoding, on)
    probe = wl.make_relation(ckeys, 1 << 20, encoding, cn)
    mode_id = abi.MODES[mode]
    info = join.join(build, probe, mode_id)
    b, p = join.fetch()
    ob_b, ob_p, oinfo = ob.oracle_join(build, probe, mode_id, n_threads=os.cpu_count() or 1)
    assert info.n_rows == oinfo["n_rows"], f"C4 SF100 {mode} {encoding}: {info.n_rows} rows, oracle {oinfo['n_rows']}"
    assert info.n_build_tuples == oinfo["n_build_tuples"] and info.n_probe_tuples == oinfo["n_probe_tuples"]
    if mode == "Left":
        assert _pair_digest(b, p) == _pair_digest(ob_b, ob_p), f"C4 SF100 Left {encoding}: result digests differ"
    else:
        assert np.array_equal(ob.canonical(b, p), ob.canonical(ob_b, ob_p)), f"C4 SF100 {mode} {encoding}: sorted RowIDs differ"
    non_null_c = int((cn == 0).sum())
    if mode == "Anti":  # quirk Q5: NULL customers are not returned; a third of the customers never ordered (CUST_MORTALITY)
        assert 0.32 * non_null_c < info.n_rows < 0.35 * non_null_c
    join.release_scratch()


# ------------------------------------------------------------------------------------------------------- full-size properties
def _device_case(name, scale):
    import torch
    import bench
    dev = torch.device("cuda", 0)
    data = bench.gen_device_workload(name, scale, dev, seed=0xB200 + 7)
    b = bench.relation_from_chunks(hj, data["build"], abi.LOC_DEVICE)
    p = bench.relation_from_chunks(hj, data["probe"], abi.LOC_DEVICE)
    return data, b, p


def _check_properties(j, data, b, p, name):
    """Exact row count, RowID checksums of both sides and key equality of 1 M sampled result rows of an Inner join, computed
    independently with torch on the PosLists where the join left them (bench.verify_result: nothing is copied to the host,
    everything chunk by chunk, so it fits beside the join's buffers at the full size of C3)."""
    import torch
    import bench
    info = j.join(b, p, abi.MODE_INNER)
    v = bench.verify_result(j, data, info, name, 0, 1, torch.device("cuda", 0))
    assert v["rows"] == v["rows_expected"] == info.n_rows, v
    assert v["checksums"] and v["sampled_key_mismatches"] == 0 and v["ok"], v
    return info


@pytest.mark.parametrize("scale", [1.0])
def test_c2_full_size_properties(join, scale):
    """C2 as published: 100 M x 1 B int32, dictionary-encoded, 1 e9 result pairs."""
    data, b, p = _device_case("c2", scale)
    info = _check_properties(join, data, b, p, "c2")
    assert info.n_rows == data["n_probe"] and info.n_probe_passes == 1
    join.release_scratch()


def test_c3_full_size_properties(join):
    """C3 as published: 200 M Zipf(1.0) int64 build x 2 B uniform int64 probe, ValueColumns (~120 GB of the 180 GB HBM with
    the join's buffers): ~2e9 pairs, the hottest key ~1e7 build duplicates."""
    import torch
    join.release_scratch()
    torch.cuda.empty_cache()
    data, b, p = _device_case("c3", 1.0)
    info = _check_properties(join, data, b, p, "c3")
    assert info.n_overflow_partitions >= 1 and 1.9e9 < info.n_rows < 2.1e9
    join.release_scratch()
    del data, b, p
    torch.cuda.empty_cache()


def test_c3_tenth_size_properties(join):
    """C3 at 1/10 (20 M x 200 M int64, Zipf build): ~2e8 pairs, hottest key ~1e6 duplicates expanded by whole warps."""
    data, b, p = _device_case("c3", 0.1)
    info = _check_properties(join, data, b, p, "c3")
    assert info.n_overflow_partitions >= 1
    join.release_scratch()
