"""Ungrouped aggregates on the device (hj_aggregate_column) through the C ABI: the reference's own cases —
on the data table and, like its test_output, behind a TableScan (device scan -> hj_gather_column -> hj_aggregate_column,
nothing but the one result row leaves the GPU) — and randomized columns against the oracle. Integer results, counts,
MIN and MAX are exact; double sums are folded in another order than the reference's row order and must agree to 1e-12
relative (tolerance of this test; the reference's own table comparison allows 1e-4, testing_assert.cpp:21).

First run on a B200 in round 2 (profiles/r02_gpu_call1_gated_suites.txt)."""
import importlib
import os
import sys

import numpy as np
import pytest

import aggregate_model as am
from oracle import aggregate_oracle as ao

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
pytestmark = pytest.mark.gpu
FIX = am.load_fixtures()
KEY_TYPES = {"int32": abi.KEY_INT32, "int64": abi.KEY_INT64, "float32": abi.KEY_FLOAT, "float64": abi.KEY_DOUBLE}
SUM_RTOL = 1e-12


@pytest.fixture(scope="module")
def ctx():
    j = hj.HashJoin(device=0)
    yield j
    j.close()


def device_column(values, is_null, chunk):
    """A device-resident value column in chunks of `chunk` rows: (relation, tensors to keep alive)."""
    import torch
    dev = torch.device("cuda", 0)
    rb, keep = hj.RelationBuilder(abi.LOC_DEVICE), []
    for c, lo in enumerate(range(0, max(len(values), 1), chunk)):
        v = torch.from_numpy(values[lo:lo + chunk].copy()).to(dev)
        z = torch.from_numpy(is_null[lo:lo + chunk].copy()).to(dev)
        keep += [v, z]
        rb.add_value(c, v, z)
    return rb.build(), keep


def close_enough(got, want):
    for k in ("count_rows", "count", "min", "max"):
        assert got[k] == want[k], (k, got[k], want[k])
    for k in ("sum", "avg"):
        if want[k] is None or isinstance(want[k], int):
            assert got[k] == want[k], (k, got[k], want[k])
        else:
            assert got[k] == pytest.approx(want[k], rel=SUM_RTOL, abs=0.0), (k, got[k], want[k])


@pytest.mark.parametrize("case", FIX["cases"], ids=[c["name"] for c in FIX["cases"]])
@pytest.mark.parametrize("variant", [0, 1], ids=["data", "reference"])
def test_reference_aggregate_case_on_gpu(ctx, case, variant):
    table = FIX["tables"][case["table"]]
    scans = am.scans_of(case)[variant]
    v, z = am.column_of(table, case["column"])
    col, keep = device_column(v, z, case["chunk_size"])
    n_rows = len(v)
    if not scans:
        # the data table itself: all rows in chunk order = the identity PosList
        import torch
        pos = torch.tensor([[c, o] for c, lo in enumerate(range(0, n_rows, case["chunk_size"]))
                            for o in range(min(case["chunk_size"], n_rows - lo))], dtype=torch.int32, device="cuda:0").reshape(-1, 2)
        pos_ptr, n = pos.data_ptr(), n_rows
    else:
        # TableScan chain on the device; every scan after the first reads its column through the previous PosList
        rel = None
        for i, s in enumerate(scans):
            sv, sz = am.column_of(table, s["column"])
            scol, k = device_column(sv, sz, case["chunk_size"])
            keep += k
            if i == 0:
                ctx.scan(scol, s["predicate"], s["value"], slot=i)
            else:
                ptr, _, n_prev, _ = ctx.scan_result_device(i - 1)
                import torch

                class W:
                    __cuda_array_interface__ = {"shape": (max(n_prev, 1), 2), "typestr": "<i4", "data": (ptr, False), "version": 2}
                view = torch.as_tensor(W(), device="cuda:0")[:n_prev]
                keep.append(view)
                rel = hj.RelationBuilder(abi.LOC_DEVICE).add_reference(0, view, scol, KEY_TYPES[sv.dtype.name]).build()
                ctx.scan(rel, s["predicate"], s["value"], slot=i, dereference=True)
        pos_ptr, _, n, _ = ctx.scan_result_device(len(scans) - 1)
    values_ptr, nulls_ptr, kt = ctx.gather_column(pos_ptr, n, col)
    assert kt == KEY_TYPES[v.dtype.name]
    got = ctx.aggregate_column(values_ptr, nulls_ptr, n, kt)
    assert am.matches_reference(got[case["function"]], case["expect"]), (got, case["expect"], case["cite"])


@pytest.mark.parametrize("dtype", [np.int32, np.int64, np.float32, np.float64])
@pytest.mark.parametrize("null_share", [0.0, 0.3, 1.0])
def test_random_columns_match_oracle(ctx, dtype, null_share):
    """Sizes around a warp, a CTA, the grid-stride unrolling and one past them; NULL-free (nulls = NULL pointer), partly and
    all NULL."""
    import torch
    rng = np.random.default_rng(0xA66)
    for n in (0, 1, 31, 33, 255, 257, 1023, 4 * 256 * 148 * 8 + 5, 3_000_001):
        if np.dtype(dtype).kind == "f":
            v = (rng.standard_normal(n) * 1e3).astype(dtype)
        else:
            v = rng.integers(-10**6, 10**6, n).astype(dtype)
        z = (rng.random(n) < null_share).astype(np.uint8)
        tv, tz = torch.from_numpy(v).cuda(), torch.from_numpy(z).cuda()
        got = ctx.aggregate_column(tv.data_ptr() if n else None, tz.data_ptr() if (n and null_share > 0) else None, n,
                                   KEY_TYPES[np.dtype(dtype).name])
        close_enough(got, ao.aggregate_column(v, z if null_share > 0 else None))


def test_aggregate_of_the_join_result_without_downloading_it(ctx):
    """scan-free pipeline at a size where the download would matter: join, read a payload column of the probe table through
    the probe-side PosList, SUM / COUNT / MIN / MAX it; compared with numpy over the fetched PosLists."""
    import torch
    rng = np.random.default_rng(9)
    chunk = 1 << 16
    bk = (rng.permutation(300_000) + 1).astype(np.int32)
    pk = rng.integers(1, 400_001, 2_000_000).astype(np.int32)
    pay = rng.integers(-1000, 1000, len(pk)).astype(np.int64)
    pay_null = (rng.random(len(pk)) < 0.05).astype(np.uint8)
    b_col, k1 = device_column(bk, np.zeros(len(bk), np.uint8), chunk)
    p_col, k2 = device_column(pk, np.zeros(len(pk), np.uint8), chunk)
    pay_col, k3 = device_column(pay, pay_null, chunk)
    info = ctx.join(b_col, p_col, abi.MODE_INNER)
    _, p_ptr, n = ctx.result_device()
    assert n == info.n_rows == int((pk <= 300_000).sum())
    values_ptr, nulls_ptr, kt = ctx.gather_column(p_ptr, n, pay_col)
    got = ctx.aggregate_column(values_ptr, nulls_ptr, n, kt)
    _, gp = ctx.fetch()
    rows = gp[:, 0].astype(np.int64) * chunk + gp[:, 1]
    close_enough(got, ao.aggregate_column(pay[rows], pay_null[rows]))


# ---- hj_aggregate_relation: the column as chunk descriptors (what the host-side Aggregate operator passes) ---------------
def host_relation(values, is_null, chunk, encoding):
    import scan_model as sm
    rb = hj.RelationBuilder(abi.LOC_HOST)
    for c, lo in enumerate(range(0, max(len(values), 1), chunk)):
        v, z = np.ascontiguousarray(values[lo:lo + chunk]), np.ascontiguousarray(is_null[lo:lo + chunk])
        if encoding == "dict":
            d = sm.dict_encode(v, z)
            rb.add_dict(c, d["dictionary"], d["attr"], d["null_value_id"])
        else:
            rb.add_value(c, v, z if z.any() else None)
    return rb.build()


@pytest.mark.parametrize("case", FIX["cases"], ids=[c["name"] for c in FIX["cases"]])
@pytest.mark.parametrize("encoding", ["value", "dict"])
@pytest.mark.parametrize("variant", [0, 1], ids=["data", "reference"])
def test_reference_aggregate_case_through_chunk_descriptors(ctx, case, encoding, variant):
    """Host-resident tables like the C++ operator sees them; behind scans the column is a REFERENCE chunk whose PosList is
    the (oracle) scan result — the reference's `Aggregate(TableScan(...))`."""
    from oracle import table_scan_oracle as tso
    from test_aggregate_oracle import PRED
    table = FIX["tables"][case["table"]]
    v, z = am.column_of(table, case["column"])
    chunk = case["chunk_size"]
    data = host_relation(v, z, chunk, encoding)
    scans = am.scans_of(case)[variant]
    if not scans:
        rel = data
    else:
        rows = np.arange(len(v))
        for s in scans:
            sv, sz = am.column_of(table, s["column"])
            pos, _ = tso.table_scan([{"kind": "value", "values": sv[rows], "nulls": sz[rows] if sz.any() else None}], [0],
                                    PRED[s["predicate"]], s["value"])
            rows = rows[pos[:, 1]]
        pos_list = np.ascontiguousarray(np.stack([rows // chunk, rows % chunk], axis=1).astype(np.uint32)).reshape(-1, 2)
        rel = hj.RelationBuilder(abi.LOC_HOST).add_reference(0, pos_list, data, KEY_TYPES[v.dtype.name]).build()
    got = ctx.aggregate_relation(rel)
    assert am.matches_reference(got[case["function"]], case["expect"]), (got, case["expect"], case["cite"])


@pytest.mark.parametrize("dtype", [np.int32, np.int64, np.float32, np.float64])
@pytest.mark.parametrize("encoding", ["value", "dict"])
def test_random_relations_match_oracle(ctx, dtype, encoding):
    rng = np.random.default_rng(0xA67)
    for n, chunk in ((0, 8), (1, 8), (4097, 4096), (100_000, 30_000), (2_000_000, 1 << 18)):
        if np.dtype(dtype).kind == "f":
            v = (rng.integers(-5000, 5000, n) / 8).astype(dtype)   # few distinct values: small dictionaries
        else:
            v = rng.integers(-5000, 5000, n).astype(dtype)
        z = (rng.random(n) < 0.2).astype(np.uint8)
        close_enough(ctx.aggregate_relation(host_relation(v, z, chunk, encoding)), ao.aggregate_column(v, z))


def test_scan_join_translate_gather_aggregate_stays_on_the_device(ctx):
    """SELECT COUNT(*), SUM(p.pay), MIN(p.pay), MAX(p.pay) FROM build b JOIN probe p ON b.key = p.key WHERE p.key >= :lit —
    the probe key column filtered by hj_scan, its PosLists joined as REFERENCE chunks, the join's probe side translated to
    rows of the probe table (hj_translate_device, A7), the payload column read through it and aggregated: only the
    aggregate comes back. Compared with numpy."""
    import torch
    rng = np.random.default_rng(41)
    chunk = 1 << 16
    dev = torch.device("cuda", 0)
    bk = (rng.permutation(300_000) + 1).astype(np.int32)
    pk = rng.integers(1, 400_001, 1_500_000).astype(np.int32)
    pay = rng.integers(-1000, 1000, len(pk)).astype(np.int32)
    pay_null = (rng.random(len(pk)) < 0.1).astype(np.uint8)
    lit = 120_000
    b_col, k1 = device_column(bk, np.zeros(len(bk), np.uint8), chunk)
    p_col, k2 = device_column(pk, np.zeros(len(pk), np.uint8), chunk)
    pay_col, k3 = device_column(pay, pay_null, chunk)
    n_chunks = (len(pk) + chunk - 1) // chunk

    ctx.scan(p_col, abi.PRED_GREATER_THAN_EQUALS, lit, slot=1)
    ptr, _, n_scan, nc = ctx.scan_result_device(1)
    assert nc == n_chunks
    begin = ctx.scan_chunk_begin(1)   # chunk boundaries (host metadata of the reference table's chunks), no PosList download

    class W:
        __cuda_array_interface__ = {"shape": (max(n_scan, 1), 2), "typestr": "<i4", "data": (ptr, False), "version": 2}
    view = torch.as_tensor(W(), device=dev)[:n_scan]
    rb = hj.RelationBuilder(abi.LOC_DEVICE)
    for c in range(n_chunks):
        if begin[c + 1] > begin[c]:
            rb.add_reference(c, view[int(begin[c]):int(begin[c + 1])], p_col, abi.KEY_INT32)
    filtered = rb.build()

    info = ctx.join(b_col, filtered, abi.MODE_INNER)
    _, p_ptr, n = ctx.result_device()
    hit = (pk >= lit) & (pk <= 300_000)
    assert n == info.n_rows == int(hit.sum())
    ctx.translate_device(p_ptr, n, filtered)           # rows of `filtered` -> rows of the probe table, in place
    values_ptr, nulls_ptr, kt = ctx.gather_column(p_ptr, n, pay_col)
    got = ctx.aggregate_column(values_ptr, nulls_ptr, n, kt)
    close_enough(got, ao.aggregate_column(pay[hit], pay_null[hit]))
    # and the translated PosList itself: exactly the probe-table rows that qualify
    _, gp = ctx.fetch()
    rows = np.sort(gp[:, 0].astype(np.int64) * chunk + gp[:, 1])
    assert np.array_equal(rows, np.flatnonzero(hit))


# ---- GROUP BY one integer column: hj_aggregate_grouped --------------------------------------------------------------------
def grouped_on_gpu(ctx, gk, gz, v, vz, max_groups=0):
    import torch
    n = len(gk)
    tg, tgz = torch.from_numpy(gk).cuda(), torch.from_numpy(gz).cuda()
    tv = None if v is None else torch.from_numpy(v).cuda()
    tvz = None if v is None else torch.from_numpy(vz).cuda()
    return ctx.aggregate_grouped(tg.data_ptr() if n else None, tgz.data_ptr() if n and gz.any() else None, KEY_TYPES[gk.dtype.name],
                                 None if (v is None or n == 0) else tv.data_ptr(), None if (v is None or n == 0 or not vz.any()) else tvz.data_ptr(),
                                 abi.KEY_INT32 if v is None else KEY_TYPES[v.dtype.name], n, max_groups)


def grouped_want(gk, gz, v, vz):
    """{group key or None: aggregates} with ONE oracle call per group (aggregate_grouped, which the CPU tests pin on the
    reference's tables, makes one call per aggregate: too slow for 100 000 groups)."""
    members = {}
    for i, (k, z) in enumerate(zip(gk.tolist(), gz.tolist())):
        members.setdefault(None if z else k, []).append(i)
    want = {}
    for key, idx in members.items():
        idx = np.array(idx, dtype=np.int64)
        want[key] = {"count_rows": len(idx)} if v is None else ao.aggregate_column(v[idx], vz[idx])
    return want


def groups_close(got, want):
    assert set(got) == set(want), (sorted(map(str, got))[:10], sorted(map(str, want))[:10])
    for key, w in want.items():
        g = got[key]
        for k, x in w.items():
            if x is None or isinstance(x, int):
                assert g[k] == x, (key, k, g[k], x)
            else:
                assert g[k] == pytest.approx(x, rel=1e-9, abs=1e-9), (key, k, g[k], x)


ONE_COLUMN_CASES = [c for c in FIX["grouped_cases"] if len(c["group_by"]) == 1 and len({a[1] for a in c["aggregates"] if a[1]}) <= 2]


@pytest.mark.parametrize("case", ONE_COLUMN_CASES, ids=[c["name"] for c in ONE_COLUMN_CASES])
def test_reference_group_by_case_on_gpu(ctx, case):
    """The reference's one-group-column cases: every aggregate column of the case is one hj_aggregate_grouped call; the
    expected table's rows are found by group key."""
    table = FIX["grouped_tables"][case["table"]]
    gk, gz = am.column_of(table, case["group_by"][0])
    expected = {row[0]: row[1:] for row in case["expected_rows"]}
    if not case["aggregates"]:
        got = grouped_on_gpu(ctx, gk, gz, None, None)
        assert set(got) == set(expected)
        return
    for j, (fn, name) in enumerate(case["aggregates"]):
        if name is None:
            got = grouped_on_gpu(ctx, gk, gz, None, None)
            for key, row in expected.items():
                assert got[key]["count_rows"] == row[j], (key, got[key], row)
            continue
        v, vz = am.column_of(table, name)
        got = grouped_on_gpu(ctx, gk, gz, v, vz)
        assert set(got) == set(expected)
        for key, row in expected.items():
            assert am.matches_reference(got[key][fn], row[j]), (key, fn, got[key], row, case["cite"])


@pytest.mark.parametrize("gdtype", [np.int32, np.int64])
@pytest.mark.parametrize("vdtype", [None, np.int32, np.int64, np.float32, np.float64])
@pytest.mark.parametrize("n_groups", [1, 7, 200, 5000, 100_000])
def test_random_group_by_matches_oracle(ctx, gdtype, vdtype, n_groups):
    """From one group (every row updates the same shared-memory slot) over more groups than the CTA tables hold (the global
    table is hit directly) to nearly one group per row; NULL group keys, NULL values, the key that is the table's EMPTY
    marker, max_groups hints that are exact, generous and absent."""
    rng = np.random.default_rng(n_groups)
    n = 150_000
    gk = rng.integers(-n_groups // 2, n_groups - n_groups // 2, n).astype(gdtype)
    if gdtype == np.int64:
        gk[rng.integers(0, n, 5)] = np.iinfo(np.int64).min   # = the EMPTY marker of the table
        gk[rng.integers(0, n, 5)] = np.iinfo(np.int64).max
    gz = (rng.random(n) < 0.05).astype(np.uint8)
    if vdtype is None:
        v = vz = None
    else:
        v = (rng.integers(-1000, 1000, n) / (4 if np.dtype(vdtype).kind == "f" else 1)).astype(vdtype)
        vz = (rng.random(n) < 0.1).astype(np.uint8)
    want = grouped_want(gk, gz, v, vz)
    for hint in (0, len(want), 4 * len(want)):
        got = grouped_on_gpu(ctx, gk, gz, v, vz, hint)
        if v is None:
            assert {k: g["count_rows"] for k, g in got.items()} == {k: w["count_rows"] for k, w in want.items()}
        else:
            groups_close(got, want)


def test_group_table_overflow_is_an_error_not_a_hang(ctx):
    gk = np.arange(100_000, dtype=np.int32)
    with pytest.raises(hj.HjError) as e:
        grouped_on_gpu(ctx, gk, np.zeros(len(gk), np.uint8), None, None, max_groups=100)   # 100 000 groups into 1024 slots
    assert e.value.status == abi.ERR_CAPACITY


@pytest.mark.parametrize("case", ONE_COLUMN_CASES, ids=[c["name"] for c in ONE_COLUMN_CASES])
@pytest.mark.parametrize("encoding", ["value", "dict"])
def test_reference_group_by_case_through_chunk_descriptors(ctx, case, encoding):
    """The same cases with host-resident chunked tables (chunk size 2 like the reference's fixtures), value and dictionary
    chunks: hj_aggregate_grouped_relation, the call the C++ Aggregate makes."""
    table = FIX["grouped_tables"][case["table"]]
    gk, gz = am.column_of(table, case["group_by"][0])
    group_rel = host_relation(gk, gz, case["chunk_size"], encoding)
    expected = {row[0]: row[1:] for row in case["expected_rows"]}
    if not case["aggregates"]:
        assert set(ctx.aggregate_grouped_relation(group_rel)) == set(expected)
        return
    for j, (fn, name) in enumerate(case["aggregates"]):
        if name is None:
            got = ctx.aggregate_grouped_relation(group_rel)
            assert {k: g["count_rows"] for k, g in got.items()} == {k: row[j] for k, row in expected.items()}
            continue
        v, vz = am.column_of(table, name)
        got = ctx.aggregate_grouped_relation(group_rel, host_relation(v, vz, case["chunk_size"], encoding))
        assert set(got) == set(expected)
        for key, row in expected.items():
            assert am.matches_reference(got[key][fn], row[j]), (key, fn, got[key], row, case["cite"])


def test_random_grouped_relations_match_oracle(ctx):
    rng = np.random.default_rng(77)
    n, chunk = 300_000, 1 << 16
    gk = rng.integers(0, 3000, n).astype(np.int32)
    gz = (rng.random(n) < 0.03).astype(np.uint8)
    v = (rng.integers(-500, 500, n) / 2).astype(np.float64)
    vz = (rng.random(n) < 0.1).astype(np.uint8)
    want = grouped_want(gk, gz, v, vz)
    for encoding in ("value", "dict"):
        got = ctx.aggregate_grouped_relation(host_relation(gk, gz, chunk, encoding), host_relation(v, vz, chunk, encoding), max_groups=4000)
        groups_close(got, want)


# ---- GROUP BY several columns: hj_aggregate_grouped_rows (tables keyed by a representative row) ---------------------------
def _as_stored(key, group_cols):
    """A fixture's group key as the column stores it: the .tbl text 350.7 of a float column is the float32 350.70001220703125
    the device (and the reference's ValueColumn<float>) groups by."""
    return tuple(None if x is None else gv.dtype.type(x).item() for x, (gv, _) in zip(key, group_cols))


def grouped_rows_on_gpu(ctx, group_cols, v, vz, max_groups=0):
    """group_cols: [(values, is_null)]; returns {tuple of group values (None = NULL): aggregate dict}, the group values read
    from the host columns at the representative row the library reports."""
    import torch
    n = len(group_cols[0][0])
    keep, cols = [], []
    for gv, gz in group_cols:
        tv, tz = torch.from_numpy(gv).cuda(), torch.from_numpy(gz).cuda()
        keep += [tv, tz]
        cols.append((tv.data_ptr() if n else None, tz.data_ptr() if n and gz.any() else None, KEY_TYPES[gv.dtype.name]))
    tvv = None if v is None else torch.from_numpy(v).cuda()
    tvz = None if v is None else torch.from_numpy(vz).cuda()
    got = ctx.aggregate_grouped_rows(cols, None if (v is None or n == 0) else tvv.data_ptr(),
                                     None if (v is None or n == 0 or not vz.any()) else tvz.data_ptr(),
                                     abi.KEY_INT32 if v is None else KEY_TYPES[v.dtype.name], n, max_groups)
    out = {}
    for row, agg in got.items():
        key = tuple(None if gz[row] else gv[row].item() for gv, gz in group_cols)
        assert key not in out, "two representative rows for one group"
        out[key] = agg
    return out


@pytest.mark.parametrize("case", FIX["grouped_cases"], ids=[c["name"] for c in FIX["grouped_cases"]])
def test_reference_group_by_case_on_gpu_by_representative_rows(ctx, case):
    """All 39 numeric group-by cases of the reference (1, 2 and 3 group-by columns, NULL group keys, COUNT(*))."""
    table = FIX["grouped_tables"][case["table"]]
    group_cols = [am.column_of(table, name) for name in case["group_by"]]
    k = len(group_cols)
    expected = {_as_stored(row[:k], group_cols): row[k:] for row in case["expected_rows"]}
    if not case["aggregates"]:
        assert set(grouped_rows_on_gpu(ctx, group_cols, None, None)) == set(expected)
        return
    for j, (fn, name) in enumerate(case["aggregates"]):
        if name is None:
            got = grouped_rows_on_gpu(ctx, group_cols, None, None)
            assert {key: g["count_rows"] for key, g in got.items()} == {key: row[j] for key, row in expected.items()}
            continue
        v, vz = am.column_of(table, name)
        got = grouped_rows_on_gpu(ctx, group_cols, v, vz)
        assert set(got) == set(expected)
        for key, row in expected.items():
            assert am.matches_reference(got[key][fn], row[j]), (key, fn, got[key], row, case["cite"])


@pytest.mark.parametrize("n_cols", [1, 2, 4])
@pytest.mark.parametrize("cardinality", [2, 40, 3000])
def test_random_multi_column_group_by_matches_oracle(ctx, n_cols, cardinality):
    """Mixed column types (int32, int64, float32, float64 group-by columns), NULLs in every column, from a handful of groups
    to ~cardinality^n_cols (far more than the shared-memory tables hold)."""
    rng = np.random.default_rng(n_cols * 1000 + cardinality)
    n = 120_000
    dtypes = [np.int32, np.int64, np.float32, np.float64][:n_cols]
    group_cols = []
    for dt in dtypes:
        gv = rng.integers(-cardinality // 2, cardinality - cardinality // 2, n).astype(dt)
        group_cols.append((gv, (rng.random(n) < 0.05).astype(np.uint8)))
    v = rng.integers(-1000, 1000, n).astype(np.int64)
    vz = (rng.random(n) < 0.1).astype(np.uint8)
    members = {}
    for i in range(n):
        members.setdefault(tuple(None if gz[i] else gv[i].item() for gv, gz in group_cols), []).append(i)
    want = {key: ao.aggregate_column(v[np.array(idx)], vz[np.array(idx)]) for key, idx in members.items()}
    got = grouped_rows_on_gpu(ctx, group_cols, v, vz)
    groups_close(got, want)


@pytest.mark.parametrize("case", FIX["grouped_cases"], ids=[c["name"] for c in FIX["grouped_cases"]])
@pytest.mark.parametrize("encoding", ["value", "dict"])
def test_reference_group_by_case_by_representative_rows_through_chunk_descriptors(ctx, case, encoding):
    """All 39 cases with host-resident chunked tables (chunk size 2): hj_aggregate_grouped_rows_relation, the call the C++
    Aggregate makes; a group's values are read from the host columns at its representative row number."""
    table = FIX["grouped_tables"][case["table"]]
    group_cols = [am.column_of(table, name) for name in case["group_by"]]
    rels = [host_relation(gv, gz, case["chunk_size"], encoding) for gv, gz in group_cols]
    k = len(group_cols)
    expected = {_as_stored(row[:k], group_cols): row[k:] for row in case["expected_rows"]}

    def keyed(got):
        return {tuple(None if gz[row] else gv[row].item() for gv, gz in group_cols): agg for row, agg in got.items()}

    if not case["aggregates"]:
        assert set(keyed(ctx.aggregate_grouped_rows_relation(rels))) == set(expected)
        return
    for j, (fn, name) in enumerate(case["aggregates"]):
        if name is None:
            got = keyed(ctx.aggregate_grouped_rows_relation(rels))
            assert {key: g["count_rows"] for key, g in got.items()} == {key: row[j] for key, row in expected.items()}
            continue
        v, vz = am.column_of(table, name)
        got = keyed(ctx.aggregate_grouped_rows_relation(rels, host_relation(v, vz, case["chunk_size"], encoding)))
        assert set(got) == set(expected)
        for key, row in expected.items():
            assert am.matches_reference(got[key][fn], row[j]), (key, fn, got[key], row, case["cite"])


# ------------------------------------------------------------------------------------------------------------------
# MIN / MAX and NaN, -0.0 / +0.0 group keys, the growing group table
# ------------------------------------------------------------------------------------------------------------------
def same_or_both_nan(a, b):
    return a == b or (isinstance(a, float) and isinstance(b, float) and np.isnan(a) and np.isnan(b))


NAN_COLUMNS = {
    # the reference seeds MIN / MAX with the first non-NULL value and replaces on plain < / > (aggregate.cpp:141-165): a NaN
    # that comes first stays, a NaN anywhere else replaces nothing; the oracle walks the rows like the reference
    "nan_first": ([np.nan, 3.0, -2.0, 8.5], [0, 0, 0, 0]),
    "nan_first_after_nulls": ([7.0, np.nan, 3.0, -2.0], [1, 0, 0, 0]),
    "nan_in_the_middle": ([3.0, np.nan, -2.0, 8.5], [0, 0, 0, 0]),
    "nan_last": ([3.0, -2.0, 8.5, np.nan], [0, 0, 0, 0]),
    "null_nan_first_is_skipped": ([np.nan, 3.0, -2.0, np.nan], [1, 0, 0, 0]),
    "only_nans": ([np.nan, np.nan], [0, 0]),
    "negative_nan_first": ([-np.nan, 1.0, 2.0], [0, 0, 0]),
}


@pytest.mark.parametrize("name", list(NAN_COLUMNS))
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("tail", [0, 70_001])
def test_min_max_keep_a_nan_only_when_it_comes_first(ctx, name, dtype, tail):
    """Ungrouped, all three entry points (device arrays, chunk descriptors, the no-allocation one); `tail` appends ordinary
    rows so that the first value and the rest are seen by different CTAs."""
    import ctypes as C
    import torch
    values, nulls = NAN_COLUMNS[name]
    rng = np.random.default_rng(len(name) + tail)
    v = np.concatenate([np.array(values, dtype=dtype), (rng.standard_normal(tail) * 100).astype(dtype)])
    z = np.concatenate([np.array(nulls, dtype=np.uint8), (rng.random(tail) < 0.2).astype(np.uint8)])
    want = ao.aggregate_column(v, z)
    assert np.isnan(want["min"]) == name.replace("negative_", "").startswith(("nan_first", "only_nans")), "the oracle follows the reference"
    tv, tz = torch.from_numpy(v).cuda(), torch.from_numpy(z).cuda()
    kt = KEY_TYPES[np.dtype(dtype).name]
    out = torch.empty(C.sizeof(abi.Aggregate), dtype=torch.uint8, device="cuda")
    ctx.aggregate_column_static(tv.data_ptr(), tz.data_ptr(), len(v), kt, out.data_ptr())
    torch.cuda.synchronize()
    rel, keep = device_column(v, z, 4096)
    for how, got in (("arrays", ctx.aggregate_column(tv.data_ptr(), tz.data_ptr(), len(v), kt)),
                     ("static", hj.HashJoin.aggregate_record(out.cpu().numpy().tobytes())),
                     ("chunks", ctx.aggregate_relation(rel))):
        for k in ("min", "max"):
            assert same_or_both_nan(got[k], want[k]), (how, k, got[k], want[k])
        assert got["count"] == want["count"], how
    del keep


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_grouped_min_max_keep_a_nan_only_when_it_comes_first_in_the_group(ctx, dtype):
    """One group per NAN_COLUMNS entry (+ ordinary groups around them), rows of the groups interleaved; through the
    one-integer-column table, the representative-row table and the chunk-descriptor entry point."""
    rng = np.random.default_rng(77)
    names = list(NAN_COLUMNS)
    longest = max(len(NAN_COLUMNS[n][0]) for n in names)
    gk, v, z = [], [], []
    for i in range(longest):               # row i of every group, group after group: a group's rows keep their order
        for g, n in enumerate(names):
            values, nulls = NAN_COLUMNS[n]
            if i < len(values):
                gk.append(g), v.append(values[i]), z.append(nulls[i])
    filler = 50_000
    gk = np.concatenate([np.array(gk, np.int32), rng.integers(100, 400, filler).astype(np.int32)])
    v = np.concatenate([np.array(v, dtype), (rng.standard_normal(filler) * 10).astype(dtype)])
    z = np.concatenate([np.array(z, np.uint8), (rng.random(filler) < 0.1).astype(np.uint8)])
    gz = np.zeros(len(gk), np.uint8)
    want = grouped_want(gk, gz, v, z)
    assert any(w["min"] is not None and np.isnan(w["min"]) for w in want.values())
    rel_g, keep_g = device_column(gk, gz, 8192)
    rel_v, keep_v = device_column(v, z, 8192)
    by_rows = grouped_rows_on_gpu(ctx, [(gk, gz)], v, z)
    for how, got in (("one column", grouped_on_gpu(ctx, gk, gz, v, z)),
                     ("representative rows", {k[0]: a for k, a in by_rows.items()}),
                     ("chunks", ctx.aggregate_grouped_relation(rel_g, rel_v))):
        assert set(got) == set(want), how
        for key, w in want.items():
            for k in ("min", "max"):
                assert same_or_both_nan(got[key][k], w[k]), (how, key, k, got[key][k], w[k])
            assert got[key]["count"] == w["count"], (how, key)
    del keep_g, keep_v


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_minus_zero_and_plus_zero_are_one_group(ctx, dtype):
    """The reference maps group-by values to ids through a map that compares with == (aggregate.cpp:254-300): -0.0 and +0.0
    are one group. The device compares bit patterns and therefore canonicalises the zero first."""
    rng = np.random.default_rng(3)
    n = 40_000
    gv = rng.integers(-3, 4, n).astype(dtype)
    gv[gv == 0] = np.where(rng.random((gv == 0).sum()) < 0.5, dtype(-0.0), dtype(0.0))
    assert np.signbit(gv[gv == 0]).any() and not np.signbit(gv[gv == 0]).all()
    other = rng.integers(0, 3, n).astype(np.int32)
    gz = np.zeros(n, np.uint8)
    v = rng.integers(-50, 50, n).astype(np.int64)
    for cols in ([(gv, gz)], [(other, gz), (gv, gz)]):
        got = grouped_rows_on_gpu(ctx, cols, v, gz)   # keys are read from the representative rows: -0.0 == 0.0 == 0 in a dict
        members = {}
        for i in range(n):
            members.setdefault(tuple(c[0][i].item() + 0.0 for c in cols), []).append(i)   # x + 0.0: -0.0 -> 0.0
        assert len(got) == len(members) == (7 if len(cols) == 1 else 21)
        for key, idx in members.items():
            hit = [a for k, a in got.items() if all(float(x) == y for x, y in zip(k, key))]
            assert len(hit) == 1 and hit[0]["count_rows"] == len(idx) and hit[0]["sum"] == int(v[idx].sum()), key


def test_group_table_grows_when_no_bound_is_given(ctx):
    """Without max_groups the table starts at 2^20 slots (64 MB) instead of 2 x n_rows; 1.2 M distinct keys do not fit,
    the library notices (bounded probing), rebuilds the table 8 x larger and repeats the pass (group_run, hj_api.cu).
    Checked vectorised on the fetched records: every key once, COUNT(*) = 1, SUM = the row's value."""
    import ctypes as C
    import torch
    n = 1_200_000
    rng = np.random.default_rng(5)
    gk = (rng.permutation(n).astype(np.int64) - n // 2) * 7
    v = rng.integers(-10**6, 10**6, n).astype(np.int64)
    tg, tv = torch.from_numpy(gk).cuda(), torch.from_numpy(v).cuda()
    for repeat in range(2):   # the second call starts small again and must grow again
        info = abi.GroupInfo()
        ctx._check(ctx._l.hj_aggregate_grouped(ctx._ctx, tg.data_ptr(), None, abi.KEY_INT64, tv.data_ptr(), None, abi.KEY_INT64, n, 0,
                                               None, C.byref(info)))
        assert info.n_groups == n and info.n_rows == n
        groups = (abi.Group * n)()
        ctx._check(ctx._l.hj_aggregate_grouped_fetch(ctx._ctx, groups, n))
        rec = np.ctypeslib.as_array(groups)
        order = np.argsort(rec["key"], kind="stable")
        by_key = np.argsort(gk, kind="stable")
        assert (rec["key"][order] == gk[by_key]).all()
        assert (rec["agg"]["n_rows"] == 1).all() and (rec["agg"]["n_values"] == 1).all()
        assert (rec["agg"]["sum_int"][order] == v[by_key]).all()
        assert (rec["agg"]["min_int"][order] == v[by_key]).all() and (rec["agg"]["max_int"][order] == v[by_key]).all()
    # a small GROUP BY right after it on the same context
    small = grouped_on_gpu(ctx, np.array([1, 2, 1], np.int32), np.zeros(3, np.uint8), np.array([5, 6, 7], np.int32), np.zeros(3, np.uint8))
    assert small[1]["sum"] == 12 and small[2]["sum"] == 6
