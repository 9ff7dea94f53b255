"""Device TableScan (hj_scan) against the oracle and the reference's own expectations, through the C ABI.

First GPU run: profiles/r01_gpu_scan_first_run.txt (312 passed: the reference's cases x encodings, the randomized
differential scans, int64, slots); the float / string / FOR / RLE scans, hj_gather_column and scan -> join with the PosLists
staying in HBM first ran in round 2 (profiles/r02_gpu_call1_gated_suites.txt).
Everything compared against (oracle/table_scan_oracle.py, tests/golden/scan_fixtures.json) is pinned on CPU by
tests/test_scan_oracle.py."""
import importlib
import os
import sys

import numpy as np
import pytest

import scan_model as sm
from oracle import table_scan_oracle as tso

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
pytestmark = pytest.mark.gpu

FIX = sm.load_fixtures()
CASES = [(c, e) for c in FIX["cases"] for e in sm.encodings_for(c)]
IDS = [f"{c['name']}-{e if isinstance(e, str) else 'chunks' + ''.join(map(str, e))}" for c, e in CASES]


@pytest.fixture(scope="module")
def ctx():
    j = hj.HashJoin(device=0)
    yield j
    j.close()


def key_type_of(chunk):
    a = chunk[{"value": "values", "dict": "dictionary", "for": "minima", "rle": "run_values"}[chunk["kind"]]]
    return {"int32": abi.KEY_INT32, "int64": abi.KEY_INT64, "float32": abi.KEY_FLOAT, "float64": abi.KEY_DOUBLE}[a.dtype.name]


def relation_of(chunks, ids, location=abi.LOC_HOST):
    rb = hj.RelationBuilder(location)
    for chunk, cid in zip(chunks, ids):
        if chunk["kind"] == "value":
            rb.add_value(cid, chunk["values"], chunk.get("nulls"))
        elif chunk["kind"] == "dict":
            rb.add_dict(cid, chunk["dictionary"], chunk["attr"], chunk["null_value_id"])
        elif chunk["kind"] == "for":
            rb.add_for(cid, chunk["minima"], chunk["offsets"], chunk.get("nulls"))
        elif chunk["kind"] == "rle":
            rb.add_rle(cid, chunk["n_rows"], chunk["run_values"], chunk["end_positions"], chunk["run_nulls"])
        else:
            referenced = relation_of(chunk["referenced"], list(range(len(chunk["referenced"]))), location)
            kt = max(key_type_of(c) for c in chunk["referenced"])
            rb.add_reference(cid, np.ascontiguousarray(chunk["pos_list"]), referenced, kt)
    return rb.build()


def make_gpu_scan(j, check_against_oracle=True):
    def scan(chunks, ids, pred, value):
        rel = relation_of(chunks, ids)
        info = j.scan(rel, pred, value)
        pos, begin = j.scan_fetch()
        assert info.n_rows == len(pos) and info.n_chunks == len(chunks) and begin[-1] == len(pos)
        if check_against_oracle:
            want_pos, want_begin = tso.table_scan(chunks, ids, pred, value)
            assert np.array_equal(begin, want_begin), "per-chunk match counts differ from the oracle"
            assert np.array_equal(pos, want_pos), "PosLists differ from the oracle"
            assert info.n_nonempty_chunks == int((np.diff(want_begin.astype(np.int64)) > 0).sum())
        return pos, begin
    return scan


@pytest.mark.parametrize("case,encode", CASES, ids=IDS)
def test_reference_scan_case_on_gpu(ctx, case, encode):
    if case.get("expect_error"):
        with pytest.raises(hj.HjError) as e:
            sm.run_case(case, FIX["tables"], make_gpu_scan(ctx), encode)
        assert e.value.status == abi.ERR_INVALID
        return
    assert sm.run_case(case, FIX["tables"], make_gpu_scan(ctx), encode) == sm.expected_of(case), case["cite"]


@pytest.mark.parametrize("style", ["for", "rle"])
@pytest.mark.parametrize("case,encode", [ce for ce in CASES if ce[1] != "none" and not ce[0].get("expect_error")],
                         ids=[i for i, ce in zip(IDS, CASES) if ce[1] != "none" and not ce[0].get("expect_error")])
def test_reference_scan_case_on_gpu_for_and_rle(ctx, case, encode, style):
    """The reference's FrameOfReference / RunLength parametrisations: decoded on the device by k_gather_refs, then scanned."""
    assert sm.run_case(case, FIX["tables"], make_gpu_scan(ctx), encode, style) == sm.expected_of(case), case["cite"]


def random_chunks(rng, n_rows, chunk, dtype, encoding, nullable, domain):
    values = rng.integers(-domain, domain, n_rows).astype(dtype)
    nulls = (rng.random(n_rows) < 0.1).astype(np.uint8) if nullable else np.zeros(n_rows, np.uint8)
    chunks = []
    for c, lo in enumerate(range(0, max(n_rows, 1), chunk)):
        v, nl = np.ascontiguousarray(values[lo:lo + chunk]), np.ascontiguousarray(nulls[lo:lo + chunk])
        enc = encoding if encoding != "mixed" else ("dict" if c % 2 else "value")
        chunks.append(sm.dict_encode(v, nl) if enc == "dict" else {"kind": "value", "values": v, "nulls": nl if nullable else None})
    return chunks


@pytest.mark.parametrize("dtype", [np.int32, np.int64])
@pytest.mark.parametrize("encoding", ["value", "dict", "mixed"])
@pytest.mark.parametrize("nullable", [False, True])
def test_random_scans_match_oracle(ctx, dtype, encoding, nullable):
    """Edge sizes (empty, one row, around a warp, around a 4096-row tile, ragged last chunk), u8/u16/u32 attribute vectors
    (domain), literals below / inside / above the value range, every predicate; bit-exact PosLists."""
    rng = np.random.default_rng(0xB2005)
    scan = make_gpu_scan(ctx)
    for n_rows, chunk, domain in [(0, 8, 5), (1, 8, 5), (31, 64, 5), (33, 32, 100), (4095, 4096, 100), (4097, 4096, 40000),
                                  (20000, 7000, 100), (300000, 131072, 100000)]:
        chunks = random_chunks(rng, n_rows, chunk, dtype, encoding, nullable, domain)
        ids = list(range(len(chunks)))
        for pred in list(range(6)) + [10, 11]:
            for lit in (-domain - 3, -1, 0, domain // 2, domain + 3):
                scan(chunks, ids, pred, lit)
        scan(chunks, ids, 0, None)  # NULL literal: nothing


def test_int64_literals_beyond_32_bits_and_int32_range_check(ctx):
    rng = np.random.default_rng(5)
    v = (rng.integers(-4, 4, 5000).astype(np.int64) << 33) + rng.integers(0, 3, 5000)
    chunks = [{"kind": "value", "values": v[:3000].copy(), "nulls": None}, sm.dict_encode(v[3000:].copy(), np.zeros(2000, np.uint8))]
    scan = make_gpu_scan(ctx)
    for pred in range(6):
        for lit in (int(v[17]), (1 << 33) + 1, -(1 << 40), 1 << 40):
            scan(chunks, [0, 1], pred, lit)
    small = [{"kind": "value", "values": np.arange(10, dtype=np.int32), "nulls": None}]
    with pytest.raises(hj.HjError) as e:  # the reference's type_cast<int32_t> fails on such a literal
        ctx.scan(relation_of(small, [0]), abi.PRED_LESS_THAN, 1 << 35)
    assert e.value.status == abi.ERR_INVALID


def test_irregular_chunk_ids_and_slots(ctx):
    rng = np.random.default_rng(9)
    chunks = random_chunks(rng, 10000, 3000, np.int32, "mixed", True, 50)
    ids = [7, 3, 100, 42]
    want_a, _ = tso.table_scan(chunks, ids, tso.GREATER_THAN, 10)
    want_b, _ = tso.table_scan(chunks, ids, tso.IS_NULL, None)
    rel = relation_of(chunks, ids)
    ctx.scan(rel, abi.PRED_GREATER_THAN, 10, slot=1)
    ctx.scan(rel, abi.PRED_IS_NULL, None, slot=2)  # IS NULL ignores the literal
    assert np.array_equal(ctx.scan_fetch(1)[0], want_a) and np.array_equal(ctx.scan_fetch(2)[0], want_b)
    ctx.scan_release(1)
    with pytest.raises(hj.HjError):
        ctx.scan_fetch(1)
    assert np.array_equal(ctx.scan_fetch(2)[0], want_b)


def test_scan_then_join_without_leaving_the_device(ctx):
    """The point of N4: both join inputs are filtered by device scans; their PosLists never leave HBM and feed hj_join as
    REFERENCE segments of device-resident relations. Checked against the oracle join of the host-filtered inputs."""
    import torch
    import oracle_binding as ob
    import workloads as wl
    rng = np.random.default_rng(11)
    chunk = 1 << 16
    bk = rng.permutation(400_000).astype(np.int32) + 1
    pk = rng.integers(1, 400_001, 1_500_000).astype(np.int32)
    dev = torch.device("cuda", 0)

    def device_column(keys):
        rb, parts = hj.RelationBuilder(abi.LOC_DEVICE), []
        for c, lo in enumerate(range(0, len(keys), chunk)):
            t = torch.from_numpy(keys[lo:lo + chunk].copy()).to(dev)
            parts.append(t)
            rb.add_value(c, t)
        return rb.build(), parts

    def device_view(ptr, n):
        class W:
            __cuda_array_interface__ = {"shape": (max(n, 1), 2), "typestr": "<i4", "data": (ptr, False), "version": 2}
        return torch.as_tensor(W(), device=dev)[:n]

    def filtered_relation(col_rel, slot, n_chunks):
        ptr, bptr, n, nc = ctx.scan_result_device(slot)
        assert nc == n_chunks
        pos, begin = ctx.scan_fetch(slot)
        view = device_view(ptr, n)
        rb = hj.RelationBuilder(abi.LOC_DEVICE)
        for c in range(n_chunks):
            if begin[c + 1] > begin[c]:
                rb.add_reference(c, view[int(begin[c]):int(begin[c + 1])], col_rel, abi.KEY_INT32)
        return rb.build(), pos, view

    b_col, keep_b = device_column(bk)
    p_col, keep_p = device_column(pk)
    ctx.scan(b_col, abi.PRED_LESS_THAN, 300_000, slot=0)
    ctx.scan(p_col, abi.PRED_GREATER_THAN_EQUALS, 100_000, slot=1)
    b_rel, b_pos, kb = filtered_relation(b_col, 0, len(keep_b))
    p_rel, p_pos, kp = filtered_relation(p_col, 1, len(keep_p))
    info = ctx.join(b_rel, p_rel, abi.MODE_INNER)
    gb, gp = ctx.fetch()
    # oracle: the same reference tables on the host
    # (the oracle restates JoinHash, not ReferenceColumn: it gets the rows a ReferenceColumnIterable would yield, gathered here)
    def host_rel(keys, pos):
        rb = hj.RelationBuilder(abi.LOC_HOST)
        for c in np.unique(pos[:, 0]):
            rows = pos[pos[:, 0] == c]
            gathered = np.ascontiguousarray(keys[rows[:, 0].astype(np.int64) * chunk + rows[:, 1]])
            rb.add_value(int(c), gathered, None, kind=abi.SEG_GATHERED)
        return rb.build()
    ob_b, ob_p, oinfo = ob.oracle_join(host_rel(bk, b_pos), host_rel(pk, p_pos), abi.MODE_INNER)
    assert info.n_rows == oinfo["n_rows"] == int(((pk >= 100_000) & (pk < 300_000)).sum())
    assert np.array_equal(ob.canonical(gb, gp), ob.canonical(ob_b, ob_p))


# ---- float / double columns (hj_scan compares the column type's values; HJ_SCAN_VALUE_IS_DOUBLE literals) ---------------
@pytest.mark.parametrize("case", FIX["float_cases"], ids=[c["name"] for c in FIX["float_cases"]])
@pytest.mark.parametrize("encode,style", [("none", "dict"), ("all", "dict"), ("all", "rle")])
def test_reference_float_scan_case_on_gpu(ctx, case, encode, style):
    assert sm.run_case(case, FIX["typed_tables"], make_gpu_scan(ctx), encode, style) == sm.expected_of(case), case["cite"]


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("encoding", ["value", "dict", "mixed"])
@pytest.mark.parametrize("nullable", [False, True])
def test_random_float_scans_match_oracle(ctx, dtype, encoding, nullable):
    from test_scan_oracle import FLOAT_LITERALS, float_chunks
    rng = np.random.default_rng(0xF10A7)
    scan = make_gpu_scan(ctx)
    for n_rows, chunk in [(0, 8), (1, 8), (33, 32), (4097, 4096), (20000, 7000), (300000, 131072)]:
        chunks = float_chunks(rng, n_rows, chunk, dtype, encoding, nullable)
        ids = list(range(len(chunks)))
        for pred in list(range(6)) + [10, 11]:
            for lit in FLOAT_LITERALS:
                scan(chunks, ids, pred, lit)
        scan(chunks, ids, 0, None)


def test_floating_literal_on_an_integer_column(ctx):
    chunks = [{"kind": "value", "values": np.array([455, 456, 457, 458, -1, 0], np.int32), "nulls": None},
              sm.dict_encode(np.array([455, 456, 457, 458, -1, 0], np.int32), np.zeros(6, np.uint8))]
    scan = make_gpu_scan(ctx)
    for pred in range(6):
        for lit in (457.9, -0.5, 0.0, -1.5, 2147483647.5):
            scan(chunks, [0, 1], pred, lit)
    with pytest.raises(hj.HjError) as e:  # boost::numeric_cast<int32_t>(3e9) throws
        ctx.scan(relation_of(chunks, [0, 1]), abi.PRED_EQUALS, 3.0e9)
    assert e.value.status == abi.ERR_INVALID


# ---- string (or any) dictionary columns scanned on value ids: hj_scan_value_ids ---------------------------------------
def make_gpu_value_id_scan(j):
    """scan_fn for scan_model.run_case: all chunks must be dictionary chunks; the dictionary search happens here on the
    host (test_scan_oracle.value_id_bounds), the attribute vectors are scanned on the device."""
    from test_scan_oracle import value_id_bounds

    def scan(chunks, ids, pred, value):
        assert all(c["kind"] == "dict" for c in chunks)
        rb = hj.RelationBuilder(abi.LOC_HOST)
        for c, cid in zip(chunks, ids):
            rb.add_value_ids(cid, c["attr"], c["null_value_id"])
        literal = tso.cast_literal(value, chunks[0]["dictionary"].dtype) if value is not None and chunks else value
        bounds = value_id_bounds(chunks, literal) if pred < 6 else None
        info = j.scan_value_ids(rb.build(), pred, bounds)
        pos, begin = j.scan_fetch()
        want_pos, want_begin = tso.table_scan(chunks, ids, pred, value)
        assert info.n_rows == len(pos) and np.array_equal(begin, want_begin) and np.array_equal(pos, want_pos)
        return pos, begin
    return scan


@pytest.mark.parametrize("case", FIX["string_cases"], ids=[c["name"] for c in FIX["string_cases"]])
def test_reference_string_scan_case_on_gpu(ctx, case):
    assert sm.run_case(case, FIX["typed_tables"], make_gpu_value_id_scan(ctx), "all") == sm.expected_of(case), case["cite"]


def test_random_string_dictionary_scans_match_oracle(ctx):
    from test_scan_oracle import string_chunks
    rng = np.random.default_rng(77)
    scan = make_gpu_value_id_scan(ctx)
    for n_rows, chunk, n_words in [(1, 8, 3), (33, 32, 5), (5000, 4096, 40), (70000, 30000, 300)]:
        chunks, words = string_chunks(rng, n_rows, chunk, n_words)
        ids = list(range(len(chunks)))
        for pred in list(range(6)) + [10, 11]:
            for lit in ("", str(words[len(words) // 2]), "zzzz", "Xé"):
                scan(chunks, ids, pred, lit)


def test_value_id_scan_refuses_other_chunk_kinds(ctx):
    small = [{"kind": "value", "values": np.arange(10, dtype=np.int32), "nulls": None}]
    with pytest.raises(hj.HjError) as e:
        ctx.scan_value_ids(relation_of(small, [0]), abi.PRED_EQUALS, np.zeros((1, 2), np.uint32))
    assert e.value.status == abi.ERR_INVALID


# ---- hj_gather_column: a column read through a device PosList (what the join's consumers do first) ----------------------
@pytest.mark.parametrize("encoding", ["value", "dict"])
@pytest.mark.parametrize("dtype", [np.int32, np.int64, np.float32])
def test_gather_column_at_the_join_result(ctx, dtype, encoding):
    """After an Inner join the key column of either side, read through that side's PosList, holds the SAME keys row by
    row (the join's defining property, at any size), and each equals the host-side gather of the fetched RowIDs. A second
    (payload) column of the probe table, dictionary-encoded with NULLs, is read the same way."""
    import torch
    rng = np.random.default_rng(23)
    chunk = 1 << 15
    dev = torch.device("cuda", 0)
    bk = (rng.permutation(200_000) + 1).astype(dtype)
    pk = rng.integers(1, 260_001, 700_000).astype(dtype)
    payload = rng.integers(-50, 50, len(pk)).astype(np.int32)
    payload_null = (rng.random(len(pk)) < 0.2).astype(np.uint8)
    keep = []

    def device_column(values, nulls, enc):
        rb = hj.RelationBuilder(abi.LOC_DEVICE)
        for c, lo in enumerate(range(0, len(values), chunk)):
            v, nl = values[lo:lo + chunk].copy(), nulls[lo:lo + chunk].copy()
            if enc == "dict":
                d = sm.dict_encode(v, nl)
                t = [torch.from_numpy(d["dictionary"]).to(dev), torch.from_numpy(d["attr"].astype(np.int32)).to(dev)]
                rb.add_dict(c, t[0], t[1], d["null_value_id"])
            else:
                t = [torch.from_numpy(v).to(dev), torch.from_numpy(nl).to(dev)]
                rb.add_value(c, t[0], t[1])
            keep.extend(t)
        return rb.build()

    def device_array(ptr, n, typestr):
        class W:
            __cuda_array_interface__ = {"shape": (max(n, 1),), "typestr": typestr, "data": (ptr, False), "version": 2}
        return torch.as_tensor(W(), device=dev)[:n].cpu().numpy()

    b_col = device_column(bk, np.zeros(len(bk), np.uint8), encoding)
    p_col = device_column(pk, np.zeros(len(pk), np.uint8), encoding)
    pay_col = device_column(payload, payload_null, "dict")
    info = ctx.join(b_col, p_col, abi.MODE_INNER)
    assert info.n_rows == int((pk <= 200_000).sum())
    b_ptr, p_ptr, n = ctx.result_device()
    gb, gp = ctx.fetch()
    typestr = {np.int32: "<i4", np.int64: "<i8", np.float32: "<f4"}[dtype]
    got = {}
    for name, ptr, col in (("build", b_ptr, b_col), ("probe", p_ptr, p_col)):
        v, nl, kt = ctx.gather_column(ptr, n, col)
        assert kt == key_type_of({"kind": "value", "values": bk})
        got[name] = device_array(v, n, typestr)       # copied before the next gather reuses the buffer
        assert not device_array(nl, n, "|u1").any()
    assert np.array_equal(got["build"], got["probe"])
    assert np.array_equal(got["probe"], pk[gp[:, 0].astype(np.int64) * chunk + gp[:, 1]])
    assert np.array_equal(got["build"], bk[gb[:, 0].astype(np.int64) * chunk + gb[:, 1]])
    v, nl, kt = ctx.gather_column(p_ptr, n, pay_col)
    rows = gp[:, 0].astype(np.int64) * chunk + gp[:, 1]
    want_null = payload_null[rows]
    assert kt == abi.KEY_INT32 and np.array_equal(device_array(nl, n, "|u1"), want_null)
    assert np.array_equal(device_array(v, n, "<i4"), np.where(want_null, 0, payload[rows]))


def test_gather_column_reports_row_ids_outside_the_column(ctx):
    import torch
    dev = torch.device("cuda", 0)
    t = torch.arange(100, dtype=torch.int32, device=dev)
    col = hj.RelationBuilder(abi.LOC_DEVICE).add_value(0, t).build()
    pos = torch.tensor([[0, 5], [0, 100]], dtype=torch.int32, device=dev)   # offset 100 is one past the end
    with pytest.raises(hj.HjError) as e:
        ctx.gather_column(pos.data_ptr(), 2, col)
    assert e.value.status == abi.ERR_INVALID
    null_row = torch.tensor([[0, 7], [-1, -1]], dtype=torch.int32, device=dev)  # NULL_ROW_ID reads as NULL
    v, nl, _ = ctx.gather_column(null_row.data_ptr(), 2, col)
    class W:
        __cuda_array_interface__ = {"shape": (2,), "typestr": "|u1", "data": (nl, False), "version": 2}
    assert torch.as_tensor(W(), device=dev).cpu().tolist() == [0, 1]
