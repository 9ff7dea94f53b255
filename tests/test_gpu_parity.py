"""GPU parity tests: libhj (through the C ABI) against the CPU oracle and the reference's golden cases.

Bar: bit-exact — the sorted multiset of result rows (build RowID, probe RowID) and the row count must equal the oracle's
for every mode, encoding and edge case. Everything here needs a B200 (`-m gpu`)."""
import importlib
import zlib

import numpy as np
import pytest

import hyrise_model as hm
import oracle_binding as ob
import workloads as wl

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
pytestmark = pytest.mark.gpu

FX = hm.Fixtures()
IN_SCOPE = [(n, c, ops) for n, c, ops in FX.cases if hm.case_in_scope(ops)]
FLOAT_CASES = [(n, c, ops) for n, c, ops in FX.cases if hm.case_in_scope(ops, numeric_floats=True) and not hm.case_in_scope(ops)]
MODES = ["Inner", "Left", "Right", "Semi", "Anti"]


@pytest.fixture(scope="module")
def joins():
    """One context per forced radix-bit setting (contexts are reused across tests like an operator's would be)."""
    ctxs = {}

    def get(radix_bits=-1, flags=0, output_capacity=0):
        key = (radix_bits, flags, output_capacity)
        if key not in ctxs:
            ctxs[key] = hj.HashJoin(device=0, radix_bits=radix_bits, flags=flags, output_capacity=output_capacity)
        return ctxs[key]

    yield get
    for c in ctxs.values():
        c.close()


def gpu_engine(j):
    def run(build, probe, mode):
        j.join(build, probe, mode)
        return j.fetch()
    return run


def assert_same(j, build, probe, mode, flags=0, what=""):
    mode_id = abi.MODES[mode]
    info = j.join(build, probe, mode_id)
    b, p = j.fetch()
    ob_b, ob_p, oinfo = ob.oracle_join(build, probe, mode_id, flags=flags)
    assert info.n_rows == oinfo["n_rows"], f"{what} {mode}: row count {info.n_rows} != oracle {oinfo['n_rows']}"
    assert info.n_build_tuples == oinfo["n_build_tuples"], f"{what} {mode}: build tuples"
    assert info.n_probe_tuples == oinfo["n_probe_tuples"], f"{what} {mode}: probe tuples"
    got, want = ob.canonical(b, p), ob.canonical(ob_b, ob_p)
    if not np.array_equal(got, want):
        diff = np.nonzero((got != want).any(axis=1))[0]
        raise AssertionError(f"{what} {mode}: {len(diff)} of {len(want)} sorted rows differ; first at {diff[0]}: "
                             f"got {got[diff[0]]} want {want[diff[0]]}")
    return info


# ------------------------------------------------------------------------------------------------ reference golden cases
@pytest.mark.parametrize("radix_bits", [-1, 0, 3, 10])
@pytest.mark.parametrize("name,cite,ops", IN_SCOPE, ids=[c[0] for c in IN_SCOPE])
def test_reference_case_on_gpu(joins, name, cite, ops, radix_bits):
    ok, msg = hm.run_case(FX, ops, gpu_engine(joins(radix_bits)))
    assert ok, f"{name} ({cite}): {msg}"


# ------------------------------------------------------------------------------------------------ randomized differential
@pytest.mark.parametrize("name,cite,ops", FLOAT_CASES, ids=[c[0] for c in FLOAT_CASES])
def test_reference_float_key_cases_on_gpu(joins, name, cite, ops):
    """join_equi_test.cpp:51-94 (InnerJoinIntFloat/IntDouble, both orders) against the reference's own expected tables
    (int_float_inner.tbl, float_int_inner.tbl, int_double_inner.tbl, double_int_inner.tbl): SURVEY §8f N3."""
    assert len(FLOAT_CASES) == 4
    ok, msg = hm.run_case(FX, ops, gpu_engine(joins()))
    assert ok, f"{name} ({cite}): {msg}"


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("encoding", ["value", "dict", "mixed", "gathered"])
@pytest.mark.parametrize("nullable", [False, True])
def test_random_small(joins, mode, encoding, nullable):
    rng = np.random.default_rng(zlib.crc32(repr((mode, encoding, nullable)).encode()))
    build, probe = wl.random_relations(rng, 3000, 9000, 2500, chunk=1000, encoding=encoding, nullable=nullable)
    assert_same(joins(), build, probe, mode, what=f"{encoding} nullable={nullable}")


@pytest.mark.parametrize("n_build,n_probe", [(0, 0), (0, 7), (7, 0), (1, 1), (31, 33), (32, 32), (33, 31), (1, 5000),
                                             (5000, 1)])
@pytest.mark.parametrize("mode", MODES)
def test_edge_sizes(joins, mode, n_build, n_probe):
    rng = np.random.default_rng(n_build * 131 + n_probe)
    build, probe = wl.random_relations(rng, n_build, n_probe, 40, chunk=16, encoding="mixed", nullable=True)
    assert_same(joins(), build, probe, mode, what=f"{n_build}x{n_probe}")


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("radix_bits", [0, 5, 8, 9, 12, 15])
def test_forced_radix_bits(joins, mode, radix_bits):
    """single pass (<= 8 bits) and two passes, far more partitions than the heuristic would pick"""
    rng = np.random.default_rng(radix_bits)
    build, probe = wl.random_relations(rng, 40000, 120000, 30000, chunk=8192, encoding="dict", nullable=True)
    info = assert_same(joins(radix_bits), build, probe, mode, what=f"bits={radix_bits}")
    assert info.radix_bits == radix_bits and info.n_partitions == 1 << radix_bits


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("dtype,probe_dtype", [(np.int64, np.int64), (np.int32, np.int64), (np.int64, np.int32)])
def test_int64_and_mixed_width(joins, mode, dtype, probe_dtype):
    rng = np.random.default_rng(11)
    build, probe = wl.random_relations(rng, 20000, 50000, 15000, chunk=4096, encoding="mixed", nullable=True,
                                       dtype=dtype, probe_dtype=probe_dtype, lo=-7000)
    assert_same(joins(), build, probe, mode, what=f"{dtype.__name__}x{probe_dtype.__name__}")


def test_int64_keys_beyond_32_bits(joins):
    rng = np.random.default_rng(12)
    base = rng.integers(-2**62, 2**62, size=5000, dtype=np.int64)
    bk = base[rng.integers(0, 5000, size=20000)]
    pk = np.concatenate([base[rng.integers(0, 5000, size=30000)], rng.integers(-2**62, 2**62, size=10000)])
    for mode in MODES:
        assert_same(joins(), wl.make_relation(bk, 4096), wl.make_relation(pk, 4096), mode, what="wide int64")


@pytest.mark.parametrize("mode", MODES)
def test_sentinel_and_extreme_keys(joins, mode):
    """INT_MIN is the hash table's EMPTY marker and takes the dedicated-slot path; 0 is what NULL decodes to."""
    for dtype in (np.int32, np.int64):
        lo, hi = np.iinfo(dtype).min, np.iinfo(dtype).max
        bk = np.array([lo, lo, hi, 0, -1, 1, 5, 5, 5, lo + 1] * 30, dtype=dtype)
        pk = np.array([lo, hi, 0, 7, -1, 5, lo + 1, hi - 1] * 100, dtype=dtype)
        bn = np.zeros(len(bk), np.uint8); bn[::7] = 1
        pn = np.zeros(len(pk), np.uint8); pn[::5] = 1
        for enc in ("value", "dict"):
            assert_same(joins(), wl.make_relation(bk, 64, enc, bn), wl.make_relation(pk, 64, enc, pn), mode,
                        what=f"extremes {dtype.__name__} {enc}")


@pytest.mark.parametrize("mode", MODES)
def test_all_duplicate_keys_overflow_partition(joins, mode):
    """One key repeated far beyond a shared-memory table: the build partition is cut into chunks (multi path)."""
    rng = np.random.default_rng(5)
    bk = np.full(40000, 42, dtype=np.int32)
    bk[:100] = rng.integers(1, 1000, 100)
    pk = rng.integers(1, 1000, 300).astype(np.int32)
    pk[::10] = 42
    pn = (rng.random(300) < 0.1).astype(np.uint8)
    info = assert_same(joins(), wl.make_relation(bk, 8192), wl.make_relation(pk, 128, "value", pn), mode, what="all-dup")
    assert info.n_overflow_partitions >= 1


@pytest.mark.parametrize("mode", MODES)
def test_skewed_build_many_chunks_with_radix(joins, mode):
    """Zipf build keys with forced partitioning: hot partitions overflow, the others do not."""
    rng = np.random.default_rng(6)
    bk = wl.zipf_keys(rng, 150000, 20000, 1.0, np.int64)
    pk = rng.integers(1, 30000, 40000).astype(np.int64)
    info = assert_same(joins(4), wl.make_relation(bk, 16384), wl.make_relation(pk, 16384), mode, what="zipf")
    assert info.n_overflow_partitions >= 1


@pytest.mark.parametrize("mode", MODES)
def test_no_match_and_all_null(joins, mode):
    rng = np.random.default_rng(8)
    bk = rng.integers(1, 1000, 5000).astype(np.int32)
    pk = rng.integers(2000, 3000, 7000).astype(np.int32)
    assert_same(joins(), wl.make_relation(bk, 1024), wl.make_relation(pk, 1024), mode, what="no-match")
    alln = np.ones(7000, np.uint8)
    for enc in ("value", "dict"):
        assert_same(joins(), wl.make_relation(bk, 1024, enc), wl.make_relation(pk, 1024, enc, alln), mode, what="null probe")
        assert_same(joins(), wl.make_relation(bk, 1024, enc, np.ones(5000, np.uint8)), wl.make_relation(pk, 1024, enc), mode,
                    what="null build")


@pytest.mark.parametrize("mode", MODES)
def test_irregular_chunks_and_chunk_ids(joins, mode):
    """Reference tables out of a TableScan have arbitrary chunk sizes (position -> RowID by binary search) and the
    shim may pass any ChunkIDs; empty chunks in the middle."""
    rng = np.random.default_rng(9)
    bk = rng.integers(1, 4000, 10000).astype(np.int32)
    pk = rng.integers(1, 4000, 21000).astype(np.int32)
    bsz = [1, 0, 4999, 17, 0, 4983]
    psz = [7000, 1, 1, 0, 13998]
    build = wl.make_relation(bk, 0, "mixed", chunk_sizes=bsz, chunk_ids=[5, 9, 2, 77, 3, 1000])
    probe = wl.make_relation(pk, 0, "mixed", (rng.random(21000) < 0.05).astype(np.uint8), chunk_sizes=psz)
    assert_same(joins(), build, probe, mode, what="irregular")


def test_gathered_row_offsets(joins):
    """HJ_SEG_GATHERED with explicit row_offsets (quirk Q2 switch): reported ChunkOffsets come from the array."""
    rng = np.random.default_rng(10)
    bk = rng.integers(1, 500, 2000).astype(np.int32)
    pk = rng.integers(1, 500, 3000).astype(np.int32)
    rb = hj.RelationBuilder()
    ro = rng.permutation(3000).astype(np.uint32)
    rb.add_value(3, pk, (rng.random(3000) < 0.1).astype(np.uint8), kind=abi.SEG_GATHERED, row_offsets=ro)
    for mode in MODES:
        assert_same(joins(), wl.make_relation(bk, 512), rb.build(), mode, what="row_offsets")


@pytest.mark.parametrize("mode", ["Left", "Right"])
@pytest.mark.parametrize("encoding", ["value", "dict", "gathered"])
def test_null_probe_literal_vs_sql(joins, mode, encoding):
    """Quirk Q4: by default a NULL probe key of an outer join is looked up as its stored value / 0 like the reference does;
    HJ_FLAG_SQL_NULL_PROBE makes it never match. Build side contains the key 0 so that the two differ."""
    rng = np.random.default_rng(13)
    bk = rng.integers(0, 50, 3000).astype(np.int32)  # includes 0
    pk = rng.integers(0, 80, 5000).astype(np.int32)
    pn = (rng.random(5000) < 0.2).astype(np.uint8)
    build, probe = wl.make_relation(bk, 1024, encoding), wl.make_relation(pk, 1024, encoding, pn)
    lit = assert_same(joins(), build, probe, mode, flags=0, what="literal")
    sql = assert_same(joins(-1, abi.FLAG_SQL_NULL_PROBE), build, probe, mode, flags=abi.FLAG_SQL_NULL_PROBE, what="sql")
    assert lit.n_rows > sql.n_rows
    assert_same(joins(9, abi.FLAG_SQL_NULL_PROBE), build, probe, mode, flags=abi.FLAG_SQL_NULL_PROBE, what="sql 2-pass")

@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("radix_bits", [9, 14, 15])
@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_two_pass_partition_paths_agree(joins, mode, radix_bits, dtype):
    """Two-pass partitioning has two implementations with identical results: the default slab-local pass 1 fused into the
    histogram kernel + fragment-gathering pass 2 (hj_fused.cuh), and HJ_FLAG_TWO_SWEEP_PARTITION (K1 + scatter1 + scatter2,
    what the sharded local join uses). Skewed keys make single pass-2 work items exceed one tile (batched)."""
    rng = np.random.default_rng(radix_bits * 7 + (dtype == np.int64))
    n_b, n_p = 150_000, 400_000
    bk = wl.zipf_keys(rng, n_b, 5000, 1.1, dtype)
    pk = rng.integers(1, 5001, n_p).astype(dtype)
    pk[: n_p // 2] = 6000  # half of the probe rows (no build partner) in ONE level-1 fragment per slab
    bn = (rng.random(n_b) < 0.03).astype(np.uint8)
    pn = (rng.random(n_p) < 0.03).astype(np.uint8)
    build, probe = wl.make_relation(bk, 1 << 15, "dict", bn), wl.make_relation(pk, 30_000, "mixed", pn)
    assert_same(joins(radix_bits), build, probe, mode, what="fused")
    assert_same(joins(radix_bits, abi.FLAG_TWO_SWEEP_PARTITION), build, probe, mode, what="two-sweep")
    if mode in ("Left", "Right"):
        f = abi.FLAG_SQL_NULL_PROBE
        assert_same(joins(radix_bits, f), build, probe, mode, flags=f, what="fused, routed NULLs")
        assert_same(joins(radix_bits, f | abi.FLAG_TWO_SWEEP_PARTITION), build, probe, mode, flags=f, what="two-sweep, routed NULLs")

@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("encoding,dtype", [("dict", np.int32), ("mixed", np.int64)])
def test_bound_output_streamed_join(mode, encoding, dtype):
    """hj_bind_output + host inputs of many chunks: the probe side is joined slice by slice while the next slice is copied in
    and the previous PosLists are copied out. Same sorted result as the oracle; rows of one slice stay contiguous."""
    import torch
    rng = np.random.default_rng(77)
    n_b, n_p, chunk = 60_000, 500_000, 20_000
    build, probe = wl.random_relations(rng, n_b, n_p, 40_000, chunk=chunk, encoding=encoding, nullable=True, dtype=dtype)
    mode_id = abi.MODES[mode]
    ob_b, ob_p, oinfo = ob.oracle_join(build, probe, mode_id)
    cap = oinfo["n_rows"] + 5
    out_b = torch.empty((cap, 2), dtype=torch.int32).pin_memory().numpy().view(np.uint32)
    out_p = torch.empty((cap, 2), dtype=torch.int32).pin_memory().numpy().view(np.uint32)
    two = mode not in ("Semi", "Anti")
    for bits in (-1, 10):
        j = hj.HashJoin(device=0, radix_bits=bits, stream_slice_rows=50_000)  # 10 slices of 2-3 chunks
        try:
            j.bind_output(out_b if two else None, out_p)
            info = j.join(build, probe, mode_id)
            assert info.n_probe_slices >= 8, "the join was not streamed"
            assert info.n_rows == oinfo["n_rows"] and info.n_probe_tuples == oinfo["n_probe_tuples"]
            b, p = j.fetch(out_b if two else None, out_p)  # no-op: already delivered
            got, want = ob.canonical(b if two else None, p), ob.canonical(ob_b, ob_p)
            assert np.array_equal(got, want), f"{mode} bits={bits}: streamed result differs"
            # the binding lapsed: the next join keeps its result on the device as usual
            info2 = j.join(build, probe, mode_id)
            b2, p2 = j.fetch()
            assert info2.n_probe_slices == 0 and np.array_equal(ob.canonical(b2, p2), want)
            # too small a binding is reported, not truncated
            j.bind_output(out_b[: max(1, cap // 2)] if two else None, out_p[: max(1, cap // 2)])
            if oinfo["n_rows"] > cap // 2:
                with pytest.raises(hj.HjError) as e:
                    j.join(build, probe, mode_id)
                assert e.value.status == abi.ERR_CAPACITY
        finally:
            j.close()


@pytest.mark.parametrize("mode", ["Inner", "Left"])
def test_output_capacity_overflow_reprobes(joins, mode):
    """Result far larger than the first buffer estimate (probe rows + one 128 Ki-row chunk per writer CTA): the library
    counts exactly, grows the buffers and probes again; the chunked writer's holes are compacted away."""
    n = 7200  # ~51 M result rows; the first estimate holds <= 39 M (Left: two writer kernels x 148 CTAs x 128 Ki rows)
    bk = np.full(n, 7, dtype=np.int32)
    bk[:50] = np.arange(100, 150)
    pk = np.full(n, 7, dtype=np.int32)
    pk[:60] = np.arange(120, 180)
    info = assert_same(joins(), wl.make_relation(bk, 2048), wl.make_relation(pk, 2048), mode, what="huge fan-out")
    assert info.n_probe_passes == 2 and info.n_rows > 50_000_000


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("encoding,dtype", [("dict", np.int32), ("value", np.int64)])
def test_medium_1m(joins, mode, encoding, dtype):
    """~1 Mi-row chunks +-1, heuristic radix bits (two passes)."""
    rng = np.random.default_rng(15)
    n_b, n_p = (1 << 20) + 1, 3 * (1 << 20) - 1
    build, probe = wl.random_relations(rng, n_b, n_p, 1 << 20, chunk=1 << 19, encoding=encoding, nullable=True,
                                       dtype=dtype, null_frac=0.01)
    info = assert_same(joins(), build, probe, mode, what="1Mi")
    assert info.radix_bits >= 8


def test_attr_widths(joins):
    rng = np.random.default_rng(16)
    n = 300000
    for domain, width in ((200, 1), (60000, 2), (400000, 4)):
        bk = rng.integers(1, domain + 1, n).astype(np.int32)
        pk = rng.integers(1, domain + 1, n).astype(np.int32)
        build = wl.make_relation(bk, n, "dict", (rng.random(n) < 0.02).astype(np.uint8))
        assert build.segs[0].attr_width == width
        probe = wl.make_relation(pk, n, "dict")
        for mode in ("Semi", "Anti") + (("Inner", "Left") if width > 1 else ()):
            assert_same(joins(), build, probe, mode, what=f"attr u{8 * width}")


def test_device_resident_inputs(joins):
    """HJ_LOC_DEVICE: segment arrays are torch CUDA tensors; result also read through hj_result_device."""
    import torch
    rng = np.random.default_rng(17)
    bk = rng.integers(1, 50000, 100000).astype(np.int32)
    pk = rng.integers(1, 50000, 300000).astype(np.int32)
    host_b, host_p = wl.make_relation(bk, 32768, "dict"), wl.make_relation(pk, 32768, "dict")
    dev = torch.device("cuda:0")
    rels = []
    for keys in (bk, pk):
        rb = hj.RelationBuilder(abi.LOC_DEVICE)
        for c, s in enumerate(range(0, len(keys), 32768)):
            d, a, nv = wl.dict_encode(keys[s:s + 32768])
            rb.add_dict(c, torch.from_numpy(d).to(dev), torch.from_numpy(a.astype(np.int32) if a.dtype == np.uint32 else a.view(np.int16)).to(dev), nv)
        rels.append(rb.build())
    torch.cuda.synchronize()
    j = joins()
    info = j.join(rels[0], rels[1], "Inner")
    b, p = j.fetch()
    ob_b, ob_p, _ = ob.oracle_join(host_b, host_p, abi.MODE_INNER)
    assert np.array_equal(ob.canonical(b, p), ob.canonical(ob_b, ob_p))
    bptr, pptr, n = j.result_device()
    assert n == info.n_rows and bptr and pptr


def test_translate(joins):
    rng = np.random.default_rng(18)
    lists = [np.stack([rng.integers(0, 9, m), rng.integers(0, 1000, m)], axis=1).astype(np.uint32) for m in (5, 0, 700, 33)]
    n = 5000
    chunk = rng.choice([0, 2, 3], size=n)
    off = np.array([rng.integers(0, len(lists[c])) for c in chunk])
    rows = np.stack([chunk, off], axis=1).astype(np.uint32)
    rows[::11] = (abi.INVALID_CHUNK_ID, abi.INVALID_CHUNK_OFFSET)
    want = rows.copy()
    for i in range(n):
        if want[i, 1] != abi.INVALID_CHUNK_OFFSET:
            want[i] = lists[want[i, 0]][want[i, 1]]
    got = joins().translate(rows.copy(), lists)
    assert np.array_equal(got, want)
    bad = rows.copy()
    bad[1] = (1, 0)  # chunk 1 is empty
    with pytest.raises(hj.HjError):
        joins().translate(bad, lists)


def test_error_paths(joins):
    j = joins()
    rng = np.random.default_rng(19)
    build, probe = wl.random_relations(rng, 10, 10, 5, chunk=4)
    for bad_mode in (abi.MODE_OUTER, abi.MODE_CROSS, 42):
        with pytest.raises(hj.HjError) as e:
            j.join(build, probe, bad_mode)
        assert e.value.status == abi.ERR_INVALID  # the shim turns this into Fail() (utils/assert.hpp:49)
    with pytest.raises(hj.HjError):
        j.fetch()  # no successful join since the failed one
    probe.n_rows += 1
    with pytest.raises(hj.HjError):
        j.join(build, probe, abi.MODE_INNER)
    probe.n_rows -= 1
    info = j.join(build, probe, abi.MODE_INNER)
    if info.n_rows > 1:  # caller buffer smaller than the result
        small = np.empty((1, 2), np.uint32)
        st = j._l.hj_fetch(j._ctx, small.ctypes.data, small.ctypes.data, 1)
        assert st == abi.ERR_CAPACITY
