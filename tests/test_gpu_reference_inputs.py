"""ReferenceColumn inputs handed over as stored (PosList + the referenced data table's key column,
HJ_SEG_REFERENCE); the library gathers the keys on the device. Checked against the oracle's join over the same rows
gathered with numpy (what the reference's ReferenceColumnIterable yields, reference_column_iterable.hpp:59-90)."""
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
MODES = ["Inner", "Left", "Right", "Semi", "Anti"]
NULL_OFF = 0xFFFFFFFF


def reference_table(rng, keys, nulls, data_chunk, n_rows, ref_chunk, encoding, null_rowid_frac=0.03):
    """(relation with HJ_SEG_REFERENCE chunks, the same rows as a GATHERED relation for the oracle)."""
    dtype = keys.dtype
    data_rel = wl.make_relation(keys, data_chunk, encoding, nulls)
    g = rng.integers(0, len(keys), n_rows)
    is_null_rid = rng.random(n_rows) < null_rowid_frac
    pos = np.stack([g // data_chunk, g % data_chunk], axis=1).astype(np.uint32)
    pos[is_null_rid] = (NULL_OFF, NULL_OFF)  # NULL_ROW_ID (types.hpp:138)
    gk = keys[g].copy()
    gn = is_null_rid.astype(np.uint8)
    if nulls is not None:
        gn |= nulls[g]
    gk[gn != 0] = 0  # NULL yields T{} (reference_column_iterable.hpp:60)
    rb = hj.RelationBuilder(abi.LOC_HOST)
    for c, start in enumerate(range(0, n_rows, ref_chunk)):
        rb.add_reference(c, np.ascontiguousarray(pos[start:start + ref_chunk]), data_rel,
                         abi.KEY_INT64 if dtype == np.int64 else abi.KEY_INT32)
    return rb.build(), wl.make_relation(gk, ref_chunk, "gathered", gn)


@pytest.fixture(scope="module")
def join():
    j = hj.HashJoin(device=0)
    yield j
    j.close()


def check(j, build, probe, o_build, o_probe, mode, what):
    mode_id = abi.MODES[mode]
    info = j.join(build, probe, mode_id)
    b, p = j.fetch()
    ob_b, ob_p, oinfo = ob.oracle_join(o_build, o_probe, mode_id)
    assert info.n_rows == oinfo["n_rows"], f"{what} {mode}: {info.n_rows} rows, oracle {oinfo['n_rows']}"
    assert info.n_build_tuples == oinfo["n_build_tuples"] and info.n_probe_tuples == oinfo["n_probe_tuples"]
    assert np.array_equal(ob.canonical(b, p), ob.canonical(ob_b, ob_p)), f"{what} {mode}: sorted result rows differ"
    return info


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("encoding", ["value", "dict", "mixed"])
@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_reference_inputs_both_sides(join, mode, encoding, dtype):
    rng = np.random.default_rng(hash((mode, encoding, dtype.__name__)) % 2**32)
    bk, pk = wl.random_keys(rng, 30_000, 4000, dtype), wl.random_keys(rng, 50_000, 4000, dtype)
    bn, pn = (rng.random(30_000) < 0.04).astype(np.uint8), (rng.random(50_000) < 0.04).astype(np.uint8)
    build, o_build = reference_table(rng, bk, bn, 7000, 20_000, 6000, encoding)
    probe, o_probe = reference_table(rng, pk, pn, 9000, 70_000, 16_384, encoding)
    check(join, build, probe, o_build, o_probe, mode, f"ref x ref {encoding}")


@pytest.mark.parametrize("mode", ["Inner", "Left", "Anti"])
def test_reference_probe_against_plain_build_two_pass(mode):
    """a reference table probing a data table (the TPC-H shape: filtered lineitem x orders), two-pass partitioning, and
    the same through a bound output (streamed over the gathered chunks)"""
    import torch
    rng = np.random.default_rng(5)
    bk = rng.permutation(200_000).astype(np.int32) + 1
    pk = wl.random_keys(rng, 300_000, 220_000, np.int32)
    build = wl.make_relation(bk, 1 << 16, "dict")
    probe, o_probe = reference_table(rng, pk, None, 1 << 16, 400_000, 50_000, "dict")
    j = hj.HashJoin(device=0, radix_bits=10, stream_slice_rows=100_000)
    try:
        info = check(j, build, probe, build, o_probe, mode, "ref probe")
        cap = info.n_rows + 3
        two = mode != "Anti"
        out_b = torch.empty((cap, 2), dtype=torch.int32).pin_memory().numpy().view(np.uint32)
        out_p = torch.empty((cap, 2), dtype=torch.int32).pin_memory().numpy().view(np.uint32)
        j.bind_output(out_b if two else None, out_p)
        info2 = j.join(build, probe, abi.MODES[mode])
        assert info2.n_probe_slices >= 2 and info2.n_rows == info.n_rows
        ob_b, ob_p, _ = ob.oracle_join(build, o_probe, abi.MODES[mode])
        assert np.array_equal(ob.canonical(out_b[:info2.n_rows] if two else None, out_p[:info2.n_rows]), ob.canonical(ob_b, ob_p))
    finally:
        j.close()


def test_reference_device_resident(join):
    """PosLists and the referenced column already on the device (HJ_LOC_DEVICE)"""
    import torch
    rng = np.random.default_rng(6)
    keys = wl.random_keys(rng, 40_000, 3000, np.int32)
    g = rng.integers(0, len(keys), 60_000)
    pos = np.stack([g // 10_000, g % 10_000], axis=1).astype(np.uint32)
    dev = torch.device("cuda", 0)
    data = hj.RelationBuilder(abi.LOC_DEVICE)
    for c in range(4):
        data.add_value(c, torch.from_numpy(keys[c * 10_000:(c + 1) * 10_000].copy()).to(dev))
    data_rel = data.build()
    rb = hj.RelationBuilder(abi.LOC_DEVICE)
    for c in range(3):
        rb.add_reference(c, torch.from_numpy(pos[c * 20_000:(c + 1) * 20_000].copy().view(np.int32)).to(dev), data_rel, abi.KEY_INT32)
    probe = rb.build()
    bk = wl.random_keys(rng, 5000, 3000, np.int32)
    build_dev = hj.RelationBuilder(abi.LOC_DEVICE).add_value(0, torch.from_numpy(bk.copy()).to(dev)).build()
    info = join.join(build_dev, probe, abi.MODE_INNER)
    b, p = join.fetch()
    ob_b, ob_p, oinfo = ob.oracle_join(wl.make_relation(bk, 5000), wl.make_relation(keys[g], 20_000, "gathered"), abi.MODE_INNER)
    assert info.n_rows == oinfo["n_rows"] and np.array_equal(ob.canonical(b, p), ob.canonical(ob_b, ob_p))


def test_reference_errors(join):
    keys = np.arange(100, dtype=np.int32)
    data_rel = wl.make_relation(keys, 50)
    bad = np.array([[0, 1], [2, 0], [1, 3]], dtype=np.uint32)  # chunk 2 does not exist
    probe = hj.RelationBuilder().add_reference(0, bad, data_rel, abi.KEY_INT32).build()
    with pytest.raises(hj.HjError) as e:
        join.join(data_rel, probe, abi.MODE_INNER)
    assert e.value.status == abi.ERR_INVALID and "outside the referenced table" in str(e.value)
    beyond = np.array([[0, 50]], dtype=np.uint32)  # offset == chunk size
    probe = hj.RelationBuilder().add_reference(0, beyond, data_rel, abi.KEY_INT32).build()
    with pytest.raises(hj.HjError):
        join.join(data_rel, probe, abi.MODE_INNER)
    # a reference to a reference is not a thing in Hyrise (and is refused)
    gathered = wl.make_relation(keys, 50, "gathered")
    probe = hj.RelationBuilder().add_reference(0, np.array([[0, 1]], dtype=np.uint32), gathered, abi.KEY_INT32).build()
    with pytest.raises(hj.HjError) as e:
        join.join(data_rel, probe, abi.MODE_INNER)
    assert e.value.status == abi.ERR_INVALID
    ok = hj.RelationBuilder().add_reference(0, np.array([[1, 49], [0, 0]], dtype=np.uint32), data_rel, abi.KEY_INT32).build()
    assert join.join(data_rel, ok, abi.MODE_INNER).n_rows == 2
