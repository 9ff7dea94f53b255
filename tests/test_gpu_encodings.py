"""FrameOfReference and RunLength key columns, decoded on the device. The encoders below restate the
reference's (frame_of_reference_encoder.hpp:30-90: 2048-row blocks, NULL stored as 0, offset = value - block minimum;
run_length_encoder.hpp:20-60: a run ends where value or NULL-ness changes, inclusive end positions). The oracle joins the
same rows as ValueColumns (both iterables yield (stored value, is_null) per row, like ValueColumnIterable)."""
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


def for_encode(values, nulls):
    v = values.copy()
    if nulls is not None:
        v[nulls != 0] = 0
    n = len(v)
    minima = np.array([v[b:b + 2048].min() for b in range(0, n, 2048)], dtype=values.dtype)
    off = (v.astype(np.int64) - np.repeat(minima.astype(np.int64), 2048)[:n])
    assert off.min(initial=0) >= 0 and off.max(initial=0) <= 0xFFFFFFFF
    width = np.uint8 if off.max(initial=0) < 256 else np.uint16 if off.max(initial=0) < 65536 else np.uint32
    return minima, off.astype(width), v


def rle_encode(values, nulls):
    n = len(values)
    nl = np.zeros(n, np.uint8) if nulls is None else nulls
    change = np.ones(n, dtype=bool)
    change[1:] = (nl[1:] != nl[:-1]) | ((nl[1:] == 0) & (values[1:] != values[:-1]))
    starts = np.nonzero(change)[0]
    ends = np.append(starts[1:] - 1, n - 1).astype(np.uint32) if n else np.zeros(0, np.uint32)
    return np.ascontiguousarray(values[starts]), ends, np.ascontiguousarray(nl[starts]), starts


bp128_pack = wl.bp128_pack


def encoded_relation(keys, nulls, chunk, encoding):
    """(relation of FOR / RLE / mixed chunks, the equivalent ValueColumn relation for the oracle)"""
    rb = hj.RelationBuilder(abi.LOC_HOST)
    plain = keys.copy()
    for c, start in enumerate(range(0, len(keys), chunk)):
        v = np.ascontiguousarray(keys[start:start + chunk])
        nl = None if nulls is None else np.ascontiguousarray(nulls[start:start + chunk])
        enc = encoding if encoding != "mixed" else ("for", "rle", "value")[c % 3]
        if enc == "dict_bp128":
            d, attr, nv = wl.dict_encode(v, nl)
            rb.add_dict_bp128(c, d, bp128_pack(attr), len(v), nv)
            if nl is not None:  # a NULL row of a dictionary column decodes to T{} (dictionary_column_iterable.hpp:80)
                stored = v.copy()
                stored[nl != 0] = 0
                plain[start:start + chunk] = stored
        elif enc == "for_bp128":
            minima, off, stored = for_encode(v, nl)
            rb.add_for_bp128(c, minima, bp128_pack(off), len(v), nl)
            plain[start:start + chunk] = stored
        elif enc == "for":
            minima, off, stored = for_encode(v, nl)
            rb.add_for(c, minima, off, nl)
            plain[start:start + chunk] = stored
        elif enc == "rle":
            rv, ends, rn, starts = rle_encode(v, nl)
            rb.add_rle(c, len(v), rv, ends, rn if nl is not None else None)
            run_of = np.searchsorted(ends, np.arange(len(v)), side="left")
            plain[start:start + chunk] = rv[run_of]  # a NULL run reports its first row's stored value
        else:
            rb.add_value(c, v, nl)
    return rb.build(), wl.make_relation(plain, chunk, "value", nulls)


def check(j, build, probe, o_build, o_probe, mode, what):
    mode_id = abi.MODES[mode]
    info = j.join(build, probe, mode_id)
    b, p = j.fetch()
    ob_b, ob_p, oinfo = ob.oracle_join(o_build, o_probe, mode_id)
    assert info.n_rows == oinfo["n_rows"], f"{what} {mode}: {info.n_rows} rows, oracle {oinfo['n_rows']}"
    assert np.array_equal(ob.canonical(b, p), ob.canonical(ob_b, ob_p)), f"{what} {mode}: sorted result rows differ"


@pytest.fixture(scope="module")
def join():
    j = hj.HashJoin(device=0)
    yield j
    j.close()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("encoding", ["for", "rle", "mixed"])
@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_for_and_rle_key_columns(join, mode, encoding, dtype):
    rng = np.random.default_rng(hash((mode, encoding, dtype.__name__)) % 2**32)
    # sorted-ish keys with repeats: runs for RLE, small per-block ranges for FOR (and one wide block)
    bk = np.sort(wl.random_keys(rng, 20_000, 3000, dtype))
    pk = np.repeat(wl.random_keys(rng, 9_000, 3500, dtype), rng.integers(1, 9, 9_000))[:40_000]
    pk[5000:5100] += 1_000_000  # one block with a wide value range (32-bit offsets)
    if dtype == np.int64:  # minima beyond 32 bits; a block's range must fit uint32 (frame_of_reference_encoder.hpp:72)
        bk += 2**40
        pk += 2**40
    bn = (rng.random(len(bk)) < 0.03).astype(np.uint8)
    pn = np.repeat((rng.random(9_000) < 0.05).astype(np.uint8), 5)[:len(pk)].copy()
    if dtype == np.int64:  # NULL rows are stored as 0, so a nullable block of such values is not FOR-encodable at all
        bn = pn = None
    build, o_build = encoded_relation(bk, bn, 4500, encoding)
    probe, o_probe = encoded_relation(pk, pn, 6000, encoding)
    check(join, build, probe, o_build, o_probe, mode, encoding)


@pytest.mark.parametrize("mode", ["Inner", "Left", "Anti"])
@pytest.mark.parametrize("encoding", ["dict_bp128", "for_bp128"])
@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_simd_bp128_vectors(join, mode, encoding, dtype):
    """SimdBp128-compressed attribute vectors (DictionaryColumn) and offset vectors (FrameOfReferenceColumn): ragged last
    meta block and last block, blocks of width 0, and wide blocks whose values straddle 32-bit words"""
    rng = np.random.default_rng(hash((mode, encoding, dtype.__name__)) % 2**32)
    bk = np.sort(wl.random_keys(rng, 9_000, 3000, dtype))
    pk = wl.random_keys(rng, 2 * 2048 + 700 + 45, 3500, dtype)   # 2 full meta blocks + 5 full blocks + 105 values
    pk[300:428] = pk[300]                                          # one block of equal values (small width)
    pk[2048:2048 + 256] += 3_000_000                               # wide offsets in two blocks
    bn = (rng.random(len(bk)) < 0.03).astype(np.uint8)
    pn = (rng.random(len(pk)) < 0.03).astype(np.uint8)
    build, o_build = encoded_relation(bk, bn, 4100, encoding)
    probe, o_probe = encoded_relation(pk, pn, len(pk), encoding)
    check(join, build, probe, o_build, o_probe, mode, encoding)


def test_simd_bp128_device_resident_and_referenced(join):
    """the stream already on the device (meta-block offsets computed by a kernel), and behind a ReferenceColumn"""
    import torch
    rng = np.random.default_rng(21)
    keys = wl.random_keys(rng, 5000, 900, np.int32)
    d, attr, nv = wl.dict_encode(keys, None)
    stream = bp128_pack(attr)
    dev = torch.device("cuda", 0)
    data_dev = hj.RelationBuilder(abi.LOC_DEVICE).add_dict_bp128(
        0, torch.from_numpy(d.copy()).to(dev), torch.from_numpy(stream.view(np.int32).copy()).to(dev), len(keys), nv).build()
    bk = wl.random_keys(rng, 700, 900, np.int32)
    build_dev = hj.RelationBuilder(abi.LOC_DEVICE).add_value(0, torch.from_numpy(bk.copy()).to(dev)).build()
    o_build, o_probe = wl.make_relation(bk, 700), wl.make_relation(keys, 5000)
    check(join, build_dev, data_dev, o_build, o_probe, "Inner", "bp128 on device")
    g = rng.integers(0, len(keys), 8000)
    pos = np.stack([np.zeros_like(g), g], axis=1).astype(np.uint32)
    data_host = hj.RelationBuilder().add_dict_bp128(0, d, stream, len(keys), nv).build()
    ref = hj.RelationBuilder().add_reference(0, pos, data_host, abi.KEY_INT32).build()
    check(join, wl.make_relation(bk, 700), ref, o_build, wl.make_relation(keys[g], 8000, "gathered"), "Inner", "ref over bp128")


def test_reference_over_encoded_table(join):
    """a reference table whose data table is FOR/RLE-encoded (point access: minima[off/2048] + offset, run search)"""
    rng = np.random.default_rng(9)
    keys = np.sort(wl.random_keys(rng, 30_000, 2000, np.int32))
    nulls = (rng.random(30_000) < 0.02).astype(np.uint8)
    data_rel, _ = encoded_relation(keys, nulls, 5000, "mixed")
    g = rng.integers(0, len(keys), 45_000)
    pos = np.stack([g // 5000, g % 5000], axis=1).astype(np.uint32)
    rb = hj.RelationBuilder()
    for c in range(3):
        rb.add_reference(c, np.ascontiguousarray(pos[c * 15_000:(c + 1) * 15_000]), data_rel, abi.KEY_INT32)
    probe = rb.build()
    gk, gn = keys[g].copy(), nulls[g]
    gk[gn != 0] = 0
    o_probe = wl.make_relation(gk, 15_000, "gathered", gn)
    bk = wl.random_keys(rng, 4000, 2000, np.int32)
    build = wl.make_relation(bk, 1000, "dict")
    for mode in ("Inner", "Left", "Anti"):
        check(join, build, probe, build, o_probe, mode, "ref over encoded")


def test_encoded_descriptor_errors(join):
    ok = wl.make_relation(np.arange(10, dtype=np.int32), 10)
    s = abi.Segment()
    s.chunk_id, s.n_rows, s.kind, s.key_type, s.attr_width = 0, 10, abi.SEG_FOR, abi.KEY_INT32, 3
    rel = abi.Relation()
    arr = (abi.Segment * 1)(s)
    import ctypes as C
    rel.segs, rel.n_segs, rel.location, rel.n_rows = C.cast(arr, C.POINTER(abi.Segment)), 1, abi.LOC_HOST, 10
    with pytest.raises(hj.HjError) as e:
        join.join(ok, rel, abi.MODE_INNER)
    assert e.value.status == abi.ERR_INVALID and "FOR" in str(e.value)
