"""SURVEY §8(f) N3 (numeric part): float/double key columns and int x float pairs. JoinHashTraits
(join_hash/hash_traits.hpp:14-35) hashes in the floating type; keys match when their values in that type have the same
bits (the reference finds partners through a hash of the bytes, join_hash.cpp:196) and are not NaN (operator==).
The expected result is computed here from that rule with numpy on the converted keys and handed to the integer oracle;
the reference's own int x float / int x double fixtures are replayed through the C++ operator (host/shim_test)."""
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


def hashed_bits(keys, hash_dtype, build_side):
    """keys converted to the hashed floating type, as the integer bit patterns two rows must share to join"""
    with np.errstate(invalid="ignore"):
        f = keys.astype(hash_dtype)
    bits = f.view(np.int32 if hash_dtype == np.float32 else np.int64).copy()
    nan = np.isnan(f)
    if hash_dtype == np.float32:
        bits[nan] = np.int32(0x7FC00001 if build_side else 0x7FC00000)
    else:
        bits[nan] = np.int64(0x7FF8000000000001 if build_side else 0x7FF8000000000000)
    return bits


def make_keys(rng, n, dtype, domain=3000):
    k = rng.integers(-domain, domain, n)
    if dtype in (np.float32, np.float64):
        k = (k / 4.0).astype(dtype)          # quarters: exactly representable, many equal values
        k[rng.random(n) < 0.01] = np.nan
        k[rng.random(n) < 0.01] = -0.0
        k[rng.random(n) < 0.01] = 0.0
        return k
    k = k.astype(dtype)
    if dtype == np.int64:
        k[:50] = 2**53 + np.arange(50)        # not exactly representable as double / float
    if dtype == np.int32:
        k[:50] = 16777217 + np.arange(50)     # > 2^24: rounds when hashed as float
    return k


@pytest.fixture(scope="module")
def join():
    j = hj.HashJoin(device=0)
    yield j
    j.close()


PAIRS = [(np.float32, np.float32), (np.float64, np.float64), (np.int32, np.float32), (np.float32, np.int32),
         (np.int64, np.float32), (np.int32, np.float64), (np.float64, np.int64), (np.float32, np.float64)]


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("bt,pt", PAIRS, ids=[f"{a.__name__}x{b.__name__}" for a, b in PAIRS])
@pytest.mark.parametrize("encoding", ["value", "dict"])
def test_float_and_mixed_numeric_keys(join, mode, bt, pt, encoding):
    rng = np.random.default_rng(hash((mode, bt.__name__, pt.__name__, encoding)) % 2**32)
    bk, pk = make_keys(rng, 12_000, bt), make_keys(rng, 30_000, pt)
    bn, pn = (rng.random(len(bk)) < 0.03).astype(np.uint8), (rng.random(len(pk)) < 0.03).astype(np.uint8)
    if encoding == "dict":  # a dictionary (sorted, unique by ==) keeps one of -0.0 / +0.0: leave the pair to "value"
        for k in (bk, pk):
            if k.dtype.kind == "f":
                k[k == 0] = 0.0
    hash_dtype = np.float64 if np.float64 in (bt, pt) else np.float32
    build, probe = wl.make_relation(bk, 5000, encoding, bn), wl.make_relation(pk, 7000, encoding, pn)
    o_build = wl.make_relation(hashed_bits(bk, hash_dtype, True), 5000, encoding, bn)
    o_probe = wl.make_relation(hashed_bits(pk, hash_dtype, False), 7000, encoding, pn)
    mode_id = abi.MODES[mode]
    info = join.join(build, probe, mode_id)
    b, p = join.fetch()
    ob_b, ob_p, oinfo = ob.oracle_join(o_build, o_probe, mode_id)
    assert info.n_rows == oinfo["n_rows"], f"{info.n_rows} rows, expected {oinfo['n_rows']}"
    assert np.array_equal(ob.canonical(b, p), ob.canonical(ob_b, ob_p))


def test_float_semantics_spot_checks(join):
    """NaN joins nothing, -0.0 does not meet +0.0, 16777217 (int) meets 16777216.0f"""
    b = wl.make_relation(np.array([np.nan, -0.0, 0.0, 16777216.0, 2.5], dtype=np.float32), 8)
    p = wl.make_relation(np.array([np.nan, 0.0, -0.0], dtype=np.float32), 8)
    info = join.join(b, p, abi.MODE_INNER)
    bs, ps = join.fetch()
    assert sorted(zip(bs[:, 1].tolist(), ps[:, 1].tolist())) == [(1, 2), (2, 1)]
    p = wl.make_relation(np.array([16777217, 16777216, 2], dtype=np.int32), 8)
    info = join.join(b, p, abi.MODE_INNER)
    bs, ps = join.fetch()
    assert sorted(zip(bs[:, 1].tolist(), ps[:, 1].tolist())) == [(3, 0), (3, 1)]
    info = join.join(b, p, abi.MODE_ANTI)  # probe rows without partner: the 2
    _, ps = join.fetch()
    assert ps[:, 1].tolist() == [2]
