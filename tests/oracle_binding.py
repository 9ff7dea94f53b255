"""ctypes binding of the CPU oracle (oracle/liboracle_joinhash.so). TEST INFRASTRUCTURE: imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs — never by the product package."""
import ctypes as C
import os as _os
import sys as _sys

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
if _ROOT not in _sys.path:
    _sys.path.insert(0, _ROOT)
import importlib
import os
import subprocess

import numpy as np

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle_joinhash.so")


class OracleResult(C.Structure):
    _fields_ = [("build_side", C.c_void_p), ("probe_side", C.c_void_p), ("n_rows", C.c_uint64),
                ("radix_bits", C.c_uint32), ("reserved", C.c_uint32),
                ("n_build_tuples", C.c_uint64), ("n_probe_tuples", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(ORACLE_DIR, "join_hash_oracle.cpp")
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle_joinhash.so"])
        l = C.CDLL(LIB)
        l.oracle_join.restype = C.c_int
        l.oracle_join.argtypes = [C.POINTER(abi.Relation), C.POINTER(abi.Relation), C.c_int, C.c_int, C.c_uint32,
                                  C.c_int, C.POINTER(OracleResult)]
        l.oracle_result_free.argtypes = [C.POINTER(OracleResult)]
        l.oracle_murmur_hash2.restype = C.c_uint32
        l.oracle_murmur_hash2.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        l.oracle_reference_radix_bits.restype = C.c_int
        l.oracle_reference_radix_bits.argtypes = [C.c_uint64, C.c_int]
        l.oracle_last_error.restype = C.c_char_p
        _lib = l
    return _lib


def oracle_join(build, probe, mode, radix_bits=-1, flags=0, n_threads=1):
    """Returns (build_side (n,2) uint32 or None, probe_side (n,2) uint32, info dict)."""
    res = OracleResult()
    st = lib().oracle_join(C.byref(build), C.byref(probe), mode, radix_bits, flags, n_threads, C.byref(res))
    if st != 0:
        raise RuntimeError(f"oracle_join status {st}: {lib().oracle_last_error().decode()}")
    n = res.n_rows
    probe_side = np.ctypeslib.as_array(C.cast(res.probe_side, C.POINTER(C.c_uint32)), shape=(max(n, 1), 2))[:n].copy()
    build_side = None
    if res.build_side:
        build_side = np.ctypeslib.as_array(C.cast(res.build_side, C.POINTER(C.c_uint32)), shape=(max(n, 1), 2))[:n].copy()
    info = dict(n_rows=n, radix_bits=res.radix_bits, n_build_tuples=res.n_build_tuples,
                n_probe_tuples=res.n_probe_tuples)
    lib().oracle_result_free(C.byref(res))
    return build_side, probe_side, info


def engine(radix_bits=-1, flags=0, n_threads=1):
    """Join engine for tests/hyrise_model.join_hash."""
    def run(build, probe, mode):
        b, p, _ = oracle_join(build, probe, mode, radix_bits, flags, n_threads)
        return b, p
    return run


def canonical(build_side, probe_side):
    """Canonical comparison form: the multiset of result rows as a lexicographically sorted (n,4|2) array."""
    if build_side is None:
        a = np.ascontiguousarray(probe_side, dtype=np.uint32).reshape(-1, 2)
    else:
        a = np.concatenate([np.asarray(build_side, np.uint32).reshape(-1, 2),
                            np.asarray(probe_side, np.uint32).reshape(-1, 2)], axis=1)
    if len(a) == 0:
        return a
    keys = tuple(a[:, c] for c in reversed(range(a.shape[1])))
    return a[np.lexsort(keys)]
