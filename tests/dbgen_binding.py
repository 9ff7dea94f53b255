"""ctypes binding of oracle/_ref/libref_dbgen.so = the reference's own tpch-dbgen (third_party/tpch-dbgen) compiled in
place by oracle/Makefile, driven like TpchDbGenerator::generate() (src/benchmarklib/tpch/tpch_db_generator.cpp:203-237).
TEST INFRASTRUCTURE: gives tests and bench.py's C1 workload the key columns "as dbgen makes them". The library is built
in the container that has /root/reference and travels to the GPU box with the snapshot; available() says whether it is
there."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libref_dbgen.so")
_lib = None


def available():
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(LIB)
        l.ref_dbgen_counts.argtypes = [C.c_float, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        l.ref_dbgen_counts.restype = None
        l.ref_dbgen_keys.argtypes = [C.c_float] + [C.c_void_p] * 5 + [C.c_int64, C.POINTER(C.c_int64)]
        l.ref_dbgen_keys.restype = C.c_int
        l.ref_dbgen_o_custkey.argtypes = [C.c_float, C.c_void_p]
        l.ref_dbgen_o_custkey.restype = C.c_int
        _lib = l
    return _lib


def tpch_keys(scale):
    """dict(c_custkey, o_orderkey, o_custkey, o_lines, l_orderkey) as int32 arrays at TPC-H scale factor `scale`.
    dbgen keeps its random streams in globals: call from one thread at a time."""
    l = lib()
    nc, no = C.c_int64(), C.c_int64()
    l.ref_dbgen_counts(scale, C.byref(nc), C.byref(no))
    c = np.zeros(nc.value, np.int32)
    ok, oc, ol = (np.zeros(no.value, np.int32) for _ in range(3))
    cap = no.value * 7  # O_LCNT_MAX lineitems per order (dss.h)
    lk = np.zeros(cap, np.int32)
    nl = C.c_int64()
    rc = l.ref_dbgen_keys(scale, c.ctypes.data, ok.ctypes.data, oc.ctypes.data, ol.ctypes.data, lk.ctypes.data, cap, C.byref(nl))
    if rc != 0:
        raise RuntimeError(f"ref_dbgen_keys: {nl.value} lineitems do not fit {cap}")
    return dict(c_custkey=c, o_orderkey=ok, o_custkey=oc, o_lines=ol, l_orderkey=lk[:nl.value].copy())


def tpch_custkeys(scale):
    """(c_custkey, o_custkey) alone — the C4 join columns — at scale factors the full generator is too slow for: c_custkey
    is 1..n (mk_cust(i + 1)), o_custkey the reference's own draw + CUST_MORTALITY loop on random stream O_CKEY_SD
    (ref_dbgen_o_custkey; equal to mk_order's column, tests/test_dbgen_keys.py)."""
    l = lib()
    nc, no = C.c_int64(), C.c_int64()
    l.ref_dbgen_counts(scale, C.byref(nc), C.byref(no))
    oc = np.zeros(no.value, np.int32)
    if l.ref_dbgen_o_custkey(scale, oc.ctypes.data) != 0:
        raise RuntimeError("ref_dbgen_o_custkey: scale factor out of range")
    return np.arange(1, nc.value + 1, dtype=np.int32), oc
