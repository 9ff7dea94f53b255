"""CPU-only checks of the drop-in boundary: libhj.so loads, exports every symbol include/hj.h declares, the ctypes
mirror matches the compiled struct layouts, and the library fails loudly (no CPU fallback) without a CUDA device."""
import ctypes as C
import importlib
import os
import re

import pytest

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_header_symbols_all_exported(built):
    with open(os.path.join(ROOT, "include", "hj.h")) as f:
        header = f.read()
    declared = set(re.findall(r"^\s*(?:const\s+)?[A-Za-z_][\w\s\*]*?\b(hj_[a-z_0-9]+)\s*\(", header, re.M))
    assert declared, "no declarations parsed from hj.h"
    lib = C.CDLL(hj.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"libhj.so does not export {name}"
    # and the Python binding covers exactly the declared surface
    assert declared == set(abi.SYMBOLS), declared ^ set(abi.SYMBOLS)


def test_struct_layouts_match(built):
    l = hj.lib()
    assert l.hj_abi_version() == abi.HJ_ABI_VERSION
    for which, struct in enumerate([abi.Segment, abi.Relation, abi.Options, abi.ResultInfo, abi.Timings, abi.RowID]):
        assert l.hj_abi_sizeof(which) == C.sizeof(struct), struct.__name__
    assert C.sizeof(abi.RowID) == 8  # types.hpp:97-99


def test_radix_bits_heuristic(built):
    l = hj.lib()
    # sized for a 228 KB shared-memory table, not the reference's 256 KB L2 (join_hash.cpp:654-668)
    assert l.hj_radix_bits(0, abi.KEY_INT32) == 0
    assert l.hj_radix_bits(1000, abi.KEY_INT32) == 0
    assert l.hj_radix_bits(1_500_000, abi.KEY_INT32) == 8     # C1
    assert l.hj_radix_bits(100_000_000, abi.KEY_INT32) == 14  # C2
    assert l.hj_radix_bits(200_000_000, abi.KEY_INT64) == 15  # C3 (capped)
    for n in (1, 10**4, 10**6, 10**8):
        assert l.hj_radix_bits(n, abi.KEY_INT64) >= l.hj_radix_bits(n, abi.KEY_INT32)


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_gpu(built):
    with pytest.raises(hj.HjError) as e:
        hj.HashJoin(device=0)
    assert e.value.status == abi.ERR_CUDA
    assert "no CPU fallback" in str(e.value)


def test_product_package_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "hyrise-1_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                with open(os.path.join(dirpath, f)) as fh:
                    text = fh.read()
                assert "oracle_binding" not in text and "liboracle" not in text and "oracle/" not in text, f
