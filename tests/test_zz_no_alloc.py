"""hj_aggregate_column_static — the entry point that allocates nothing (scratch in a file-scope `static __device__` array of
libhj, no reference counterpart) — against the existing entry point
hj_aggregate_column, through the C ABI, two calls in a row. Collected last (test_zz_*).

Bar: identical results. Both entry points run the same per-CTA kernel on the same grid and fold the partials in the same
order, so even the double sums agree bit for bit; the test still allows the 1e-12 relative tolerance tests/
test_gpu_aggregate.py uses for double sums."""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
pytestmark = pytest.mark.gpu
KEY_TYPES = {"int32": abi.KEY_INT32, "int64": abi.KEY_INT64, "float32": abi.KEY_FLOAT, "float64": abi.KEY_DOUBLE}
SUM_RTOL = 1e-12
MAX_ROWS = 1 << 31  # HJ_STATIC_AGG_MAX_ROWS


@pytest.fixture(scope="module")
def ctx():
    j = hj.HashJoin(device=0)
    yield j
    j.close()


def same(got, want):
    for k in want:
        a, b = got[k], want[k]
        if isinstance(b, float) and b is not None and a is not None:
            assert a == b or abs(a - b) <= SUM_RTOL * max(abs(a), abs(b)), (k, got, want)
        else:
            assert a == b, (k, got, want)


def run_static(ctx, v, z, n, key_type, out):
    import torch
    out.fill_(0x5A)  # stale bytes: the call must overwrite the whole record
    ctx.aggregate_column_static(v.data_ptr() if n else None, None if z is None else z.data_ptr(), n, key_type, out.data_ptr())
    torch.cuda.synchronize()
    return hj.HashJoin.aggregate_record(out.cpu().numpy().tobytes())


@pytest.mark.parametrize("dtype", ["int32", "int64", "float32", "float64"])
def test_static_scratch_aggregate_equals_the_allocating_entry_point_twice_in_a_row(ctx, dtype):
    import torch
    rng = np.random.default_rng(0xA110C + len(dtype))
    out = torch.empty(C.sizeof(abi.Aggregate), dtype=torch.uint8, device="cuda")
    # two calls in a row on DIFFERENT inputs (the second must not see anything of the first in the shared scratch), then
    # the first input again, an all-NULL column and an empty one
    cases = [(1_000_003, True), (300_001, False), (1_000_003, True), (5000, "all"), (0, False)]
    first = None
    for n, nullable in cases:
        sub = np.random.default_rng(n)
        if dtype.startswith("int"):
            host = sub.integers(-(1 << 30), 1 << 30, n).astype(dtype)
        else:
            host = (sub.standard_normal(n) * 1e3).astype(dtype)
        nulls = None
        if nullable == "all":
            nulls = np.ones(n, np.uint8)
        elif nullable:
            nulls = (sub.random(n) < 0.1).astype(np.uint8)
        v = torch.from_numpy(host).cuda()
        z = None if nulls is None else torch.from_numpy(nulls).cuda()
        want = ctx.aggregate_column(v.data_ptr() if n else None, None if z is None else z.data_ptr(), n, KEY_TYPES[dtype])
        got = run_static(ctx, v, z, n, KEY_TYPES[dtype], out)
        same(got, want)
        again = run_static(ctx, v, z, n, KEY_TYPES[dtype], out)  # back to back on the same stream: same answer
        assert again == got
        if n == 1_000_003:
            if first is None:
                first = got
            else:
                assert got == first, "the shared scratch leaked state between calls"
        # and against numpy for the exact parts
        valid = np.ones(n, bool) if nulls is None else nulls == 0
        assert got["count_rows"] == n and got["count"] == int(valid.sum())
        if dtype.startswith("int") and valid.any():
            assert got["sum"] == int(host[valid].astype(np.int64).sum()) and got["min"] == int(host[valid].min()) and got["max"] == int(host[valid].max())
    del rng


def test_static_scratch_aggregate_rejects_sizes_beyond_its_maximum(ctx):
    import torch
    v = torch.zeros(16, dtype=torch.int32, device="cuda")
    out = torch.zeros(C.sizeof(abi.Aggregate), dtype=torch.uint8, device="cuda")
    with pytest.raises(hj.HjError) as e:
        ctx.aggregate_column_static(v.data_ptr(), None, MAX_ROWS + 1, abi.KEY_INT32, out.data_ptr())  # refused before any launch
    assert e.value.status == abi.ERR_TOO_LARGE and "HJ_STATIC_AGG_MAX_ROWS" in str(e.value)
    torch.cuda.synchronize()
    assert int(out.sum()) == 0, "a refused call must not touch the output"
    with pytest.raises(hj.HjError):
        ctx.aggregate_column_static(v.data_ptr(), None, 16, abi.KEY_INT32, None)
