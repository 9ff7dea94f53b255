"""Pins the CPU oracle against the reference's own JoinHash expectations (CPU only).

Every in-scope case of join_equi_test.cpp / join_null_test.cpp / join_semi_anti_test.cpp is replayed through the
restated algorithm (oracle/join_hash_oracle.cpp) with the test-side Hyrise model, for several radix-bit settings and
thread counts (the reference's result must not depend on either), and MurmurHash2 is checked against known answers
produced by the reference's own murmur_hash.cpp."""
import ctypes as C
import json
import os
import struct

import pytest

import hyrise_model as hm
import oracle_binding as ob

FX = hm.Fixtures()
IN_SCOPE = [(n, c, ops) for n, c, ops in FX.cases if hm.case_in_scope(ops)]
OUT_OF_SCOPE = [(n, c, ops) for n, c, ops in FX.cases if not hm.case_in_scope(ops)]


def test_fixture_inventory():
    # 53 reference cases; 7 use float/double/string keys (out of v1 scope)
    assert len(FX.cases) == 53
    assert len(OUT_OF_SCOPE) == 7
    assert len(FX.tables) == 48


@pytest.mark.parametrize("name,cite,ops", IN_SCOPE, ids=[c[0] for c in IN_SCOPE])
@pytest.mark.parametrize("radix_bits,n_threads", [(-1, 1), (0, 1), (3, 4), (9, 2)])
def test_reference_case(name, cite, ops, radix_bits, n_threads):
    ok, msg = hm.run_case(FX, ops, ob.engine(radix_bits=radix_bits, n_threads=n_threads))
    assert ok, f"{name} ({cite}): {msg}"


def test_murmur_known_answers():
    with open(os.path.join(hm.GOLDEN, "murmur_kat.json")) as f:
        kat = json.load(f)
    fn = ob.lib().oracle_murmur_hash2
    for k, h in kat["int32"]:
        assert fn(C.create_string_buffer(struct.pack("<i", k), 4), 4, kat["seed"]) == h
    for k, h in kat["int64"]:
        assert fn(C.create_string_buffer(struct.pack("<q", k), 8), 8, kat["seed"]) == h


def test_reference_radix_bits():
    # C1 (1.5 M int32 build rows) -> 9 bits, C2 (100 M) -> 15, C3 (200 M int64) -> 16
    rb = ob.lib().oracle_reference_radix_bits
    assert rb(1_500_000, 0) == 9
    assert rb(100_000_000, 0) == 15
    assert rb(200_000_000, 1) == 16
    assert rb(0, 0) == 0 and rb(10, 0) == 0


def test_simd_bp128_layout_matches_reference():
    """The numpy SimdBp128 packer the GPU tests feed libhj with (workloads.bp128_pack) against the reference's own
    SimdBp128Packing::pack_block / unpack_block, compiled in place into oracle/_ref/libref_bp128.so (all 33 bit widths)."""
    import numpy as np
    import workloads as wl
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    so = os.path.join(root, "oracle", "_ref", "libref_bp128.so")
    if not os.path.exists(so):
        pytest.skip("oracle/_ref/libref_bp128.so not built (needs /root/reference at build time)")
    ref = C.CDLL(so)
    rng = np.random.default_rng(128)
    for width in range(33):
        vals = (rng.integers(0, 2**63, 128, dtype=np.uint64) & np.uint64((1 << width) - 1)).astype(np.uint32)
        if width:
            vals[5] = (1 << width) - 1  # make sure the block needs exactly `width` bits
        words = np.zeros(max(width, 1) * 4 + 4, dtype=np.uint32)
        buf = np.ascontiguousarray(np.concatenate([np.zeros(0, np.uint32), words]))
        aligned = np.zeros(len(buf) + 4, dtype=np.uint32)
        off = (-aligned.ctypes.data // 4) % 4          # 16-byte aligned start for _mm_store_si128
        out = aligned[off:off + max(width, 1) * 4]
        ref.ref_bp128_pack_block(vals.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), C.c_uint8(width))
        mine = wl.bp128_pack(vals)                      # [meta word][width words]
        assert bytes(mine[:4].view(np.uint8))[0] == width and not any(bytes(mine[:4].view(np.uint8))[1:])
        assert np.array_equal(mine[4:], out[:width * 4]), f"width {width}: packed words differ from the reference's"
        back = np.zeros(128, dtype=np.uint32)
        ref.ref_bp128_unpack_block(out.ctypes.data_as(C.c_void_p), back.ctypes.data_as(C.c_void_p), C.c_uint8(width))
        assert np.array_equal(back, vals)
