"""Seeded synthetic join inputs as hj_relation descriptors (host numpy arrays). Shared by tests, smoke() and bench.py.

Encodings follow the reference's storage formats: ValueColumn = values (+ 1 byte/row null flags); DictionaryColumn =
sorted unique dictionary + FixedSizeByteAligned attribute vector, NULL = dictionary size
(storage/dictionary_column/dictionary_encoder.hpp:58-130, fixed_size_byte_aligned_compressor.cpp:21-30)."""
import os as _os
import sys as _sys

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
if _ROOT not in _sys.path:
    _sys.path.insert(0, _ROOT)
import importlib

import numpy as np

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi


def dict_encode(values, nulls=None):
    """(dictionary, attribute vector in the narrowest of u8/u16/u32, null_value_id) for one chunk."""
    valid = np.ones(len(values), bool) if nulls is None else (nulls == 0)
    dictionary = np.unique(values[valid]).astype(values.dtype)
    attr = np.searchsorted(dictionary, values).astype(np.uint32)
    attr[~valid] = len(dictionary)
    max_value = len(dictionary) + 1
    width = np.uint8 if max_value <= 0xFF else (np.uint16 if max_value <= 0xFFFF else np.uint32)
    return dictionary, attr.astype(width), len(dictionary)


def make_relation(keys, chunk, encoding="value", nulls=None, chunk_sizes=None, chunk_ids=None):
    """Cuts `keys` (1-D int32/int64) into chunks and describes them.
    encoding: "value" | "dict" | "mixed" (alternating) | "gathered".
    chunk_sizes overrides the regular `chunk` cut (list summing to len(keys)); chunk_ids overrides 0,1,2,..."""
    rb = hj.RelationBuilder(abi.LOC_HOST)
    n = len(keys)
    if chunk_sizes is None:
        chunk_sizes = [min(chunk, n - s) for s in range(0, n, chunk)]
    start = 0
    for c, sz in enumerate(chunk_sizes):
        v = np.ascontiguousarray(keys[start:start + sz])
        nl = None if nulls is None else np.ascontiguousarray(nulls[start:start + sz])
        cid = c if chunk_ids is None else chunk_ids[c]
        enc = encoding if encoding != "mixed" else ("dict" if c % 2 else "value")
        if enc == "dict":
            d, a, nv = dict_encode(v, nl)
            rb.add_dict(cid, d, a, nv)
        elif enc == "gathered":
            vv = v.copy()
            if nl is not None:
                vv[nl != 0] = 0
            rb.add_value(cid, vv, nl, kind=abi.SEG_GATHERED)
        else:
            rb.add_value(cid, v, nl)
        start += sz
    return rb.build()


def random_keys(rng, n, domain, dtype=np.int32, lo=1):
    return rng.integers(lo, lo + domain, size=n, dtype=np.int64).astype(dtype)


def random_relations(rng, n_build, n_probe, domain, chunk, encoding="value", nullable=False, dtype=np.int32,
                     probe_dtype=None, null_frac=0.05, unique_build=False, lo=1):
    """A build and a probe relation over a shared key domain [lo, lo+domain)."""
    probe_dtype = probe_dtype or dtype
    if unique_build:
        assert domain >= n_build
        bk = (rng.permutation(domain)[:n_build] + lo).astype(dtype)
    else:
        bk = random_keys(rng, n_build, domain, dtype, lo)
    pk = random_keys(rng, n_probe, domain, probe_dtype, lo)
    bn = pn = None
    if nullable:
        bn = (rng.random(n_build) < null_frac).astype(np.uint8)
        pn = (rng.random(n_probe) < null_frac).astype(np.uint8)
    return (make_relation(bk, chunk, encoding, bn), make_relation(pk, chunk, encoding, pn))


def zipf_keys(rng, n, domain, s=1.0, dtype=np.int64):
    """Zipf(s) ranks over 1..domain by inverse-CDF on the harmonic weights, scrambled by a fixed bijection."""
    ranks = np.arange(1, domain + 1, dtype=np.float64)
    cdf = np.cumsum(ranks ** (-s))
    cdf /= cdf[-1]
    r = np.searchsorted(cdf, rng.random(n)).astype(np.int64)  # 0-based rank
    return ((r * 2654435761) % domain + 1).astype(dtype) if np.gcd(2654435761, domain) == 1 else (r + 1).astype(dtype)


def bp128_pack(values):
    """SimdBp128Compressor (simd_bp128_compressor.cpp:33-110) + Pack128Bit (simd_bp128_packing.cpp:30-95): per meta block
    of 2048 values (zero-padded) one 16-byte word of 16 bit widths, then per packed block of 128 values `width` 16-byte
    words; value j sits in 32-bit lane j % 4, at bit (j / 4) * width of the lane's little-endian bit stream."""
    v = np.asarray(values, dtype=np.uint64)
    n = len(v)
    out = []
    for m0 in range(0, n, 2048):
        meta = np.zeros(2048, dtype=np.uint64)
        cnt = min(2048, n - m0)
        meta[:cnt] = v[m0:m0 + cnt]
        widths = [int(np.bitwise_or.reduce(meta[b * 128:(b + 1) * 128])).bit_length() for b in range(16)]
        out.append(np.frombuffer(bytes(widths), dtype=np.uint32))
        for b in range((cnt + 127) // 128):
            w = widths[b]
            if w == 0:
                continue
            words = np.zeros((w, 4), dtype=np.uint32)
            for lane in range(4):
                stream = 0
                for k, x in enumerate(meta[b * 128 + lane:(b + 1) * 128:4]):
                    stream |= int(x) << (k * w)
                for i in range(w):
                    words[i, lane] = (stream >> (32 * i)) & 0xFFFFFFFF
            out.append(words.reshape(-1))
    return np.concatenate(out) if out else np.zeros(4, dtype=np.uint32)
