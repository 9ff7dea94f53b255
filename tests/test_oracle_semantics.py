"""A second, independent pin of the CPU oracle (CPU only): the RowID-pair multiset the reference's JoinHash DEFINES, written
down as a few lines of dictionary lookups, against what oracle/join_hash_oracle.cpp — the line-by-line restatement with the
reference's radix partitioning, hash tables and task structure — actually returns.

The golden fixtures (tests/test_oracle_golden.py) pin the oracle on the reference's own expected TABLES, i.e. on values; the
RowIDs behind them were pinned only through the restatement itself. This file closes that: the
semantics below follow SURVEY.md §8a (modes table and quirks Q4 / Q5, with the reference lines they come from) and share
no code with the oracle:

  Inner        (b, p) for every pair of non-NULL equal keys                                  join_hash.cpp:362-431
  Left / Right the same + (NULL_ROW_ID, p) for every probe row without a partner; build NULLs are dropped, a NULL probe
               key is LOOKED UP as what materialisation left in its slot — the stored value of a ValueColumn, T{} = 0 of a
               DictionaryColumn (quirk Q4, :253, :409) — or matches nothing with HJ_FLAG_SQL_NULL_PROBE
  Semi / Anti  every non-NULL probe row with / without a partner, once; NULL-key probe rows never appear (Q5, :468-527)

The engine-level call has no side selection (A0 is the operator's job, tests/hyrise_model.py): `probe` is the outer side."""
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

import oracle_binding as ob
import workloads as wl

NULL_ROW = (0xFFFFFFFF, 0xFFFFFFFF)


def row_ids(n, chunk):
    return [(i // chunk, i % chunk) for i in range(n)]


def defined_result(bk, bn, pk, pn, chunk, mode, encoding, sql_null_probe):
    """Sorted list of result rows — (build chunk, build offset, probe chunk, probe offset), or (probe chunk, probe offset)
    for Semi / Anti — straight from the definition above."""
    partners = {}
    for rid, k, z in zip(row_ids(len(bk), chunk), bk.tolist(), bn.tolist()):
        if not z:
            partners.setdefault(k, []).append(rid)
    out = []
    outer = mode in ("Left", "Right")
    for c, (rid, k, z) in enumerate(zip(row_ids(len(pk), chunk), pk.tolist(), pn.tolist())):
        if z:
            if not outer:
                continue                                   # Inner: no partner; Semi / Anti: dropped at materialisation (Q5)
            if sql_null_probe:
                found = []
            else:                                          # Q4: the slot's content is looked up
                slot = k if encoding == "value" or (encoding == "mixed" and (c // chunk) % 2 == 0) else 0
                found = partners.get(slot, [])
        else:
            found = partners.get(k, [])
        if mode == "Semi":
            if found:
                out.append(rid)
        elif mode == "Anti":
            if not found:
                out.append(rid)
        else:
            out += [b + rid for b in found]
            if outer and not found:
                out.append(NULL_ROW + rid)
    return sorted(out)


def oracle_result(build, probe, mode, **kw):
    b, p, _ = ob.oracle_join(build, probe, abi.MODES[mode], **kw)
    return [tuple(int(x) for x in row) for row in ob.canonical(b, p)]


@pytest.mark.parametrize("mode", ["Inner", "Left", "Right", "Semi", "Anti"])
@pytest.mark.parametrize("encoding", ["value", "dict", "mixed"])
@pytest.mark.parametrize("nullable", [False, True])
def test_oracle_returns_the_defined_rowid_pairs(mode, encoding, nullable):
    rng = np.random.default_rng(sum(map(ord, mode + encoding)) * 2 + int(nullable))   # fixed per case, independent of PYTHONHASHSEED
    chunk = 37
    for n_build, n_probe, domain, lo in ((0, 50, 10, 1), (60, 0, 10, 1), (1, 1, 1, 1), (200, 700, 60, 0), (500, 300, 2000, -5),
                                         (300, 900, 5, 0)):
        bk = rng.integers(lo, lo + domain, n_build).astype(np.int32)
        pk = rng.integers(lo, lo + domain, n_probe).astype(np.int32)
        bn = (rng.random(n_build) < 0.15).astype(np.uint8) if nullable else np.zeros(n_build, np.uint8)
        pn = (rng.random(n_probe) < 0.15).astype(np.uint8) if nullable else np.zeros(n_probe, np.uint8)
        # what sits in a NULL row's slot matters (Q4): junk that also occurs as a build key, and zero
        bk[bn != 0] = rng.integers(lo, lo + domain, int(bn.sum()))
        pk[pn != 0] = np.where(rng.random(int(pn.sum())) < 0.5, 0, rng.integers(lo, lo + domain, int(pn.sum())))
        build = wl.make_relation(bk, chunk, encoding, bn if nullable else None)
        probe = wl.make_relation(pk, chunk, encoding, pn if nullable else None)
        for flags in ((0, abi.FLAG_SQL_NULL_PROBE) if nullable and mode in ("Left", "Right") else (0,)):
            want = defined_result(bk, bn, pk, pn, chunk, mode, encoding, bool(flags))
            for radix_bits, n_threads in ((-1, 1), (0, 1), (4, 3), (9, 2)):
                got = oracle_result(build, probe, mode, radix_bits=radix_bits, flags=flags, n_threads=n_threads)
                assert got == want, (mode, encoding, nullable, n_build, n_probe, flags, radix_bits,
                                     [r for r in got if r not in set(want)][:5], [r for r in want if r not in set(got)][:5])


def test_int64_keys_and_duplicates_on_both_sides():
    rng = np.random.default_rng(64)
    chunk = 50
    bk = (rng.integers(0, 40, 400).astype(np.int64) << 33) + 7      # equal low words, different high words
    pk = (rng.integers(0, 50, 1000).astype(np.int64) << 33) + 7
    build, probe = wl.make_relation(bk, chunk, "value"), wl.make_relation(pk, chunk, "dict")
    for mode in ("Inner", "Left", "Semi", "Anti"):
        want = defined_result(bk, np.zeros(len(bk), np.uint8), pk, np.zeros(len(pk), np.uint8), chunk, mode, "value", False)
        assert oracle_result(build, probe, mode, radix_bits=5, n_threads=2) == want, mode
