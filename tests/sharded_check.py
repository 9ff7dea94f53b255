#!/usr/bin/env python3
"""Multi-rank check of the sharded join, run under torchrun with WORLD_SIZE >= 2:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/sharded_check.py --backend nccl|gloo

Every rank holds a shard (a contiguous run of chunks, GLOBAL chunk ids) of the same seeded global relations.
  --backend nccl : the product path — hj_join_sharded on one GPU per rank (exchange kernels + NCCL all-to-all + local join).
  --backend gloo : CPU model of the same protocol — rows are routed with the library's own host function
                   hj_exchange_rank_of, exchanged with gloo, and each rank joins what it received with the oracle. This
                   pins the host-side logic (partition function, NULL routing, "per-rank results concatenate") without a GPU.
In both cases rank 0 gathers all per-rank results and compares the sorted multiset with the oracle's join of the GLOBAL
relations, for every mode. Exit code 0 = all equal."""
import argparse
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np
import torch
import torch.distributed as dist

import oracle_binding as ob
import workloads as wl

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
MODES = ["Inner", "Left", "Right", "Semi", "Anti"]


def global_case(seed, n_build, n_probe, domain, dtype, nullable):
    rng = np.random.default_rng(seed)
    bk = wl.random_keys(rng, n_build, domain, dtype)
    pk = wl.random_keys(rng, n_probe, domain, dtype)
    bn = pn = None
    if nullable:
        bn = (rng.random(n_build) < 0.05).astype(np.uint8)
        pn = (rng.random(n_probe) < 0.05).astype(np.uint8)
    return bk, bn, pk, pn


def shard_chunks(n_rows, chunk, rank, world):
    """Chunk ranges [(chunk_id, start, stop)] of this rank: chunks are dealt out in contiguous runs."""
    n_chunks = (n_rows + chunk - 1) // chunk
    per = (n_chunks + world - 1) // world
    return [(c, c * chunk, min(n_rows, (c + 1) * chunk)) for c in range(rank * per, min(n_chunks, (rank + 1) * per))]


def make_shard(keys, nulls, chunk, encoding, rank, world):
    ids = shard_chunks(len(keys), chunk, rank, world)
    if not ids:
        return wl.make_relation(keys[:0], chunk, encoding, None if nulls is None else nulls[:0])
    lo, hi = ids[0][1], ids[-1][2]
    return wl.make_relation(keys[lo:hi], chunk, encoding, None if nulls is None else nulls[lo:hi],
                            chunk_ids=[c for c, _, _ in ids])


def model_join(rank, world, bk, bn, pk, pn, chunk, encoding, mode, wide):
    """CPU model of hj_join_sharded for this rank: route my shard's rows, exchange with gloo, oracle-join what arrived."""
    outer = mode in ("Left", "Right")
    key_type = abi.KEY_INT64 if wide else abi.KEY_INT32
    received = []
    for keys, nulls, keep_nulls in ((bk, bn, False), (pk, pn, outer)):
        outbox = [[] for _ in range(world)]  # per destination: (key, chunk_id, offset)
        for c, lo, hi in shard_chunks(len(keys), chunk, rank, world):
            for i in range(lo, hi):
                is_null = nulls is not None and nulls[i] != 0
                if is_null and not keep_nulls:
                    continue  # kNullDrop (join_hash.cpp:253)
                # kNullAsValue (quirk Q4): a NULL probe key of an outer join travels as its stored value (T{} = 0 for
                # dictionary/gathered columns) and is looked up like any other key
                k = int(keys[i]) if not (is_null and encoding in ("dict", "gathered")) else 0
                outbox[hj.exchange_rank_of(k, key_type, world)].append((k, c, i - lo))
        gathered = [None] * world
        dist.all_gather_object(gathered, outbox)
        mine = [row for src in range(world) for row in gathered[src][rank]]
        received.append(mine)

    def as_relation(rows, dtype):
        rb = hj.RelationBuilder(abi.LOC_HOST)
        rows = sorted(rows, key=lambda r: (r[1], r[2]))
        by_chunk = {}
        for k, c, o in rows:
            by_chunk.setdefault(c, []).append((k, o))
        for c in sorted(by_chunk):
            ks = np.array([k for k, _ in by_chunk[c]], dtype=dtype)
            offs = np.array([o for _, o in by_chunk[c]], dtype=np.uint32)
            rb.add_value(c, ks, None, kind=abi.SEG_GATHERED, row_offsets=offs)
        return rb.build()

    b_rel, p_rel = as_relation(received[0], bk.dtype), as_relation(received[1], pk.dtype)
    b, p, _ = ob.oracle_join(b_rel, p_rel, abi.MODES[mode])
    return b, p


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="nccl", choices=["nccl", "gloo"])
    ap.add_argument("--big", action="store_true", help="nccl only: add a 2 M x 8 M case")
    ap.add_argument("--staged", action="store_true", help="nccl only: HJ_FLAG_XCHG_STAGED (send buffers + ncclSend/Recv)")
    ap.add_argument("--slice-rows", type=int, default=0,
                    help="nccl only: hj_options.stream_slice_rows — small values force the sliced exchange (probe slices "
                         "exchanged while the previous slice is joined) on these small cases")
    ap.add_argument("--radix-bits", type=int, default=-1,
                    help="nccl only: hj_options.radix_bits; > 8 makes the local join two-pass, which lets shards with regular "
                         "chunks (2^k rows) take the PUSH exchange (hj_xpush.cuh) on these small cases")
    ap.add_argument("--bound", action="store_true",
                    help="nccl only: hj_bind_output before every join (host outputs); with --radix-bits > 8 and a small --slice-rows "
                         "the regular shards take the STREAMED push exchange (probe slices: copy-in, exchange + join, copy-out overlap)")
    ap.add_argument("--general", action="store_true", help="nccl only: HJ_FLAG_XCHG_GENERAL (never the push exchange)")
    ap.add_argument("--c5", type=float, default=0.0,
                    help="nccl only: add config C5 at this scale (0.01 = 40 M x 40 M int64, permutation build x uniform probe, "
                         "ValueColumns of 2^20-row chunks), Inner, compared bit-exactly with the oracle's join of the global relations")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    if args.backend == "nccl":
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        dist.init_process_group("gloo")

    j = None
    if args.backend == "nccl":
        j = hj.HashJoin(device=local_rank, flags=(abi.FLAG_XCHG_STAGED if args.staged else 0) | (abi.FLAG_XCHG_GENERAL if args.general else 0),
                        stream_slice_rows=args.slice_rows, radix_bits=args.radix_bits)
        uid = [hj.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        j.comm_init(uid[0], world, rank)

    cases = [  # (name, n_build, n_probe, domain, dtype, nullable, encoding, chunk)
        ("int32 value", 3000, 9000, 2500, np.int32, True, "value", 700),
        ("int32 dict", 4000, 12000, 3000, np.int32, True, "dict", 1000),
        ("int64 gathered", 2500, 6000, 2000, np.int64, True, "gathered", 512),
        ("empty build", 0, 500, 100, np.int32, False, "value", 128),
    ]
    if args.backend == "nccl":
        cases += [  # regular shards (chunks of 2^k rows): the push exchange when the local join is two-pass (--radix-bits > 8)
            ("int32 dict 2^k", 40_000, 130_000, 30_000, np.int32, True, "dict", 4096),
            ("int32 value 2^k", 50_000, 90_000, 45_000, np.int32, True, "value", 8192),
            ("int64 value 2^k", 30_000, 100_000, 20_000, np.int64, True, "value", 2048),
            ("int64 mixed 2^k", 25_000, 60_000, 1_000, np.int64, False, "mixed", 1024),
            ("empty probe 2^k", 3000, 0, 100, np.int32, False, "value", 1024),
            ("one chunk each", 900, 1500, 500, np.int32, True, "dict", 1024),
        ]
    if args.backend == "gloo":
        cases = [(n, b // 4, p // 4, max(1, d // 4), t, nu, e, max(16, c // 4)) for n, b, p, d, t, nu, e, c in cases]
    elif args.big:
        cases.append(("2M x 8M dict", 2_000_000, 8_000_000, 2_000_000, np.int32, False, "dict", 1 << 18))

    if args.backend == "nccl" and args.c5 > 0:
        n5 = int(4_000_000_000 * args.c5)
        cases.append(("C5 x %g" % args.c5, n5, n5, n5, np.int64, False, "value", 1 << 20))

    failures = 0
    paths = {0: "-", 1: "general", 2: "push"}
    for ci, (name, nb, npr, domain, dtype, nullable, encoding, chunk) in enumerate(cases):
        if name.startswith("C5"):
            rng = np.random.default_rng(0xB205)
            bk, bn, pk, pn = (rng.permutation(nb) + 1).astype(np.int64), None, rng.integers(1, nb + 1, npr, dtype=np.int64), None
        else:
            bk, bn, pk, pn = global_case(1000 + ci, nb, npr, domain, dtype, nullable)
        wide = dtype == np.int64
        for mode in (["Inner"] if name.startswith("C5") else MODES):
            if args.backend == "nccl":
                b_rel = make_shard(bk, bn, chunk, encoding, rank, world)
                p_rel = make_shard(pk, pn, chunk, encoding, rank, world)
                if args.bound:
                    cap = 10_000_000
                    out_b = None if mode in ("Semi", "Anti") else np.empty((cap, 2), np.uint32)
                    out_p = np.empty((cap, 2), np.uint32)
                    j.bind_output(out_b, out_p)
                    info = j.join_sharded(b_rel, p_rel, mode)
                    b, p = j.fetch(out_b, out_p)
                else:
                    info = j.join_sharded(b_rel, p_rel, mode)
                    b, p = j.fetch()
            else:
                b, p = model_join(rank, world, bk, bn, pk, pn, chunk, encoding, mode, wide)
            parts = [None] * world
            dist.all_gather_object(parts, (None if b is None else np.asarray(b).copy(), np.asarray(p).copy()))
            if rank == 0:
                gb = None if parts[0][0] is None else np.concatenate([x[0].reshape(-1, 2) for x in parts])
                gp = np.concatenate([x[1].reshape(-1, 2) for x in parts])
                full_b = wl.make_relation(bk, chunk, encoding, bn)
                full_p = wl.make_relation(pk, chunk, encoding, pn)
                ob_b, ob_p, oinfo = ob.oracle_join(full_b, full_p, abi.MODES[mode])
                got, want = ob.canonical(gb, gp), ob.canonical(ob_b, ob_p)
                ok = got.shape == want.shape and np.array_equal(got, want)
                slices = info.n_probe_slices if args.backend == "nccl" else 0
                path = paths.get(j.timings()["xchg_path"], "?") if args.backend == "nccl" else "model"
                print(f"[{args.backend} x{world}] {name:16s} {mode:5s} rows={len(want):8d} per-rank={[len(x[1]) for x in parts]} "
                      f"slices={slices} exchange={path} {'OK' if ok else 'MISMATCH'}", flush=True)
                failures += 0 if ok else 1
    res = [failures]
    dist.broadcast_object_list(res, src=0)
    if j is not None:
        j.close()
    dist.destroy_process_group()
    return 1 if res[0] else 0


if __name__ == "__main__":
    sys.exit(main())
