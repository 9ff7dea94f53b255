#!/usr/bin/env python3
"""bench.py — JoinHash input tuples/s on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c1] [--scale F]

A "step" is one complete hash join (hj_join: materialize+histogram -> radix scatter x2 -> shared-memory build+probe ->
PosLists) over the synthetic workload BASELINE.json's metric is quoted on: C2 = 100 M x 1 B int32 rows, uniform keys,
dictionary-encoded 1 Mi-row chunks, Inner. `value` is measured with the key columns already resident in
HBM; `e2e` runs the same join through the C ABI from pinned HOST buffers (H2D of the columns and D2H of both PosLists
inside the timed region). Inputs (8.8 GB) are far larger than L2, so no L2 flush is needed between steps.
`--impl reference` times the reference's CPU algorithm (the oracle restatement, all host threads) on a bounded sample.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np

CHUNK = 1 << 20  # rows per chunk (1 048 576)
WORKLOADS = {
    # name: (n_build, n_probe, key dtype, encoding, description)
    "c2": (100_000_000, 1_000_000_000, "int32", "dict",
           "C2 synthetic 1:N int32 inner equi-join, 100M build x 1B probe, uniform keys, dictionary-encoded"),
    "c3": (200_000_000, 2_000_000_000, "int64", "value",
           "C3 synthetic int64 join, Zipf(1.0) build keys, 200M x 2B, ValueColumns"),
    "c5": (4_000_000_000, 4_000_000_000, "int64", "value",
           "C5 radix-partitioned 4B x 4B int64 join over 2/4/8 GPUs, permutation build x uniform probe, ValueColumns "
           "(STRONG scaling: the published row counts are the global relations, every rank holds 1/N)"),
    "c1": (1_500_000, 6_001_215, "int32", "dict",
           "C1 TPC-H SF1 lineitem x orders on l_orderkey = o_orderkey (dbgen keys when oracle/_ref/libref_dbgen.so is present, "
           "else 1.5M unique x 6M uniform), dictionary-encoded"),
}


def csrc_hash():
    """sha256 over the CUDA sources of libhj (what profiles/traffic_c2_full.json was captured from)."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "hyrise-1_b200", "csrc")
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh")):
            with open(os.path.join(d, name), "rb") as f:
                h.update(name.encode() + b"\0" + f.read())
    return h.hexdigest()


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------------------ data
STRONG = {"c5"}  # workloads whose published row counts are the GLOBAL relations (each rank holds 1/world of them)


def perm_multiplier(domain):
    """Odd multiplier A coprime to `domain` with x * A < 2^63 for every x < domain: x -> (x * A) mod domain + 1 is the
    global permutation the build keys of a multi-rank run are cut from (every rank can generate its slice on its own)."""
    import math
    a = (2654435761 if domain < (1 << 31) else 1_000_000_007) % domain or 1
    while math.gcd(a, domain) != 1:
        a += 1
    return a


def gen_device_workload(name, scale, device, seed, rank=0, world=1):
    """Builds this rank's shard of the workload's key columns on `device` with torch: `scale` x the published row counts
    PER RANK (weak scaling; STRONG workloads: 1/world of them), keys drawn from the GLOBAL domain of all ranks' build rows,
    chunk ids global. Returns dict with per-relation lists of (chunk_id, dictionary|values tensor, attr tensor|None) and
    row counts."""
    import torch
    n_build, n_probe, dtype, encoding, _ = WORKLOADS[name]
    if name in STRONG:
        n_build, n_probe = n_build // world, n_probe // world
    n_build, n_probe = max(1, int(n_build * scale)), max(1, int(n_probe * scale))
    tdt = torch.int32 if dtype == "int32" else torch.int64
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    domain = n_build * world
    if name == "c3":
        # Zipf(1.0) ranks by inverse CDF of the continuous approximation, scrambled by a fixed odd multiplier
        u = torch.rand(n_build, device=device, generator=g, dtype=torch.float64)
        ranks = torch.exp(u * np.log(domain + 1.0)).floor().clamp_(1, domain).to(torch.int64)
        build_keys = ((ranks * 0x9E3779B1) % domain + 1).to(tdt)
        del u, ranks
    elif world == 1 and name not in STRONG:
        build_keys = (torch.randperm(domain, device=device, generator=g) + 1).to(tdt)  # unique: 1:N join
    else:
        # this rank's slice of ONE global permutation of 1..domain: x -> (x * A) mod domain with gcd(A, domain) = 1
        a = perm_multiplier(domain)
        x = torch.arange(rank * n_build, (rank + 1) * n_build, device=device, dtype=torch.int64)
        build_keys = ((x * a) % domain + 1).to(tdt)
        del x

    def chunks_of(keys_fn, n):
        out = []
        n_chunks = (n + CHUNK - 1) // CHUNK
        for c, start in enumerate(range(0, n, CHUNK)):
            k = keys_fn(start, min(CHUNK, n - start))
            cid = rank * n_chunks + c
            if encoding == "dict":
                d, inv = torch.unique(k, sorted=True, return_inverse=True)
                # FixedSizeByteAligned width rule (fixed_size_byte_aligned_compressor.cpp:21-30); C2's 1 Mi-row chunks
                # hold ~1 M distinct keys -> 4-byte attribute vectors
                m = len(d) + 1
                adt = torch.uint8 if m <= 0xFF else (torch.int16 if m <= 0xFFFF else torch.int32)
                out.append((cid, d.contiguous(), inv.to(adt).contiguous()))
            else:
                out.append((cid, k.contiguous(), None))
        return out

    if name == "c1" and world == 1 and scale == 1.0:
        # C1 on the key columns the reference's own dbgen generates at SF1 (oracle/_ref/libref_dbgen.so, test
        # infrastructure built where /root/reference exists; it only GENERATES the input here). Absent: C1-shaped synthetic.
        import dbgen_binding as db
        if db.available():
            k = db.tpch_keys(1.0)
            okeys = torch.from_numpy(k["o_orderkey"]).to(device)
            lkeys = torch.from_numpy(k["l_orderkey"]).to(device)
            return dict(build=chunks_of(lambda s, m: okeys[s:s + m], len(okeys)),
                        probe=chunks_of(lambda s, m: lkeys[s:s + m], len(lkeys)), n_build=len(okeys), n_probe=len(lkeys),
                        dtype=dtype, encoding=encoding, source="dbgen (reference's tpch-dbgen, SF1: orders x lineitem keys)")
    build = chunks_of(lambda s, m: build_keys[s:s + m], n_build)
    probe = chunks_of(lambda s, m: torch.randint(1, domain + 1, (m,), device=device, generator=g, dtype=torch.int64).to(tdt),
                      n_probe)
    del build_keys
    return dict(build=build, probe=probe, n_build=n_build, n_probe=n_probe, dtype=dtype, encoding=encoding)


def relation_from_chunks(hj, chunks, location, attr_small=False):
    rb = hj.RelationBuilder(location)
    for cid, a, b in chunks:
        if b is None:
            rb.add_value(cid, a)
        else:
            rb.add_dict(cid, a, b)
    return rb.build()


def to_pinned_host(chunks):
    """Pinned host copies of the device chunks (what a host-resident Hyrise table would hand the shim)."""
    import torch
    out = []
    for cid, a, b in chunks:
        ha = torch.empty(a.shape, dtype=a.dtype, pin_memory=True).copy_(a)
        hb = None if b is None else torch.empty(b.shape, dtype=b.dtype, pin_memory=True).copy_(b)
        out.append((cid, ha, hb))
    return out


def input_bytes(chunks):
    return sum(a.numel() * a.element_size() + (0 if b is None else b.numel() * b.element_size()) for _, a, b in chunks)


# ------------------------------------------------------------------------------------------------------------ verify
class _DevArray:
    """A raw device pointer as something torch.as_tensor understands (no copy)."""
    def __init__(self, ptr, n, typestr="<i8"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def _mulmod(k, m, d):
    """(k * m) mod d for int64 tensors k < d <= 2^32 and a Python int m < d, without leaving 63 bits."""
    hi, lo = m >> 16, m & 0xFFFF
    return (((k * hi) % d) * 65536 + k * lo) % d


def _chunk_keys(chunk):
    """Decoded keys of one (chunk_id, values | dictionary, attr | None) chunk as int64."""
    import torch
    _, a, at = chunk
    if at is None:
        return a.to(torch.int64)
    mask = {torch.int32: 0xFFFFFFFF, torch.int16: 0xFFFF, torch.uint8: 0xFF}[at.dtype]
    return a[at.long() & mask].to(torch.int64)


def _keys_at(chunks, local_pos):
    """Keys at local row positions (chunk index * CHUNK + offset) of a chunk list, without concatenating the column."""
    import torch
    out = torch.empty_like(local_pos)
    c = local_pos // CHUNK
    order = torch.argsort(c)
    counts = torch.bincount(c, minlength=len(chunks)).tolist()
    start = 0
    for ci, n in enumerate(counts):
        if n:
            sel = order[start:start + n]
            out[sel] = _chunk_keys(chunks[ci])[local_pos[sel] % CHUNK]
            start += n
    return out


def verify_result(j, data, info, name, rank, world, dev, sample=1_000_000):
    """Checks the PosLists the last step left in HBM (hj_result_device; nothing is fetched) against values computed
    independently with torch, reduced over the ranks: exact global row count, probe-side and build-side RowID checksums
    (sum of chunk_id * 2^20 + offset over all result rows, mod 2^64) and key equality of sampled result rows (across
    ranks: the sampled RowIDs are sent to the rank that holds the probe row). Inner joins only. Works chunk by chunk so
    that it fits next to the join's buffers at the full size of C3 / C5."""
    import torch
    import torch.distributed as dist
    nb, npr = data["n_build"], data["n_probe"]
    nc_b, nc_p = (nb + CHUNK - 1) // CHUNK, (npr + CHUNK - 1) // CHUNK
    bptr, pptr, n = j.result_device()
    rb = torch.as_tensor(_DevArray(bptr, n), device=dev) if n else torch.zeros(0, dtype=torch.int64, device=dev)
    rp = torch.as_tensor(_DevArray(pptr, n), device=dev) if n else torch.zeros(0, dtype=torch.int64, device=dev)
    return verify_poslists(rb, rp, data, name, rank, world, dev, sample)


def verify_poslists(rb, rp, data, name, rank, world, dev, sample=1_000_000):
    """verify_result over the two PosLists as int64 tensors (packed RowIDs: chunk_id | offset << 32)."""
    import torch
    import torch.distributed as dist
    n = len(rp)
    nb, npr = data["n_build"], data["n_probe"]
    nc_b, nc_p = (nb + CHUNK - 1) // CHUNK, (npr + CHUNK - 1) // CHUNK

    def pos_sum(r):  # sum of chunk_id * CHUNK + offset, in slices (chunk ids are global: rank * chunks_per_rank + c)
        tot = torch.zeros((), dtype=torch.int64, device=dev)
        for lo in range(0, len(r), 1 << 28):
            x = r[lo:lo + (1 << 28)]
            tot += ((x & 0xFFFFFFFF) * CHUNK + (x >> 32)).sum()
        return tot
    got = torch.stack([torch.tensor(n, device=dev), pos_sum(rb), pos_sum(rp)])
    permutation = name != "c3"
    zero = lambda: torch.zeros((), dtype=torch.int64, device=dev)
    if permutation and (world > 1 or name in STRONG):
        # build key of global row x is (x * A) mod D + 1: the build row a probe key joins with is x = (key - 1) * A^-1 mod D
        domain = nb * world
        ainv = pow(perm_multiplier(domain), -1, domain)
        want_b, want_p = zero(), zero()
        for ci, ch in enumerate(data["probe"]):
            pk = _chunk_keys(ch)
            x = _mulmod(pk - 1, ainv, domain)
            want_b += ((x // nb * nc_b + x % nb // CHUNK) * CHUNK + x % nb % CHUNK).sum()
            want_p += (rank * nc_p + ci) * CHUNK * len(pk) + len(pk) * (len(pk) - 1) // 2
        want = torch.stack([torch.tensor(npr, device=dev), want_b, want_p])
    elif world == 1:
        domain = int(max(max(int(_chunk_keys(ch).max()) for ch in data["build"]), max(int(_chunk_keys(ch).max()) for ch in data["probe"]))) + 1
        cb = torch.zeros(domain, dtype=torch.int64, device=dev)
        cp = torch.zeros(domain, dtype=torch.int64, device=dev)
        for ch in data["build"]:
            k = _chunk_keys(ch)
            cb.index_add_(0, k, torch.ones_like(k))
        for ch in data["probe"]:
            k = _chunk_keys(ch)
            cp.index_add_(0, k, torch.ones_like(k))
        want_n, want_b, want_p = zero(), zero(), zero()
        for ci, ch in enumerate(data["probe"]):
            k = _chunk_keys(ch)
            m = cb[k]  # matches per probe row
            want_n += m.sum()
            want_p += ((ci * CHUNK + torch.arange(len(k), device=dev)) * m).sum()
        for ci, ch in enumerate(data["build"]):
            k = _chunk_keys(ch)
            want_b += ((ci * CHUNK + torch.arange(len(k), device=dev)) * cp[k]).sum()
        want = torch.stack([want_n, want_b, want_p])
        del cb, cp
    else:
        return None
    # sampled rows: the keys behind both RowIDs must be equal
    k = min(sample, n)
    bad = torch.zeros(1, dtype=torch.int64, device=dev)
    lpos = lambda r, nc: ((r & 0xFFFFFFFF) - rank * nc) * CHUNK + (r >> 32)
    if world == 1:
        if k:
            idx = torch.randint(0, n, (k,), device=dev)
            bad += (_keys_at(data["build"], lpos(rb[idx], nc_b)) != _keys_at(data["probe"], lpos(rp[idx], nc_p))).sum()
    else:
        idx = torch.randint(0, max(n, 1), (sample,), device=dev)
        smp = torch.full((sample, 2), -1, dtype=torch.int64, device=dev)  # (probe RowID, key the build RowID stands for)
        if n and permutation:
            xb = rb[idx]
            xg = ((xb & 0xFFFFFFFF) // nc_b) * nb + ((xb & 0xFFFFFFFF) % nc_b) * CHUNK + (xb >> 32)  # global build row
            smp = torch.stack([rp[idx], _mulmod(xg, perm_multiplier(nb * world), nb * world) + 1], dim=1)
        allsmp = [torch.empty_like(smp) for _ in range(world)]
        dist.all_gather(allsmp, smp)
        q = torch.cat(allsmp)
        chunk = q[:, 0] & 0xFFFFFFFF
        mine = (q[:, 0] != -1) & (chunk >= rank * nc_p) & (chunk < (rank + 1) * nc_p)
        q = q[mine]
        bad += (_keys_at(data["probe"], lpos(q[:, 0], nc_p)) != q[:, 1]).sum()
        k = sample
    if world > 1:
        dist.all_reduce(got)
        dist.all_reduce(want)
        dist.all_reduce(bad)
    ok_rows, ok_b, ok_p = (bool(got[i] == want[i]) for i in range(3))
    return {"rows": int(got[0]), "rows_expected": int(want[0]), "rows_ok": ok_rows, "checksums": ok_b and ok_p,
            "build_checksum_ok": ok_b, "probe_checksum_ok": ok_p, "sampled_rows_per_rank": int(k),
            "sampled_key_mismatches": int(bad.item()), "ok": ok_rows and ok_b and ok_p and int(bad.item()) == 0,
            "how": "device-side check of the last timed step's PosLists (hj_result_device), reduced over ranks"}


# ------------------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------------------ CPU arm
def cpu_reference_sample(name, scale_of_full, n_threads, steps, warmup, seed=0xB200 + 2):
    """The reference's CPU JoinHash (oracle restatement, reference task granularity) on a bounded sample of the
    workload; returns (tuples/s best-of-steps, sample description, mean seconds per step)."""
    import oracle_binding as ob
    import workloads as wl
    n_build, n_probe, dtype, encoding, _ = WORKLOADS[name]
    nb, npr = max(1, int(n_build * scale_of_full)), max(1, int(n_probe * scale_of_full))
    rng = np.random.default_rng(seed)
    npdt = np.int32 if dtype == "int32" else np.int64
    if name == "c3":
        bk = wl.zipf_keys(rng, nb, nb, 1.0, npdt)
    else:
        bk = (rng.permutation(nb) + 1).astype(npdt)
    pk = rng.integers(1, nb + 1, npr, dtype=np.int64).astype(npdt)
    build = wl.make_relation(bk, CHUNK, encoding)
    probe = wl.make_relation(pk, CHUNK, encoding)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        _, _, info = ob.oracle_join(build, probe, 0, n_threads=n_threads)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    best = min(times)
    sample = f"{nb} x {npr} rows ({scale_of_full:g} of {name.upper()}), {encoding} chunks of {CHUNK}, Inner, {info['n_rows']} pairs"
    return (nb + npr) / best, sample, float(np.mean(times))


# ------------------------------------------------------------------------------------------------------------ main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the workload's row counts (1.0 = as published)")
    ap.add_argument("--cpu-scale", type=float, default=None,
                    help="fraction of the workload the CPU baseline runs (default: BASELINE.md 4.3 — 0.1 of C2/C3, 0.01 of C5, all of C1)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--slice-rows", type=int, default=0, help="hj_options.stream_slice_rows (0 = library default)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the device-side result check after the timed region")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n_build, n_probe, dtype, encoding, desc = WORKLOADS[args.workload]
    metric = "JoinHash input tuples/s"
    unit = "tuples/s"
    n_threads = os.cpu_count() or 1
    if args.cpu_scale is None:
        args.cpu_scale = {"c2": 0.1, "c3": 0.1, "c5": 0.01, "c1": 1.0}[args.workload]
    scaling = "strong" if args.workload in STRONG else "weak"

    def cpu_baseline(steps, warmup):
        """All host threads on the bounded sample + the single-threaded figure (what the reference's own
        join_benchmark.cpp:62-79 measures: no scheduler, jobs run inline) on a tenth of that sample."""
        val, sample, sec = cpu_reference_sample(args.workload, args.cpu_scale, n_threads, steps, warmup)
        val1, sample1, _ = cpu_reference_sample(args.workload, args.cpu_scale / 10, 1, 1, 1)
        return {"value": val, "unit": unit, "cores": n_threads, "kind": "port", "sample": sample,
                "value_1thread": val1, "sample_1thread": sample1}, sec

    if args.impl == "reference":
        if rank != 0:
            return 0
        cpu, sec = cpu_baseline(args.steps, args.warmup)
        line = {"impl": "reference", "metric": metric, "value": cpu["value"], "unit": unit, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
                "dtype": dtype, "data": "synthetic",
                "config": {"workload": desc, "sample": cpu["sample"]},
                "cpu_baseline": cpu,
                "e2e": {"value": cpu["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return 0

    import torch
    import torch.distributed as dist
    hj = importlib.import_module("hyrise-1_b200")
    abi = hj.abi
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- workload: weak scaling — every rank holds `scale` x C2 rows of ONE global join (global key domain and chunk ids)
    t0 = time.time()
    data = gen_device_workload(args.workload, args.scale, dev, seed=0xB200 + 2 + 1000 * rank, rank=rank, world=world)
    nb, npr = data["n_build"], data["n_probe"]
    build_dev = relation_from_chunks(hj, data["build"], abi.LOC_DEVICE)
    probe_dev = relation_from_chunks(hj, data["probe"], abi.LOC_DEVICE)
    torch.cuda.synchronize()
    log(f"[rank {rank}] generated {nb} x {npr} rows in {time.time() - t0:.1f}s; inputs "
        f"{(input_bytes(data['build']) + input_bytes(data['probe'])) / 1e9:.2f} GB")

    j = hj.HashJoin(device=local_rank, flags=abi.FLAG_TIMINGS, stream_slice_rows=args.slice_rows)
    if world > 1:
        uid = [hj.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        j.comm_init(uid[0], world, rank)
    # the library's kernels are launched on torch's current stream so that torch.cuda.Event brackets them
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        if world > 1:
            return j.join_sharded(build_dev, probe_dev, abi.MODE_INNER, stream=stream)
        return j.join(build_dev, probe_dev, abi.MODE_INNER, stream=stream)

    W = max(3, args.warmup)
    for _ in range(W):
        info = step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    phase_ms, launches = [], 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        info = step()
        tm = j.timings()
        phase_ms.append(tm)
        launches += tm["n_kernel_launches"]
    ev1.record()
    barrier()
    elapsed = ev0.elapsed_time(ev1) * 1e-3  # device time of exactly K steps, CUDA events on the launching stream
    clocks = sampler.stop()
    dev_ms = float(np.mean([t["total"] for t in phase_ms]))  # the library's own CUDA events
    if world > 1:
        tt = torch.tensor([elapsed], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        elapsed = float(tt.item())
        cnt = torch.tensor([nb + npr, info.n_rows], device=dev, dtype=torch.int64)
        dist.all_reduce(cnt)
        total_tuples, total_results = int(cnt[0]), int(cnt[1])
    else:
        total_tuples, total_results = nb + npr, info.n_rows
    ms_per_step = elapsed / args.steps * 1e3
    value = total_tuples / (elapsed / args.steps)

    # ---- roofline of the dominant kernel (the phase with the largest live CUDA-event time), DESIGN.md §5:
    # algorithmic bytes of a phase = the bytes that phase must move in the SURVEY §8d model, for this rank's rows
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        with open(peaks_path) as f:
            peak, peak_src = json.load(f)["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs, burst)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    k = 4 if dtype == "int32" else 8
    t_bytes = k + 8
    w_in = {"dict": 4 + 4, "value": k}[encoding]  # attr + amortised dictionary bytes per row for 1 Mi-row chunks
    mean_ms = {p_: float(np.mean([t[p_] for t in phase_ms])) for p_ in abi.PHASES}
    n_local = info.n_build_tuples + info.n_probe_tuples  # rows this rank partitions and joins (after the exchange)
    t_real = 8 if dtype == "int32" else 16  # this implementation's partitioned tuple: (key, 4-byte position)
    phase_bytes = {"materialize": n_local * w_in, "scatter1": n_local * (w_in + t_real), "scatter2": n_local * 2 * t_real,
                   "build_probe": n_local * t_real + info.n_rows * 16}
    phase_kernel = {"materialize": "k_materialize_hist", "scatter1": "k_scatter1", "scatter2": "k_scatter2",
                    "build_probe": "k_build_probe"}
    xpath = int(phase_ms[-1].get("xchg_path", 0))  # 0 single GPU, 1 general exchange, 2 push exchange
    if xpath == 2:
        w_in = t_real  # the local join's pass 1 reads the finished tuples the exchange delivered, not the columns
    if mean_ms["scatter1"] < 0.05 * mean_ms["scatter2"]:
        # default two-pass path (hj_fused.cuh): pass 1 runs inside the histogram kernel — the "materialize" phase reads the
        # input and writes the slab-local tuples, there is no separate scatter1 kernel, pass 2 gathers fragments
        phase_bytes["materialize"] = n_local * (w_in + t_real)
        del phase_bytes["scatter1"]
        phase_kernel.update({"materialize": "k_part1p<XCH>" if xpath == 2 else "k_part1", "scatter2": "k_scatter2t"})
    dom = max(phase_bytes, key=lambda p_: mean_ms[p_])
    traffic = None
    # dram bytes per launch from the committed `ncu --set full` capture — only while the kernels are the ones that were
    # profiled: the file records a hash of hyrise-1_b200/csrc/*, a stale one is refused (traffic stays null)
    tpath = os.path.join(ROOT, "profiles", "traffic_c2_full.json")
    if os.path.exists(tpath) and args.workload == "c2" and args.scale == 1.0 and world == 1:
        with open(tpath) as f:
            tj = json.load(f)
        if tj.get("csrc_sha256") == csrc_hash():
            traffic = tj.get(phase_kernel[dom])  # per launch set: the kernel's launches of one join (build + probe side)
        else:
            log("profiles/traffic_c2_full.json was captured from other kernel sources: roofline.traffic = null")
    whole_bytes = (nb + npr) * (2 * w_in + 2 * t_bytes) + info.n_rows * 16
    ach = phase_bytes[dom] / (mean_ms[dom] * 1e-3) / 1e9 if mean_ms[dom] else None
    roofline = {"bound": "hbm", "kernel": phase_kernel[dom], "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak if ach else None, "traffic": traffic, "peak_source": peak_src, "kernel_ms": mean_ms[dom],
                "algorithmic_bytes_per_launch_set": phase_bytes[dom],
                "per_phase": {p_: {"ms": mean_ms[p_], "algorithmic_bytes": phase_bytes[p_],
                                   "frac": phase_bytes[p_] / (mean_ms[p_] * 1e-3) / 1e9 / peak if mean_ms[p_] else None}
                              for p_ in phase_bytes},
                "whole_join": {"algorithmic_bytes": whole_bytes, "achieved": whole_bytes / (dev_ms * 1e-3) / 1e9,
                               "frac": whole_bytes / (dev_ms * 1e-3) / 1e9 / peak, "frac_of_nominal_8TBs":
                               whole_bytes / (dev_ms * 1e-3) / 1e9 / 8000.0, "device_ms": dev_ms},
                "phase_ms": mean_ms}
    if world > 1:
        sent = float(np.mean([t["bytes_nvlink_sent"] for t in phase_ms]))
        xms = mean_ms["xchg_transfer"] + (mean_ms["xchg_partition"] if xpath == 2 else 0.0)
        roofline["nvlink"] = {"bytes_sent_per_rank": sent, "transfer_ms": xms, "achieved": sent / (xms * 1e-3) / 1e9 if xms else None,
                              "peak": 900.0, "unit": "GB/s per direction", "frac": sent / (xms * 1e-3) / 1e9 / 900.0 if xms else None,
                              "exchange": {1: "general (count sweep + RowID-carrying tuples, hj_comm.cuh)",
                                           2: "push (one sweep, finished tuples stored into peer regions, hj_xpush.cuh)"}.get(xpath),
                              "note": "transfer_ms is the exchange kernels' time: decode + hash + staging + NVLink stores, "
                                      "both relations (push), or the probe side's scatter (general)"}

    # ---- the last timed step's PosLists are still in HBM: check them (outside the timed region, every N)
    verified = None
    if not args.no_verify:
        verified = verify_result(j, data, info, args.workload, rank, world, dev)
        torch.cuda.empty_cache()
        if rank == 0:
            log(f"verified: {verified}")

    # ---- e2e: same join through the C ABI from pinned host buffers, PosLists fetched to pinned host memory
    e2e = None
    if not args.no_e2e:
        ok = 1
        try:
            hb, hp = to_pinned_host(data["build"]), to_pinned_host(data["probe"])
            build_host = relation_from_chunks(hj, [(c, a.numpy(), None if b is None else b.numpy()) for c, a, b in hb], abi.LOC_HOST)
            probe_host = relation_from_chunks(hj, [(c, a.numpy(), None if b is None else b.numpy()) for c, a, b in hp], abi.LOC_HOST)
            out_rows = int(info.n_rows * 1.05) + 1024  # per-rank result sizes vary a little between exchanges of equal data
            out_b = torch.empty((out_rows, 2), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
            out_p = torch.empty((out_rows, 2), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
            h2d = input_bytes(hb) + input_bytes(hp)
        except (RuntimeError, MemoryError) as exc:  # not enough pinnable host memory for N ranks x 25 GB
            log(f"[rank {rank}] e2e leg: cannot pin the host buffers: {exc}")
            ok = 0
        if world > 1:  # every rank takes the same branch
            okt = torch.tensor([ok], device=dev, dtype=torch.int32)
            dist.all_reduce(okt, op=dist.ReduceOp.MIN)
            ok = int(okt.item())
        if ok:
            def e2e_step():
                # the operator's call sequence for host tables: bind the PosLists, join (streamed: copy-in, [exchange +] join
                # and copy-out of probe slices overlap), fetch (a no-op after a bound join)
                j.bind_output(out_b, out_p)
                if world > 1:
                    inf_ = j.join_sharded(build_host, probe_host, abi.MODE_INNER, stream=stream)
                else:
                    inf_ = j.join(build_host, probe_host, abi.MODE_INNER, stream=stream)
                j.fetch(out_b, out_p)
                return inf_
            e2e_step()
            barrier()
            n_e2e = max(1, min(args.steps, 3))
            t1 = time.perf_counter()
            for _ in range(n_e2e):
                inf = e2e_step()
            barrier()
            e2e_s = (time.perf_counter() - t1) / n_e2e
            if world > 1:
                tt = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                e2e_s = float(tt.item())
            e2e = {"value": total_tuples / e2e_s, "unit": unit, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": inf.n_rows * 16,
                   "ms_per_step": e2e_s * 1e3, "steps": n_e2e, "probe_slices": int(inf.n_probe_slices)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu, _ = cpu_baseline(2, 1)

    if rank == 0:
        line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps,
                "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling,
                "vs_baseline": None, "dtype": dtype, "data": data.get("source", "synthetic"),
                "config": {"workload": desc + (f"; {world} ranks each hold that many rows of one global join (hash exchange over NVLink)"
                                               if world > 1 and scaling == "weak" else ""),
                           "scale": args.scale, "n_build": nb * world, "n_probe": npr * world,
                           "chunk_rows": CHUNK, "mode": "Inner", "result_rows": total_results,
                           "radix_bits": info.radix_bits, "l2": "inputs larger than L2, no flush"},
                "clocks": clocks, "gpu_launches": launches, "verified": verified, "roofline": roofline, "cpu_baseline": cpu,
                "e2e": e2e}
        print(json.dumps(line), flush=True)
    j.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
