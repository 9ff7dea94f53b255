"""Multi-GPU path: the partition function on the host, a world-size-2 gloo model of the exchange on CPU,
and — on a box with >= 2 GPUs — the real hj_join_sharded over NCCL. All compare against the oracle's join of the GLOBAL
relations (tests/sharded_check.py)."""
import importlib
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

hj = importlib.import_module("hyrise-1_b200")
abi = hj.abi
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _torchrun(n, *args, timeout=900):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "sharded_check.py"), *args]
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


def test_exchange_rank_of_is_a_balanced_partition(built):
    for key_type, keys in ((abi.KEY_INT32, np.arange(-5000, 5000)), (abi.KEY_INT64, np.arange(2**40, 2**40 + 10000))):
        for n_ranks in (1, 2, 3, 8, 16):
            dest = np.array([hj.exchange_rank_of(int(k), key_type, n_ranks) for k in keys])
            assert dest.min() >= 0 and dest.max() < n_ranks
            counts = np.bincount(dest, minlength=n_ranks)
            assert counts.min() > 0.8 * len(keys) / n_ranks, (key_type, n_ranks, counts)
            again = np.array([hj.exchange_rank_of(int(k), key_type, n_ranks) for k in keys[:100]])
            assert np.array_equal(again, dest[:100])
    # a key that fits int32 is routed by the HASHED type of the join: int32 and int64 joins may disagree, which is why all
    # ranks agree on the type first (hj_join_sharded all-gathers "I hold an int64 column")
    assert any(hj.exchange_rank_of(k, abi.KEY_INT32, 8) != hj.exchange_rank_of(k, abi.KEY_INT64, 8) for k in range(100))


def test_sharded_model_world2_gloo(built):
    r = _torchrun(2, "--backend", "gloo")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count(" OK") == 20 and "MISMATCH" not in r.stdout


def test_sharded_entry_points_fail_loudly_without_comm(built):
    # no GPU here: hj_create itself refuses (no CPU fallback); on a GPU box hj_join_sharded without hj_comm_init is HJ_ERR_COMM
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(hj.HjError):
            hj.HashJoin(device=0)
        return
    j = hj.HashJoin(device=0)
    import workloads as wl
    rel = wl.make_relation(np.arange(10, dtype=np.int32), 4)
    with pytest.raises(hj.HjError) as e:
        j.join_sharded(rel, rel, "Inner")
    assert e.value.status == abi.ERR_COMM
    j.close()


@pytest.mark.gpu
def test_sharded_join_nccl_2gpus():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    # general exchange peer-mapped (default for these shapes), NCCL-staged, and sliced (slice s+1 travels while slice s is
    # joined) forced onto these small cases; then the PUSH exchange: --radix-bits 10 makes the local join two-pass, so the
    # shards with regular chunks take it (the irregular / gathered ones stay on the general path), and the same with
    # HJ_FLAG_XCHG_GENERAL as the control
    # ... and the STREAMED push exchange: bound host outputs, probe slices of >= 3000 rows (copy-in, exchange + join, copy-out overlap)
    for extra in ([], ["--staged"], ["--slice-rows", "700"], ["--radix-bits", "10"], ["--radix-bits", "10", "--general"],
                  ["--radix-bits", "10", "--bound", "--slice-rows", "3000"]):
        r = _torchrun(2, "--backend", "nccl", "--big", *extra)
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
        assert "MISMATCH" not in r.stdout and r.stdout.count(" OK") == 55
        if extra and extra[0] == "--slice-rows":
            assert "slices=8" in r.stdout and "slices=0" in r.stdout  # big cases sliced, the empty-build case not
        if "--bound" in extra:
            assert r.stdout.count("exchange=push") == 40 and "slices=21" in r.stdout and "slices=24" in r.stdout, r.stdout[-3000:]
        elif extra == ["--radix-bits", "10"]:
            # regular shards: the six 2^k cases, "empty build" (128-row chunks) and the 2M x 8M case
            assert r.stdout.count("exchange=push") == 40 and r.stdout.count("exchange=general") == 15, r.stdout[-3000:]
        else:
            assert "exchange=push" not in r.stdout  # single-pass local joins / HJ_FLAG_XCHG_GENERAL
