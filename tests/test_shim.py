"""The C++ drop-in operator (hyrise-1_b200/host: opossum::JoinHash with the reference's class surface, calling the C ABI)
replays the reference's own JoinHash test cases; see host/shim_test_main.cpp."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "hyrise-1_b200", "host", "shim_test")
FIXTURES = os.path.join(ROOT, "tests", "golden", "join_fixtures.txt")


def test_shim_fails_loudly_without_a_gpu(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by test_shim_replays_reference_cases")
    r = subprocess.run([EXE, FIXTURES, "JoinEquiTest.LeftJoin"], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0
    assert "no CPU fallback" in r.stdout or "no CUDA device" in r.stdout, r.stdout[-2000:]


@pytest.mark.gpu
def test_shim_replays_reference_cases(built):
    r = subprocess.run([EXE, FIXTURES], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert "53 cases run, 0 failed, 0 out of scope" in r.stdout, r.stdout[-1000:]
