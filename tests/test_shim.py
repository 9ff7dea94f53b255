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


@pytest.mark.gpu
def test_shim_replays_reference_cases_with_device_scans(built):
    """Same 53 cases, but every `scan` the device operator covers (integer column vs integer literal) runs through the C++
    opossum::TableScan -> hj_scan, so the filtered join inputs are produced on the GPU."""
    env = dict(os.environ, HJ_SHIM_DEVICE_SCAN="1")
    r = subprocess.run([EXE, FIXTURES], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert "53 cases run, 0 failed, 0 out of scope" in r.stdout, r.stdout[-1000:]
    n = int(r.stdout.rsplit("shim_test: ", 1)[1].split(" scans ran on the device")[0])
    assert n >= 1, r.stdout[-500:]


@pytest.mark.gpu
def test_shim_replays_reference_float_and_string_scan_cases(built):
    """table_scan_test.cpp DoubleScan (float column behind a reference table) and table_scan_string_test.cpp ScanEquals /
    ScanNotEquals / ScanLessThan (dictionary-encoded string column: dictionary searched on the host, attribute vectors
    scanned on the device by hj_scan_value_ids) through the C++ operator."""
    env = dict(os.environ, HJ_SHIM_DEVICE_SCAN="1")
    fixtures = os.path.join(ROOT, "tests", "golden", "scan_shim_fixtures.txt")
    r = subprocess.run([EXE, fixtures], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert "5 cases run, 0 failed" in r.stdout, r.stdout[-1000:]
    n = int(r.stdout.rsplit("shim_test: ", 1)[1].split(" scans ran on the device")[0])
    assert n == 7, r.stdout[-500:]


@pytest.mark.gpu
def test_shim_replays_reference_ungrouped_aggregate_cases(built):
    """aggregate_test.cpp NoGroupbySingleAggregate{Max,Min,Sum,Avg,Count} — on the data table (value and dictionary chunks)
    and, like the reference's test_output, behind a TableScan — TwoAggregateEmptyTable, and all 39 numeric group-by cases (1, 2
    and 3 group-by columns, value and dictionary chunks, with and without NULLs, COUNT(*)), through the C++ operators
    opossum::TableScan -> opossum::Aggregate (hj_scan, hj_aggregate_relation, hj_aggregate_grouped_rows_relation)."""
    env = dict(os.environ, HJ_SHIM_DEVICE_SCAN="1")
    fixtures = os.path.join(ROOT, "tests", "golden", "aggregate_shim_fixtures.txt")
    r = subprocess.run([EXE, fixtures], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert "94 cases run, 0 failed" in r.stdout, r.stdout[-1000:]
    n = int(r.stdout.rsplit("shim_test: ", 1)[1].split(" scans ran on the device")[0])
    assert n == 6, r.stdout[-500:]
