"""The C++ host mirror of the reference's Go API (include/b200hnsw_host.hpp) through its own test program
(tests/cpp/host_test.cpp): the reference's validation / maxLevel tests without a device, and on the GPU the distance
KATs, TestGraph_DefaultCosine and Search / BatchSearch / Delete parity against the oracle."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
CPP = ROOT / "tests" / "cpp"


def _build():
    from oracle import binding as oracle
    oracle.build()  # liboracle.so: the checker the test program links
    subprocess.run(["make", "-C", str(CPP)], check=True, capture_output=True)
    return CPP / "host_test"


def test_cpp_host_validation_and_loud_failure_without_device():
    exe = _build()
    r = subprocess.run([str(exe), "--cpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([str(exe), "--cpu-nodevice"], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr


@pytest.mark.gpu
def test_cpp_host_matches_oracle_on_gpu():
    exe = _build()
    r = subprocess.run([str(exe), "--gpu"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "--gpu: 0 failure(s)" in r.stdout
