"""GPU: the C++ drop-in class (n-body-simulation_b200/host/CudaNaiveSimulation.*) exercised through
the reference's abstract interface by a small C++ program (tests/cpp/test_cuda_naive.cpp)."""
import os
import subprocess

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_cpp_host_class():
    exe = os.path.join(ROOT, "tests", "cpp", "test_cuda_naive")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", os.path.join(ROOT, "n-body-simulation_b200", "host")], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and "PASS" in r.stdout
