import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """This repo's C restatement (oracle/nbody_oracle.c), built on demand."""
    from oracle import Oracle

    return Oracle()


@pytest.fixture(scope="session")
def reference():
    """The reference's own classes (oracle/_ref/libnbody_ref.so); skipped where it was never built."""
    from oracle import Reference

    if not Reference.available():
        pytest.skip("oracle/_ref/libnbody_ref.so not built (needs /root/reference at build time)")
    return Reference()


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


GOLDEN_CASES = ["disk_n256", "plummer_n512", "clusters_4x64", "solar_9", "disk_n300"]


@pytest.fixture(scope="session")
def nb():
    import nbody_b200

    return nbody_b200


def have_gpu():
    import nbody_b200

    try:
        return nbody_b200.load_library().nbody_cuda_device_count() > 0
    except Exception:
        return False


def rel_err(a, b):
    """Per-particle ||a-b||_2 / ||b||_2 for (n,3) arrays, in float64."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    num = np.linalg.norm(a - b, axis=1)
    den = np.linalg.norm(b, axis=1)
    den = np.where(den == 0, 1.0, den)
    return num / den
