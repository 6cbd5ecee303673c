"""CPU: the C-ABI library loads and exports exactly what include/nbody_cuda.h declares; without a
GPU it refuses to work (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, have_gpu


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "nbody_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nbody_cuda_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree(nb):
    assert declared_symbols() == sorted(nb.SIGNATURES)


def test_library_exports_every_declared_symbol(nb):
    lib = nb.load_library()
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_record_size_and_version(nb):
    lib = nb.load_library()
    assert lib.nbody_cuda_particle_bytes() == 40 == nb.PARTICLE_DTYPE.itemsize
    assert b"sm_100a" in lib.nbody_cuda_version()
    assert [nb.PARTICLE_DTYPE.fields[k][1] for k in ("position", "velocity", "acceleration", "mass")] == [0, 12, 24, 36]


def test_library_is_built_for_sm_100a_only(nb):
    import subprocess

    out = subprocess.run(["cuobjdump", "-lelf", nb.LIB_PATH], capture_output=True, text=True).stdout
    if not out:
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in out and not re.search(r"sm_(?!100a)\d+", out)


def test_no_cpu_fallback(nb):
    """Without a device, creating the simulation fails loudly with the library's message."""
    if have_gpu():
        pytest.skip("a GPU is present")
    with pytest.raises(nb.NBodyCudaError) as e:
        nb.CudaNaiveSimulation()
    assert "no CPU fallback" in str(e.value)
    lib = nb.load_library()
    h = C.c_void_p()
    assert lib.nbody_cuda_create(C.byref(h), 0, 0) < 0 and not h.value


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may reference it."""
    pkg = os.path.join(ROOT, "n-body-simulation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in src and "libnbody_ref" not in src, f
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
