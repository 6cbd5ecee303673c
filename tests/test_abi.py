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


def test_null_handle_calls_are_rejected_not_fatal(nb):
    """Every entry point validates its handle before touching CUDA: callable without a GPU."""
    lib = nb.load_library()
    null = C.c_void_p(None)
    f, d, i = C.c_float(), C.c_double(), C.c_int()
    sz = C.c_size_t()
    assert lib.nbody_cuda_configure(null, 10, 0, 1) == -1
    assert lib.nbody_cuda_slice(null, C.byref(sz), None, None, None) == -1
    assert lib.nbody_cuda_upload_aos(null, None, 0) == -1
    assert lib.nbody_cuda_download_aos(null, None, 0) == -1
    assert lib.nbody_cuda_download_positions(null, None, 0) == -1
    assert lib.nbody_cuda_step(null, 1.0, 0.5, 1.0, 1) == -1
    assert lib.nbody_cuda_step_local(null, 1.0, 0.5, 1.0) == -1
    assert lib.nbody_cuda_compute_step_aos(null, None, 0, 1.0, 0.5, 1.0) == -1
    assert lib.nbody_cuda_synchronize(null) == -1
    assert lib.nbody_cuda_last_step_ms(null, C.byref(f)) == -1
    assert lib.nbody_cuda_profile_force(null, 1) == -1
    assert lib.nbody_cuda_force_time(null, C.byref(d), C.byref(i), 0) == -1
    assert lib.nbody_cuda_plan(null, None, None, None, None) == -1
    assert lib.nbody_cuda_energy(null, 1.0, 1.0, C.byref(d), C.byref(d)) == -1
    assert lib.nbody_cuda_generate(null, 0, 100.0, -1.0, 1.0, 1, C.byref(f)) == -1
    assert lib.nbody_cuda_render_frame(null, 8, 8, 800.0, 0.0, 0.0, None) == -1
    assert lib.nbody_cuda_peer_export(null, None) == -1
    assert lib.nbody_cuda_peer_attach(null, 0, None) == -1
    assert lib.nbody_cuda_peer_attach_local(null, 0, null) == -1
    assert lib.nbody_cuda_exchange(null) == -1
    assert lib.nbody_cuda_measure_fma_peak(null, 1, C.byref(d)) == -1
    assert lib.nbody_cuda_launch_count(null) == 0
    assert lib.nbody_cuda_stream(null) is None and lib.nbody_cuda_positions(null) is None
    lib.nbody_cuda_destroy(null)                                     # a no-op
    # the allocation-free entry point checks its size limit before anything else
    err = C.create_string_buffer(200)
    assert lib.nbody_cuda_accel_noalloc(None, None, 65537, 1.0, 1.0, None, err, len(err)) == -1
    assert b"NBODY_NOALLOC_MAX_N" in err.value
    assert lib.nbody_cuda_accel_noalloc(None, None, 0, 1.0, 1.0, None, err, len(err)) == 0
