"""GPU: the allocation-free entry point `nbody_cuda_accel_noalloc`, called through the C ABI.

It exists for hosts that cannot let this library allocate device memory (a fixed arena owns the
GPU, or cudaMalloc's device-wide synchronisation is unaffordable in their loop), so its scratch is
a fixed array inside the library.  That array is shared by every call and never cleared -- hence
this test makes TWO calls in a row with different systems and checks each against the ordinary,
allocating path (upload -> one step -> download, `acceleration` field), at the tolerance the fast
mode is held to elsewhere in this suite (2e-5 relative per particle).  The device buffers handed to
the entry point are allocated here, by the caller, with torch.

Collected last (file name) so that a failure here cannot hide the other tests under `-x`.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu

TOL = 2e-5


def reference_path_acc(nb, p, G, eps2):
    """Accelerations from the existing entry points (allocating handle, fast mode)."""
    sim = nb.CudaNaiveSimulation(mode="fast")
    q = p.copy()
    sim.computeStep(q, nb.SimParams(G=G, dt=0.0, min_r2=eps2))     # dt = 0: positions untouched, acceleration written
    sim.close()
    return q["acceleration"]


def call_noalloc(nb, torch, p, G, eps2, stream=None, in_place=False):
    """One call through the C ABI.  `stream`: a torch.cuda.Stream to enqueue on (None: the default
    stream).  `in_place`: hand the position buffer in as the output buffer too."""
    lib = nb.load_library()
    n = p.shape[0]
    xyzm = np.concatenate([p["position"], p["mass"][:, None]], axis=1).astype(np.float32)
    d_pos = torch.from_numpy(xyzm).cuda().contiguous()
    d_acc = d_pos if in_place else torch.full((max(n, 1), 4), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    err = C.create_string_buffer(256)
    rc = lib.nbody_cuda_accel_noalloc(C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_acc.data_ptr()), n, G, eps2,
                                      C.c_void_p(stream.cuda_stream) if stream is not None else None, err, len(err))
    assert rc == 0, err.value
    torch.cuda.synchronize()
    return d_acc[:n, :3].cpu().numpy(), d_acc[:n, 3].cpu().numpy()


def truth_acc(oracle, p, G, eps2):
    """FP64 accelerations from the CPU oracle (oracle/nbody_oracle.c), the checker of this suite."""
    return oracle.accel_f64(p, G, eps2)


def test_noalloc_matches_existing_path_on_two_calls_in_a_row(nb):
    import torch

    torch.cuda.set_device(0)
    G = 1.25
    first, eps1 = nb.datasets.plummer_sphere(5000, seed=31)
    first["mass"] = np.random.default_rng(7).uniform(0.5, 2.0, 5000).astype(np.float32)
    second, eps2 = nb.datasets.galaxy_disk(3001, seed=32)          # different size, different system
    for p, eps in ((first, eps1), (second, eps2)):
        got, w = call_noalloc(nb, torch, p, G, eps)
        want = reference_path_acc(nb, p, G, eps)
        assert np.isfinite(got).all() and np.all(w == 0)
        assert rel_err(got, want).max() < TOL


def test_noalloc_against_the_oracle_on_a_side_stream_and_in_place(nb, oracle):
    """The same two systems, checked against the FP64 CPU oracle rather than against this library's own
    fast path; the first call is enqueued on a non-default stream, the second writes the accelerations
    over the positions it was given (the sources are copied into the built-in array first)."""
    import torch

    torch.cuda.set_device(0)
    G = 0.75
    a, eps_a = nb.datasets.plummer_sphere(4099, seed=41)
    a["mass"] = np.random.default_rng(9).uniform(0.5, 2.0, 4099).astype(np.float32)
    b, eps_b = nb.datasets.galaxy_disk(2500, seed=42)
    side = torch.cuda.Stream()
    got_a, _ = call_noalloc(nb, torch, a, G, eps_a, stream=side)
    got_b, w_b = call_noalloc(nb, torch, b, G, eps_b, in_place=True)
    assert np.all(w_b == 0)
    assert rel_err(got_a, truth_acc(oracle, a, G, eps_a)).max() < TOL
    assert rel_err(got_b, truth_acc(oracle, b, G, eps_b)).max() < TOL


def test_noalloc_really_allocates_nothing(nb):
    """The property the entry point exists for: device memory in use is the same before and after a
    call (cudaMemGetInfo through torch), once the first call on the device has done its look-up."""
    import torch

    torch.cuda.set_device(0)
    p, eps = nb.datasets.plummer_sphere(3000, seed=43)
    call_noalloc(nb, torch, p, 1.0, eps)                                # first call on the device: set-up
    lib = nb.load_library()
    xyzm = np.concatenate([p["position"], p["mass"][:, None]], axis=1).astype(np.float32)
    d_pos = torch.from_numpy(xyzm).cuda().contiguous()
    d_acc = torch.empty_like(d_pos)
    err = C.create_string_buffer(256)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        rc = lib.nbody_cuda_accel_noalloc(C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_acc.data_ptr()), 3000, 1.0, eps,
                                          None, err, len(err))
        assert rc == 0, err.value
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before


def test_noalloc_limits_and_errors(nb):
    import torch

    torch.cuda.set_device(0)
    lib = nb.load_library()
    err = C.create_string_buffer(256)
    dummy = torch.zeros(4, device="cuda")
    too_many = 65536 + 1
    rc = lib.nbody_cuda_accel_noalloc(C.c_void_p(dummy.data_ptr()), C.c_void_p(dummy.data_ptr()), too_many, 1.0, 1.0,
                                      None, err, len(err))
    assert rc == -1 and b"NBODY_NOALLOC_MAX_N" in err.value          # refused before anything is launched
    assert lib.nbody_cuda_accel_noalloc(None, None, 0, 1.0, 1.0, None, err, len(err)) == 0      # empty system
    assert lib.nbody_cuda_accel_noalloc(None, None, 10, 1.0, 1.0, None, err, len(err)) == -1
    # the largest size it accepts, against the allocating path
    p, eps = nb.datasets.plummer_sphere(65536, seed=33)
    got, _ = call_noalloc(nb, torch, p, 1.0, eps)
    assert rel_err(got, reference_path_acc(nb, p, 1.0, eps)).max() < TOL
