"""GPU parity tests: the CUDA path, called through the C ABI (ctypes -> libnbody_cuda.so), against
the oracle and the committed golden vectors of the reference.

Bars:
  strict mode : BIT-EXACT with the reference's NaiveSimulation -- positions, velocities,
                accelerations -- on every input and step count tested.
  fast mode   : FP32 with a different summation order and MUFU.RSQ; single-step acceleration error
                against the FP64 truth no larger than the reference's own FP32 error against that
                truth (max within 1.5x to allow for a different worst particle, median no larger),
                and within 2e-5 relative of the reference's FP32 result in the median.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import GOLDEN_CASES, load_golden, rel_err

pytestmark = pytest.mark.gpu


def same_bits(a, b):
    return a.tobytes() == b.tobytes()


def fields_equal(a, b):
    return {k: same_bits(np.ascontiguousarray(a[k]), np.ascontiguousarray(b[k]))
            for k in ("position", "velocity", "acceleration", "mass")}


@pytest.fixture(scope="module")
def strict(nb):
    s = nb.CudaNaiveSimulation(mode="strict")
    yield s
    s.close()


@pytest.fixture(scope="module")
def fast(nb):
    s = nb.CudaNaiveSimulation(mode="fast")
    yield s
    s.close()


def run_compute_steps(sim, nb, p, G, dt, eps2, steps):
    q = p.copy()
    params = nb.SimParams(G=G, dt=dt, min_r2=eps2)
    for _ in range(steps):
        sim.computeStep(q, params)
    return q


# ------------------------------------------------------------------------------------------------
# strict mode: bit parity
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_strict_matches_reference_golden_bitwise(strict, nb, case):
    g = load_golden(case)
    for k in [k for k in g if k.startswith("naive_step")]:
        steps = int(k[len("naive_step"):])
        got = run_compute_steps(strict, nb, g["initial"], float(g["G"]), float(g["dt"]), float(g["min_r2"]), steps)
        assert all(fields_equal(got, g[k]).values()), (case, k, fields_equal(got, g[k]))


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_strict_device_resident_equals_per_step_round_trips(strict, nb, case):
    g = load_golden(case)
    k = sorted([k for k in g if k.startswith("naive_step")], key=lambda s: int(s[10:]))[-1]
    steps = int(k[len("naive_step"):])
    q = g["initial"].copy()
    strict.upload(q)
    strict.step(nb.SimParams(float(g["G"]), float(g["dt"]), float(g["min_r2"])), steps)
    strict.download(q)
    assert same_bits(q, g[k])


def test_strict_config1_disk_4096_100_steps(strict, nb, oracle):
    """BASELINE configs[0] trajectory: N=4096 galaxy disk (reference generator, seed 12345 when the
    reference build is present, else this repo's generator), 100 steps, bit for bit."""
    from oracle import Reference

    if Reference.available():
        p, eps = Reference().disk(4096, 12345)
    else:
        p, eps = nb.datasets.galaxy_disk(4096, seed=12345)
    want = oracle.step_rows(p, 1.0, 0.5, eps, 100)
    q = p.copy()
    strict.upload(q)
    strict.step(nb.SimParams(1.0, 0.5, eps), 100)
    strict.download(q)
    assert all(fields_equal(q, want).values()), fields_equal(q, want)


@pytest.mark.parametrize("kind", ["plummer", "disk"])
def test_strict_config2_n65536_three_steps(strict, nb, oracle, kind):
    """BASELINE configs[1] size: bit parity at N=65536 (3 steps; the oracle needs ~1.3e10 pair terms)."""
    if kind == "plummer":
        p, eps = nb.datasets.plummer_sphere(65536, seed=12345)
    else:
        p, eps = nb.datasets.galaxy_disk(65536, seed=12345)
    want = oracle.step_rows(p, 1.0, 0.5, eps, 3)
    q = p.copy()
    strict.upload(q)
    strict.step(nb.SimParams(1.0, 0.5, eps), 3)
    strict.download(q)
    assert all(fields_equal(q, want).values()), fields_equal(q, want)


# ------------------------------------------------------------------------------------------------
# edge cases the reference's interface allows (any size, size 0, size changing per call)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", ["strict", "fast"])
@pytest.mark.parametrize("n", [0, 1, 2, 3, 31, 255, 256, 257, 511, 513, 1000, 1025, 2049])
def test_sizes_around_tile_edges(nb, oracle, strict, fast, mode, n):
    sim = strict if mode == "strict" else fast
    p, eps = nb.datasets.plummer_sphere(max(n, 1), seed=n + 1)
    p = p[:n].copy()
    p["mass"] = np.linspace(0.5, 2.0, n, dtype=np.float32) if n else p["mass"]
    want = oracle.step_rows(p, 1.3, 0.25, eps, 1)
    got = run_compute_steps(sim, nb, p, 1.3, 0.25, eps, 1)
    assert got.shape == p.shape and same_bits(np.ascontiguousarray(got["mass"]), np.ascontiguousarray(p["mass"]))
    if mode == "strict":
        assert same_bits(got, want)
    elif n > 1:
        assert rel_err(got["acceleration"], want["acceleration"]).max() < 2e-5
        assert rel_err(got["position"], want["position"]).max() < 1e-6


def test_size_changes_between_calls(nb, oracle, strict):
    """computeStep must tolerate a different size() on every call (headless_main.cpp:262)."""
    for n in (700, 64, 3000, 0, 700):
        p, eps = nb.datasets.galaxy_disk(max(n, 2), seed=n)
        p = p[:n].copy()
        want = oracle.step_rows(p, 1.0, 0.5, eps, 1)
        assert same_bits(run_compute_steps(strict, nb, p, 1.0, 0.5, eps, 1), want)


@pytest.mark.parametrize("mode", ["strict", "fast"])
def test_zero_mass_particles_and_zero_softening(nb, oracle, strict, fast, mode):
    sim = strict if mode == "strict" else fast
    p, _ = nb.datasets.plummer_sphere(777, seed=5)
    p["mass"][::3] = 0.0                     # massless tracers feel forces but exert none
    want = oracle.step_rows(p, 1.0, 0.5, 0.0, 1)      # min_r2 = 0: self term must be skipped by index
    got = run_compute_steps(sim, nb, p, 1.0, 0.5, 0.0, 1)
    assert np.isfinite(got["acceleration"]).all()
    if mode == "strict":
        assert same_bits(got, want)
    else:
        assert rel_err(got["acceleration"], want["acceleration"]).max() < 2e-5


def test_tiny_default_softening(nb, oracle, fast):
    """SimParams default min_r2 = 1e-8 (SimulationAlgorithm.h:8): the self term must not poison the sum."""
    p, _ = nb.datasets.plummer_sphere(1024, seed=8)
    want = oracle.step_rows(p, 1.0, 0.05, 1e-8, 1)
    got = run_compute_steps(fast, nb, p, 1.0, 0.05, 1e-8, 1)
    assert np.isfinite(got["acceleration"]).all()
    assert rel_err(got["acceleration"], want["acceleration"]).max() < 2e-5


def test_upload_download_round_trip(nb, fast):
    p, _ = nb.datasets.galaxy_disk(3001, seed=3)
    p["acceleration"] = 123.0                 # must be ignored on upload, zero before the first step
    q = p.copy()
    fast.upload(q)
    out = np.zeros_like(q)
    fast.download(out)
    for k in ("position", "velocity", "mass"):
        assert same_bits(np.ascontiguousarray(out[k]), np.ascontiguousarray(p[k])), k
    assert np.all(out["acceleration"] == 0)


def test_bad_arguments_are_reported_not_fatal(nb, fast):
    p, _ = nb.datasets.plummer_sphere(100, seed=1)
    fast.upload(p)
    with pytest.raises(nb.NBodyCudaError):
        fast.download(np.zeros(99, nb.PARTICLE_DTYPE))
    lib = nb.load_library()
    h = C.c_void_p()
    assert lib.nbody_cuda_create(C.byref(h), 0, 7) == -1          # unknown mode
    assert lib.nbody_cuda_create(C.byref(h), 4096, 0) == -1       # device out of range
    assert b"device" in lib.nbody_cuda_last_error(None)


# ------------------------------------------------------------------------------------------------
# fast mode: accuracy against FP64 truth and against the reference's FP32 result
# ------------------------------------------------------------------------------------------------
def truth_acc(oracle, p, eps):
    return oracle.accel_f64(p, 1.0, eps)


def accuracy_report(nb, oracle, fast, p, eps):
    truth = oracle.accel_f64(p, 1.0, eps)
    ref32 = oracle.accel_rows(p, 1.0, eps)
    got = run_compute_steps(fast, nb, p, 1.0, 0.5, eps, 1)
    e_ref = rel_err(ref32, truth)
    e_got = rel_err(got["acceleration"], truth)
    e_vs = rel_err(got["acceleration"], ref32)
    return dict(ref_max=e_ref.max(), ref_med=np.median(e_ref), got_max=e_got.max(), got_med=np.median(e_got),
                vs_max=e_vs.max(), vs_med=np.median(e_vs)), got


@pytest.mark.parametrize("kind,n", [("plummer", 4096), ("disk", 4096), ("plummer", 16384), ("disk", 16384)])
def test_fast_accuracy_against_fp64_truth(nb, oracle, fast, kind, n):
    p, eps = (nb.datasets.plummer_sphere if kind == "plummer" else nb.datasets.galaxy_disk)(n, seed=12345)
    r, got = accuracy_report(nb, oracle, fast, p, eps)
    print(kind, n, {k: float(f"{v:.3g}") for k, v in r.items()})
    assert r["got_med"] <= r["ref_med"] * 1.05 + 1e-7      # no worse than the reference's own FP32 sum
    assert r["got_max"] <= r["ref_max"] * 1.5 + 1e-6
    assert r["vs_med"] <= 2e-5                             # and close to the reference's FP32 result
    assert r["vs_max"] <= 2.0 * r["ref_max"] + 2e-6
    # integrator: position after the step against the FP64 truth x + (v + a*dt)*dt
    dt = 0.5
    x64 = p["position"].astype(np.float64) + (p["velocity"].astype(np.float64) + truth_acc(oracle, p, eps) * dt) * dt
    want = oracle.step_rows(p, 1.0, dt, eps, 1)
    e_got, e_ref = rel_err(got["position"], x64), rel_err(want["position"], x64)
    assert e_got.max() <= e_ref.max() * 1.5 + 2e-7 and np.median(e_got) <= np.median(e_ref) * 1.05 + 1e-8


def test_fast_is_deterministic(nb, fast):
    p, eps = nb.datasets.plummer_sphere(20000, seed=4)
    a = run_compute_steps(fast, nb, p, 1.0, 0.5, eps, 2)
    b = run_compute_steps(fast, nb, p, 1.0, 0.5, eps, 2)
    assert same_bits(a, b)


def test_fast_plummer_100_steps_tracks_reference(nb, oracle, fast):
    """Multi-step parity on the non-chaotic input: N=4096 Plummer, 100 steps,
    positions within 1e-4 of the reference trajectory (the reference's own FP32-vs-FP64 drift over
    the same run is ~4e-6)."""
    p, eps = nb.datasets.plummer_sphere(4096, seed=12345)
    want = oracle.step_rows(p, 1.0, 0.5, eps, 100)
    q = p.copy()
    fast.upload(q)
    fast.step(nb.SimParams(1.0, 0.5, eps), 100)
    fast.download(q)
    e = rel_err(q["position"], want["position"])
    print("100-step position error max/median", e.max(), np.median(e))
    assert e.max() < 1e-4 and np.median(e) < 1e-5


def test_energy_matches_oracle_and_drift_parity(nb, oracle, fast):
    p, eps = nb.datasets.plummer_sphere(4096, seed=21)
    params = nb.SimParams(1.0, 0.5, eps)
    fast.upload(p.copy())
    k0, u0 = fast.energy(params)
    ko, uo = oracle.energy(p, 1.0, eps)
    assert k0 == pytest.approx(ko, rel=1e-10) and u0 == pytest.approx(uo, rel=1e-9)
    fast.step(params, 20)
    k1, u1 = fast.energy(params)
    ref = oracle.step_rows(p, 1.0, 0.5, eps, 20)
    kr, ur = oracle.energy(ref, 1.0, eps)
    drift_ours = (k1 + u1 - k0 - u0) / abs(k0 + u0)
    drift_ref = (kr + ur - ko - uo) / abs(ko + uo)
    print("energy drift ours/ref", drift_ours, drift_ref)
    assert drift_ours == pytest.approx(drift_ref, rel=0.01, abs=1e-7)


# ------------------------------------------------------------------------------------------------
# full-size, size-independent properties (BASELINE configs 2-3 sizes)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [65536, 262144])
def test_full_size_properties(nb, oracle, fast, strict, n):
    p, eps = nb.datasets.plummer_sphere(n, seed=12345)
    p["mass"] = np.random.default_rng(1).uniform(0.5, 2.0, n).astype(np.float32)
    params = nb.SimParams(1.0, 0.5, eps)
    q = p.copy()
    fast.upload(q)
    fast.step(params, 1)
    fast.download(q)
    acc = q["acceleration"].astype(np.float64)
    m = p["mass"].astype(np.float64)[:, None]
    # Newton's third law as a checksum: sum m_i a_i = 0 up to rounding
    net = np.abs((m * acc).sum(axis=0)).max()
    scale = (m * np.abs(acc)).sum()
    assert net / scale < 1e-6
    # a sample of rows against the FP64 truth and the reference-order FP32 rows
    rows = np.random.default_rng(2).choice(n, 64, replace=False)
    for i in rows[:16]:
        truth = oracle.accel_f64(p, 1.0, eps, int(i), int(i) + 1)
        assert rel_err(acc[i:i + 1], truth).max() < 2e-5
    # strict and fast agree at full size (strict is bit-equal to the reference by the tests above)
    s = p.copy()
    strict.upload(s)
    strict.step(params, 1)
    strict.download(s)
    e = rel_err(q["acceleration"], s["acceleration"])
    assert np.median(e) < 2e-5 and e.max() < 1e-3
    for i in rows[:8]:
        want = oracle.accel_rows(p, 1.0, eps, int(i), int(i) + 1)
        assert same_bits(np.ascontiguousarray(s["acceleration"][i:i + 1]), want)


def test_uniform_mass_path_agrees_with_general_path(nb, oracle, fast):
    """Equal masses (every Plummer input) take a shorter inner loop (m_j applied once to the sum);
    one differing mass switches back to the general loop.  Both must satisfy the fast-mode bar."""
    p, eps = nb.datasets.plummer_sphere(6000, seed=44)
    p["mass"] = np.float32(0.75)
    want = oracle.step_rows(p, 1.5, 0.5, eps, 1)
    got_u = run_compute_steps(fast, nb, p, 1.5, 0.5, eps, 1)
    assert rel_err(got_u["acceleration"], want["acceleration"]).max() < 2e-5
    q = p.copy()
    q["mass"][4321] = np.float32(0.7500001)
    want_q = oracle.step_rows(q, 1.5, 0.5, eps, 1)
    got_g = run_compute_steps(fast, nb, q, 1.5, 0.5, eps, 1)
    assert rel_err(got_g["acceleration"], want_q["acceleration"]).max() < 2e-5
    z = p.copy()
    z["mass"] = 0.0                                   # all-zero masses: uniform, and every force vanishes
    got_z = run_compute_steps(fast, nb, z, 1.5, 0.5, eps, 1)
    assert np.all(got_z["acceleration"] == 0)


@pytest.mark.parametrize("kind,n", [("plummer", 1000), ("plummer", 4096), ("disk", 4099), ("disk", 10007), ("plummer", 20000)])
def test_resident_kernel_equals_two_kernel_path_bitwise(nb, kind, n):
    """Small systems run all n_steps in one cooperative launch (grid barriers instead of kernel
    boundaries, csrc/nbody_kernels.cuh: resident_steps_kernel).  Same plan, same partial sums, same
    order: every field must come out bit-identical to the force + finish kernel pair, for odd and
    even step counts (the double-buffered position arrays end on either side)."""
    p, eps = (nb.datasets.plummer_sphere if kind == "plummer" else nb.datasets.galaxy_disk)(n, seed=77)
    params = nb.SimParams(1.0, 0.5, eps)
    out = {}
    for resident in (0, 2):
        sim = nb.CudaNaiveSimulation(mode="fast")
        sim.set_option("resident_kernel", resident)
        q = p.copy()
        sim.upload(q)
        launches0 = sim.launch_count()
        sim.step(params, 1); sim.step(params, 5); sim.step(params, 2)
        launches = sim.launch_count() - launches0
        sim.download(q)
        k, u = sim.energy(params)
        sim.close()
        out[resident] = (q, launches, k + u)
    assert same_bits(out[0][0], out[2][0])
    assert out[0][2] == out[2][2]
    assert out[2][1] < out[0][1]           # 3 calls -> at most 2 launches each, against 2 kernels per step


def test_launch_counter_and_timers(nb, fast):
    p, eps = nb.datasets.plummer_sphere(8192, seed=1)
    fast.upload(p)
    before = fast.launch_count()
    fast.profile_force(True)
    fast.step(nb.SimParams(1.0, 0.5, eps), 5)
    ms, n = fast.force_time()
    fast.profile_force(False)
    assert n == 5 and ms > 0
    # force + finish per step and one mass check after the upload; until the host has seen the mass flag
    # both force loops are enqueued (the one that does not apply exits at once)
    assert 11 <= fast.launch_count() - before <= 16
    assert fast.last_step_ms() >= ms * 0.9


# ------------------------------------------------------------------------------------------------
# data formats around the path: trajectory records straight from the device, exact restart
# ------------------------------------------------------------------------------------------------
def test_trajectory_from_device_positions_matches_reference_writer(nb, reference, strict, tmp_path):
    n, steps = 2500, 6
    p, eps = reference.disk(n, 77)
    params = nb.SimParams(1.0, 0.5, eps)
    ours, theirs = str(tmp_path / "ours.bin"), str(tmp_path / "ref.bin")
    w = nb.TrajectoryWriter(ours, p["mass"], steps)
    reference.checkpoint_begin(theirs, p, steps)
    strict.upload(p.copy())
    q = p
    for _ in range(steps):
        strict.step(params, 1)
        w.write_step(strict.download_positions())          # 12 B/particle, no AoS
        q = reference.naive_steps(q, 1.0, 0.5, eps, 1)
        reference.checkpoint_write_step(theirs, q)
    assert open(ours, "rb").read() == open(theirs, "rb").read()


def test_async_trajectory_records_equal_synchronous_ones(nb, reference, tmp_path):
    """Records taken with download_positions_async (two staging buffers, copy stream; the next step is
    queued before the record has left the device) build the same file, byte for byte, as the
    synchronous path and as the reference's CheckpointManager -- also when three records are requested
    back to back without waiting (the third makes the device, not the host, wait for the first)."""
    n, steps = 3000, 7
    p, eps = reference.disk(n, 78)
    params = nb.SimParams(1.0, 0.5, eps)
    files = {k: str(tmp_path / f"{k}.bin") for k in ("sync", "async", "ref")}
    sim = nb.CudaNaiveSimulation(mode="strict")
    # synchronous
    w = nb.TrajectoryWriter(files["sync"], p["mass"], steps)
    sim.upload(p.copy())
    for _ in range(steps):
        sim.step(params, 1)
        w.write_step(sim.download_positions())
    # asynchronous: every record is requested and the next step queued before anything is awaited
    bufs = [nb.PinnedArray((n, 3)) for _ in range(steps)]
    sim.upload(p.copy())
    for k in range(steps):
        sim.step(params, 1)
        sim.download_positions_async(bufs[k].array)
    sim.wait_records()
    w = nb.TrajectoryWriter(files["async"], p["mass"], steps)
    for b in bufs:
        w.write_step(b.array)
        b.close()
    sim.close()
    # the reference
    reference.checkpoint_begin(files["ref"], p, steps)
    q = p
    for _ in range(steps):
        q = reference.naive_steps(q, 1.0, 0.5, eps, 1)
        reference.checkpoint_write_step(files["ref"], q)
    data = {k: open(f, "rb").read() for k, f in files.items()}
    assert data["async"] == data["sync"] == data["ref"]


@pytest.mark.parametrize("mode", ["strict", "fast"])
def test_restart_record_resumes_bit_exactly(nb, strict, fast, mode, tmp_path):
    sim = strict if mode == "strict" else fast
    p, eps = nb.datasets.plummer_sphere(9000, seed=12)
    params = nb.SimParams(1.0, 0.5, eps)
    whole = p.copy()
    sim.upload(whole); sim.step(params, 12); sim.download(whole)
    half = p.copy()
    sim.upload(half); sim.step(params, 5); sim.download(half)
    path = str(tmp_path / "run.restart")
    nb.write_restart(path, half, params, 5)
    again, params2, step = nb.read_restart(path)
    sim.upload(again); sim.step(params2, 12 - step); sim.download(again)
    assert again.tobytes() == whole.tobytes()


# ------------------------------------------------------------------------------------------------
# optional FP64 variant: forces summed in double, state stays FP32 (reported separately in bench.py)
# ------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def fp64(nb):
    s = nb.CudaNaiveSimulation(mode="fp64")
    yield s
    s.close()


@pytest.mark.parametrize("kind,n", [("plummer", 5000), ("disk", 4096)])
def test_fp64_mode_matches_fp64_truth_to_one_float_rounding(nb, oracle, fp64, kind, n):
    p, eps = (nb.datasets.plummer_sphere if kind == "plummer" else nb.datasets.galaxy_disk)(n, seed=3)
    truth = oracle.accel_f64(p, 1.0, eps)
    got = run_compute_steps(fp64, nb, p, 1.0, 0.5, eps, 1)
    e = rel_err(got["acceleration"], truth)
    assert e.max() < 1.2e-7            # the only error left is the final rounding to binary32 (2^-24 per component)


def test_fast_mode_against_gpu_fp64_at_config3_size(nb, fast, fp64):
    """Accuracy of the fast kernel at N=262144 (BASELINE configs[2]), where the CPU truth is too slow:
    the FP64 GPU variant (itself pinned on the CPU FP64 oracle above) is the yardstick."""
    n = 262144
    p, eps = nb.datasets.plummer_sphere(n, seed=12345)
    params = nb.SimParams(1.0, 0.5, eps)
    a = run_compute_steps(fast, nb, p, 1.0, 0.5, eps, 1)["acceleration"]
    b = run_compute_steps(fp64, nb, p, 1.0, 0.5, eps, 1)["acceleration"]
    e = rel_err(a, b)
    print("fast vs fp64 at N=262144: max %.3g median %.3g" % (e.max(), np.median(e)))
    assert np.median(e) < 2e-6 and e.max() < 5e-5


def test_config5_size_round_trip_and_limits(nb, fast):
    """BASELINE configs[4] size through the boundary (no stepping: 1.8e13 interactions per step belong
    to the 8-GPU run): 4,194,304 particles up, back, and as a trajectory record; sizes past the slot
    index range are refused with a message, not attempted."""
    n = 4194304
    rng = np.random.default_rng(3)
    p = np.zeros(n, nb.PARTICLE_DTYPE)
    p["position"] = rng.standard_normal((n, 3)).astype(np.float32)
    p["velocity"] = rng.standard_normal((n, 3)).astype(np.float32)
    p["mass"] = rng.uniform(0.5, 2.0, n).astype(np.float32)
    fast.upload(p)
    out = np.zeros_like(p)
    fast.download(out)
    for k in ("position", "velocity", "mass"):
        assert same_bits(np.ascontiguousarray(out[k]), np.ascontiguousarray(p[k])), k
    assert same_bits(fast.download_positions(), np.ascontiguousarray(p["position"]))
    lib = nb.load_library()
    assert lib.nbody_cuda_configure(fast._h, (1 << 30) + 1, 0, 1) == -1
    assert b"2^30" in lib.nbody_cuda_last_error(fast._h)
    assert lib.nbody_cuda_configure(fast._h, 100, 3, 2) == -1           # rank >= world
    fast.n_total = None                                                  # the handle is reconfigured by the next upload
