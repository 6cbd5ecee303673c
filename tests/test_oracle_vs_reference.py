"""CPU: pins the oracle on the reference's own compiled classes (oracle/_ref), beyond the small
committed golden vectors.  Skipped where oracle/_ref was never built (needs /root/reference at
build time; the built .so travels to the GPU box)."""
import os

import numpy as np
import pytest


def same_bits(a, b):
    return a.tobytes() == b.tobytes()


def test_layout_probes(reference):
    lay = reference.layout()
    assert lay == dict(sizeof_particle=40, velocity=12, acceleration=24, mass=36, sizeof_simparams=12, sizeof_header=24)
    assert reference.default_simparams() == pytest.approx((1.0, 0.05, 1e-8))


def test_generators_are_seed_reproducible(reference):
    a, ea = reference.disk(1000, 5)
    b, eb = reference.disk(1000, 5)
    c, _ = reference.disk(1000, 6)
    assert same_bits(a, b) and ea == eb and not same_bits(a, c)
    assert a["mass"][0] == 1.0e6 and np.all(a["position"][0] == 0)       # Simulation.cpp:93-98
    p, eps = reference.plummer(2000, 5)
    assert np.isfinite(p["position"]).all() and np.all(p["mass"] == 1.0)
    # min_r2 at the configs' sizes
    assert reference.disk(4096, 1)[1] == pytest.approx(2.28385, rel=1e-5)
    assert reference.plummer(4096, 1)[1] == pytest.approx(13.155, rel=1e-4)


@pytest.mark.parametrize("kind,n,steps", [("disk", 1500, 3), ("plummer", 2048, 3), ("clusters", 1024, 4)])
def test_oracle_equals_reference_naive(reference, oracle, kind, n, steps):
    if kind == "disk":
        p, eps = reference.disk(n, 12345)
    elif kind == "plummer":
        p, eps = reference.plummer(n, 12345)
    else:
        p, eps = reference.clusters(4, n // 4, 12345), 2.0
    want = reference.naive_steps(p, 1.0, 0.5, eps, steps)
    assert same_bits(oracle.step_pairs(p, 1.0, 0.5, eps, steps), want)
    assert same_bits(oracle.step_rows(p, 1.0, 0.5, eps, steps), want)


def test_config1_disk_4096_100_steps(reference, oracle):
    """BASELINE.json configs[0]: sequential NaiveSimulation, N=4096 disk, 100 steps -- the row-order
    oracle reproduces the reference trajectory bit for bit."""
    p, eps = reference.disk(4096, 12345)
    assert eps == pytest.approx(2.28385, rel=1e-5)
    want = reference.naive_steps(p, 1.0, 0.5, eps, 100)
    got = oracle.step_rows(p, 1.0, 0.5, eps, 100)
    assert same_bits(got, want)


@pytest.mark.parametrize("nranks", [1, 2, 5, 8])
def test_oracle_equals_reference_mpi_order(reference, oracle, nranks):
    p, eps = reference.disk(1003, 77)          # n % nranks != 0 for 2, 5, 8
    want = reference.mpi_naive_steps(p, 1.0, 0.5, eps, 2, nranks=nranks)
    assert same_bits(oracle.step_mpi_order(p, nranks, 1.0, 0.5, eps, 2), want)


def test_mpi_equals_seq_on_equal_masses(reference):
    """with G = 1 and all masses 1 (Plummer) the two groupings coincide for np=1."""
    p, eps = reference.plummer(1024, 3)
    a = reference.naive_steps(p, 1.0, 0.5, eps, 1)
    b = reference.mpi_naive_steps(p, 1.0, 0.5, eps, 1, nranks=1)
    assert same_bits(a["acceleration"], b["acceleration"])


def test_facade_step_equals_algorithm(reference):
    p, eps = reference.plummer(512, 9)
    assert same_bits(reference.naive_steps(p, 1.0, 0.5, eps, 2), reference.naive_steps(p, 1.0, 0.5, eps, 2, facade=True))


def test_checkpoint_round_trip(reference, tmp_path):
    """Trajectory file of src/CheckpointManager.cpp: 24-byte header, masses, then xyz per record."""
    p, eps = reference.plummer(300, 11)
    path = str(tmp_path / "simulation_output.bin")
    reference.checkpoint_begin(path, p, 5)
    states = []
    q = p
    for _ in range(3):
        q = reference.naive_steps(q, 1.0, 0.5, eps, 1)
        reference.checkpoint_write_step(path, q)
        states.append(q)
    assert os.path.getsize(path) == 24 + 4 * 300 + 12 * 300 * 3
    assert reference.checkpoint_read_header(path) == (300, 5, 3)
    assert same_bits(reference.checkpoint_read_masses(path, 300), np.ascontiguousarray(p["mass"]))
    for k, s in enumerate(states):
        pos, got = reference.checkpoint_read_step(path, 300, k)
        assert got == 300 and same_bits(pos, np.ascontiguousarray(s["position"]))
    assert reference.checkpoint_read_step(path, 300, 3)[1] == 0
