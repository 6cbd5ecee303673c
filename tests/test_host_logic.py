"""CPU: host-side logic of the package -- slice arithmetic, slot layout, generators, argument checks."""
import numpy as np
import pytest


def test_slot_layout(nb):
    lay = nb.SlotLayout(10, 3)
    assert lay.counts == [4, 3, 3] and lay.displs == [0, 4, 7]
    assert lay.slot_stride == 256 and lay.n_slots == 768
    assert [lay.slot_of(i) for i in (0, 3, 4, 6, 7, 9)] == [0, 3, 256, 258, 512, 514]
    assert [lay.owner_of(i) for i in (0, 3, 4, 9)] == [0, 0, 1, 2]
    big = nb.SlotLayout(1048576, 8)
    assert big.counts == [131072] * 8 and big.slot_stride == 131072 and big.n_slots == 1048576
    odd = nb.SlotLayout(1000003, 8)
    assert sum(odd.counts) == 1000003 and odd.slot_stride % 256 == 0 and odd.slot_stride >= max(odd.counts)
    assert nb.SlotLayout(0, 2).slot_stride == 256


def test_slice_counts_match_reference_rule(nb):
    # src/apps/headless_main.cpp:249-259
    assert nb.slice_counts(10, 4) == ([3, 3, 2, 2], [0, 3, 6, 8])
    assert nb.slice_counts(8, 8) == ([1] * 8, list(range(8)))
    assert nb.slice_counts(3, 4) == ([1, 1, 1, 0], [0, 1, 2, 3])


def test_simparams_defaults(nb):
    p = nb.SimParams()            # include/SimulationAlgorithm.h:4-9
    assert (p.G, p.dt, p.min_r2) == (1.0, 0.05, 1e-8)


def test_plummer_dataset_shape(nb):
    p, eps = nb.datasets.plummer_sphere(20000, seed=1)
    assert p.dtype == nb.PARTICLE_DTYPE and p.shape == (20000,)
    assert np.isfinite(p["position"]).all() and np.isfinite(p["velocity"]).all()
    assert np.all(p["mass"] == 1.0)
    r = np.linalg.norm(p["position"].astype(np.float64), axis=1)
    assert r.min() >= 10.0                                   # u >= 0.001  =>  r >= a / sqrt(99)
    assert 100.0 < np.median(r) < 160.0                      # half-mass radius of a Plummer sphere ~ 1.305 a
    assert abs(p["velocity"].astype(np.float64).mean(axis=0)).max() < 1e-3
    # src/Simulation.cpp:186-195 at n = 65536
    assert nb.datasets.plummer_sphere(65536, seed=1)[1] == pytest.approx(2.07178, rel=1e-4)


def test_disk_dataset_shape(nb):
    p, eps = nb.datasets.galaxy_disk(4096, seed=2)
    assert p["mass"][0] == 1.0e6 and np.all(p["position"][0] == 0)
    assert eps == pytest.approx(2.28385, rel=1e-4)
    r = np.hypot(p["position"][1:, 0], p["position"][1:, 2])
    assert r.max() <= 300.0 and 0.5 <= p["mass"][1:].min() and p["mass"][1:].max() < 2.0


def test_particles_argument_is_checked(nb):
    from nbody_b200.simulation import _as_particles

    with pytest.raises(TypeError):
        _as_particles(np.zeros((4, 10), np.float32))
    with pytest.raises(ValueError):
        _as_particles(np.zeros(8, nb.PARTICLE_DTYPE)[::2])
    assert _as_particles(np.zeros(0, nb.PARTICLE_DTYPE)).shape == (0,)
