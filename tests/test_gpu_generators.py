"""GPU: on-device input generators (nbody_cuda_generate).  They follow the reference
generators' distributions (src/Simulation.cpp:88-136, :173-267) from a different random stream, so
the checks are statistical -- against the reference's own generator where the compiled reference
is available, else against this repo's NumPy generator -- plus determinism and slice-independence."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def device_plummer(nb, n, seed, world=1, rank=0):
    sim = nb.CudaNaiveSimulation(mode="fast", exchange="nccl")     # "nccl": configure() does not look for peers
    if world > 1:
        sim.rank, sim.world = rank, world                          # a slice handle without a process group
    eps = sim.generate("plummer", n, 100.0, -1.0, 1.0, seed)
    counts, _ = nb.slice_counts(n, world)
    out = np.zeros(counts[rank], nb.PARTICLE_DTYPE)
    sim.download(out)
    sim.close()
    return out, eps


def test_plummer_statistics_match_the_reference_generator(nb):
    from oracle import Reference

    n = 200000
    got, eps = device_plummer(nb, n, 1)
    if Reference.available():
        want, eps_ref = Reference().plummer(n, 1)
    else:
        want, eps_ref = nb.datasets.plummer_sphere(n, seed=1)
    assert eps == pytest.approx(eps_ref, rel=1e-5)
    assert np.all(got["mass"] == 1.0) and np.isfinite(got["position"]).all() and np.isfinite(got["velocity"]).all()
    rg, rw = np.linalg.norm(got["position"], axis=1), np.linalg.norm(want["position"], axis=1)
    vg, vw = np.linalg.norm(got["velocity"], axis=1), np.linalg.norm(want["velocity"], axis=1)
    qs = [0.05, 0.25, 0.5, 0.75, 0.95]
    np.testing.assert_allclose(np.quantile(rg, qs), np.quantile(rw, qs), rtol=0.03)     # radial profile
    np.testing.assert_allclose(np.quantile(vg, qs), np.quantile(vw, qs), rtol=0.03)     # speed distribution
    assert rg.min() >= 10.0                                                              # u >= 0.001
    assert np.abs(got["velocity"].astype(np.float64).mean(axis=0)).max() < 1e-3          # bulk velocity removed
    assert np.abs(got["position"].astype(np.float64).mean(axis=0)).max() < 3.0           # isotropic


def test_disk_statistics(nb):
    n = 100000
    sim = nb.CudaNaiveSimulation(mode="fast")
    eps = sim.generate("disk", n, 300.0, 50.0, 1.0, 9)
    p = np.zeros(n, nb.PARTICLE_DTYPE)
    sim.download(p)
    sim.close()
    want, eps_np = nb.datasets.galaxy_disk(n, seed=9)
    assert eps == pytest.approx(eps_np, rel=1e-5)
    assert p["mass"][0] == 1.0e6 and np.all(p["position"][0] == 0) and np.all(p["velocity"][0] == 0)
    r = np.hypot(p["position"][1:, 0], p["position"][1:, 2]).astype(np.float64)
    assert r.max() <= 300.0 and np.median(r) == pytest.approx(300.0 / np.sqrt(2.0), rel=0.02)   # r = R sqrt(U)
    assert p["position"][1:, 1].std() == pytest.approx(15.0, rel=0.03)                           # y ~ N(0, 0.3 T)
    vt = (-p["velocity"][1:, 0] * p["position"][1:, 2] + p["velocity"][1:, 2] * p["position"][1:, 0]) / r
    np.testing.assert_allclose(np.median(vt * np.sqrt(r)), 1000.0, rtol=0.02)                    # v_orb = sqrt(G 1e6 / r)
    m = p["mass"][1:]
    assert 0.5 <= m.min() and m.max() < 2.0 and m.mean() == pytest.approx(1.25, rel=0.01)


def test_generation_is_deterministic_and_slice_independent(nb):
    n = 30001
    a, _ = device_plummer(nb, n, 77)
    b, _ = device_plummer(nb, n, 77)
    c, _ = device_plummer(nb, n, 78)
    assert a.tobytes() == b.tobytes() and a.tobytes() != c.tobytes()
    counts, displs = nb.slice_counts(n, 3)
    for r in range(3):
        part, _ = device_plummer(nb, n, 77, world=3, rank=r)
        assert part.tobytes() == a[displs[r]:displs[r] + counts[r]].tobytes()


def test_generated_system_steps_like_an_uploaded_one(nb):
    n = 20000
    sim = nb.CudaNaiveSimulation(mode="fast")
    eps = sim.generate("plummer", n, 100.0, -1.0, 1.0, 5)
    start = np.zeros(n, nb.PARTICLE_DTYPE)
    sim.download(start)
    params = nb.SimParams(1.0, 0.5, eps)
    sim.step(params, 3)
    a = np.zeros(n, nb.PARTICLE_DTYPE)
    sim.download(a)
    b = start.copy()
    sim.upload(b); sim.step(params, 3); sim.download(b)
    sim.close()
    assert a.tobytes() == b.tobytes()
