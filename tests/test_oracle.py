"""CPU: the oracle (oracle/nbody_oracle.c) against the golden vectors that the reference itself
produced (tests/golden/make_golden.py), plus its internal consistency.  Bit-exact throughout."""
import numpy as np
import pytest

from conftest import GOLDEN_CASES, load_golden


def same_bits(a, b):
    return a.tobytes() == b.tobytes()


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_pairs_order_matches_reference_golden(oracle, case):
    g = load_golden(case)
    for k in [k for k in g if k.startswith("naive_step")]:
        steps = int(k[len("naive_step"):])
        got = oracle.step_pairs(g["initial"], float(g["G"]), float(g["dt"]), float(g["min_r2"]), steps)
        assert same_bits(got, g[k]), (case, k)


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_row_order_matches_reference_golden(oracle, case):
    """one accumulator per target, sources ascending, skip self == NaiveSimulation, bit for bit."""
    g = load_golden(case)
    for k in [k for k in g if k.startswith("naive_step")]:
        steps = int(k[len("naive_step"):])
        got = oracle.step_rows(g["initial"], float(g["G"]), float(g["dt"]), float(g["min_r2"]), steps)
        assert same_bits(got, g[k]), (case, k)


@pytest.mark.parametrize("case", GOLDEN_CASES)
@pytest.mark.parametrize("nranks", [1, 3, 4])
def test_mpi_order_matches_reference_golden(oracle, case, nranks):
    g = load_golden(case)
    got = oracle.step_mpi_order(g["initial"], nranks, float(g["G"]), float(g["dt"]), float(g["min_r2"]), 2)
    assert same_bits(got, g[f"mpi_np{nranks}_step2"])


def test_layout_facts():
    lay = load_golden("layout")
    assert int(lay["sizeof_particle"]) == 40
    assert (int(lay["velocity"]), int(lay["acceleration"]), int(lay["mass"])) == (12, 24, 36)
    assert int(lay["sizeof_simparams"]) == 12 and int(lay["sizeof_header"]) == 24
    np.testing.assert_allclose(lay["default_simparams"], [1.0, 0.05, 1e-8], rtol=1e-6)


def test_two_body_closed_form(oracle, nb):
    p = np.zeros(2, nb.PARTICLE_DTYPE)
    p["position"][1] = [3.0, 4.0, 0.0]
    p["mass"] = [2.0, 5.0]
    G, eps2 = 1.5, 0.0
    a = oracle.accel_rows(p, G, eps2)
    # |r| = 5: a_0 = G m_1 r / 125, a_1 = -G m_0 r / 125
    np.testing.assert_allclose(a[0], np.array([3, 4, 0]) * G * 5.0 / 125.0, rtol=1e-6)
    np.testing.assert_allclose(a[1], -np.array([3, 4, 0]) * G * 2.0 / 125.0, rtol=1e-6)
    a64 = oracle.accel_f64(p, G, eps2)
    np.testing.assert_allclose(a64, a, rtol=1e-6)


def test_row_slices_are_independent(oracle):
    g = load_golden("plummer_n512")
    p, G, eps2 = g["initial"], float(g["G"]), float(g["min_r2"])
    full = oracle.accel_rows(p, G, eps2)
    parts = [oracle.accel_rows(p, G, eps2, a, b) for a, b in [(0, 100), (100, 101), (101, 512)]]
    assert same_bits(np.concatenate(parts), full)


def test_slice_bounds_follow_headless_rule(oracle, nb):
    for n in [0, 1, 7, 256, 1000, 65536, 1048577]:
        for P in [1, 2, 3, 4, 8]:
            counts, displs = nb.slice_counts(n, P)
            assert sum(counts) == n
            for r in range(P):
                assert oracle.slice_bounds(n, P, r) == (displs[r], counts[r])


def test_energy_against_numpy(oracle):
    g = load_golden("solar_9")
    p, G, eps2 = g["initial"], float(g["G"]), float(g["min_r2"])
    k, u = oracle.energy(p, G, eps2)
    pos = p["position"].astype(np.float64)
    m = p["mass"].astype(np.float64)
    v = p["velocity"].astype(np.float64)
    k_np = 0.5 * np.sum(m * np.sum(v * v, axis=1))
    u_np = 0.0
    for i in range(len(p)):
        for j in range(i + 1, len(p)):
            u_np -= G * m[i] * m[j] / np.sqrt(np.sum((pos[i] - pos[j]) ** 2) + eps2)
    assert k == pytest.approx(k_np, rel=1e-12) and u == pytest.approx(u_np, rel=1e-12)


def test_fp32_error_against_fp64_truth_is_small(oracle):
    """States the yardstick used for the fast kernel: the reference's own FP32 sum vs FP64."""
    from conftest import rel_err

    g = load_golden("plummer_n512")
    p, G, eps2 = g["initial"], float(g["G"]), float(g["min_r2"])
    e = rel_err(oracle.accel_rows(p, G, eps2), oracle.accel_f64(p, G, eps2))
    assert e.max() < 1e-5
