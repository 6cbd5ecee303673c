"""Synthetic particle sets for benchmarks and examples (NumPy, seeded).

These follow the distributions of the reference's generators so that the named workloads have
the reference's shape, but they draw from NumPy's generator, not from the reference's
``std::mt19937`` stream: they are NOT bit-identical to the reference's inputs.  Parity tests use
inputs produced by the reference's own generators (tests/golden/, oracle/_ref) instead.
Throughput of the direct sum does not depend on the particle values.
"""
from __future__ import annotations

import numpy as np

from .simulation import PARTICLE_DTYPE


def _on_sphere(rng, radius):
    """Uniform directions scaled by ``radius`` (include/Random.hpp:51-67, Marsaglia)."""
    n = radius.shape[0]
    z = rng.uniform(-1.0, 1.0, n)
    phi = rng.uniform(0.0, 2.0 * np.pi, n)
    s = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    return np.stack([radius * s * np.cos(phi), radius * s * np.sin(phi), radius * z], axis=1)


def plummer_sphere(n: int, scale_radius: float = 100.0, total_mass: float | None = None, G: float = 1.0,
                   seed: int = 12345):
    """Plummer model as src/Simulation.cpp:173-267.  Returns (particles, min_r2)."""
    rng = np.random.default_rng(seed)
    if total_mass is None or total_mass < 0:
        total_mass = float(n)
    a = float(scale_radius)
    # softening: 15 % of the mean spacing inside 2.4 a  (Simulation.cpp:186-195)
    volume = (4.0 / 3.0) * np.pi * (2.4 * a) ** 3
    softening = 0.15 * np.cbrt(volume / n)
    min_r2 = float(np.float32(softening) ** 2)
    u = rng.uniform(0.001, 1.0, n)                                  # :216
    r = a / np.sqrt(u ** (-2.0 / 3.0) - 1.0)                        # :217
    pos = _on_sphere(rng, r)
    # speed ratio q = v / v_esc by rejection on q^2 (1-q^2)^3.5  (:229-242)
    q = np.empty(n)
    todo = np.arange(n)
    while todo.size:
        x = rng.uniform(0.0, 1.0, todo.size)
        y = rng.uniform(0.0, 0.1, todo.size)
        ok = y < x * x * (1.0 - x * x) ** 3.5
        q[todo[ok]] = x[ok]
        todo = todo[~ok]
    v_esc = np.sqrt(2.0 * G * total_mass / np.sqrt(r * r + a * a))  # :245-246
    vel = _on_sphere(rng, q * v_esc)
    vel -= vel.mean(axis=0)                                          # equal masses: bulk velocity (:256-265)
    p = np.zeros(n, PARTICLE_DTYPE)
    p["position"] = pos.astype(np.float32)
    p["velocity"] = vel.astype(np.float32)
    p["mass"] = np.float32(total_mass / n)
    return p, min_r2


def galaxy_disk(n: int, radius: float = 300.0, thickness: float = 50.0, G: float = 1.0, seed: int = 12345):
    """Central mass + thin disk as src/Simulation.cpp:88-136.  Returns (particles, min_r2)."""
    rng = np.random.default_rng(seed)
    central = 1.0e6
    p = np.zeros(n, PARTICLE_DTYPE)
    if n > 0:
        p["mass"][0] = central
    m = max(0, n - 1)
    r = radius * np.sqrt(rng.uniform(0.0, 1.0, m))
    r = np.maximum(r, 1e-3)
    th = rng.uniform(0.0, 2.0 * np.pi, m)
    y = rng.normal(0.0, thickness * 0.3, m)
    v_orb = np.sqrt(G * central / r)
    disp = 0.1 * v_orb
    p["position"][1:] = np.stack([r * np.cos(th), y, r * np.sin(th)], axis=1).astype(np.float32)
    p["velocity"][1:] = np.stack(
        [-v_orb * np.sin(th) + rng.normal(0.0, 1.0, m) * disp,
         rng.normal(0.0, 1.0, m) * disp * 0.5,
         v_orb * np.cos(th) + rng.normal(0.0, 1.0, m) * disp], axis=1).astype(np.float32)
    p["mass"][1:] = rng.uniform(0.5, 2.0, m).astype(np.float32)
    spacing = np.cbrt(np.pi * radius * radius * thickness / max(n, 1))   # :126-131
    min_r2 = float(np.float32(0.1 * spacing) ** 2)
    return p, min_r2


def uniform_random(n: int, W: float = 2400.0, H: float = 2400.0, D: float = 2400.0, min_mass: float = 1.0,
                   max_mass: float = 1000.0, seed: int = 12345):
    """Uniform box, zero velocity, as src/Simulation.cpp:19-39 (World.h: 2400^3)."""
    rng = np.random.default_rng(seed)
    p = np.zeros(n, PARTICLE_DTYPE)
    p["position"] = (rng.uniform(0.0, 1.0, (n, 3)) * np.array([W, H, D])).astype(np.float32)
    p["mass"] = rng.uniform(min_mass, max_mass, n).astype(np.float32)
    return p
