"""The data formats on either side of the hot path.

1. The reference's trajectory file, written by ``CheckpointManager`` (src/CheckpointManager.cpp)
   and read by its viewer and ``checkpoint_to_vtk``: little-endian
   ``[u64 n_particles][u64 target_steps][u64 passed_steps]`` (include/SimulationOutput.h:8-13),
   ``float mass[n]`` (CheckpointManager.cpp:27-48), then one ``float xyz[n][3]`` record per saved
   step (:76-104), the header's ``passed_steps`` being rewritten after every record (:50-74,106).
   ``TrajectoryWriter`` / ``TrajectoryReader`` produce and consume byte-identical files
   (tests/test_trajectory.py checks that against the reference's own writer), taking positions as
   an (n,3) float32 array -- what ``CudaNaiveSimulation.download_positions`` returns without ever
   building the 40-byte AoS records the reference's API wants.

2. A restart record, which the reference lacks: its file has no velocities, so a run cannot be
   resumed from it.  ``write_restart`` / ``read_restart`` keep a sidecar with the
   full particle state, SimParams and the step index; resuming from it reproduces the
   uninterrupted run bit for bit (the kernels are deterministic).
"""
from __future__ import annotations

import os
import struct

import numpy as np

from .simulation import PARTICLE_DTYPE, SimParams

HEADER = struct.Struct("<QQQ")          # include/SimulationOutput.h:8-13 (size_t x3 on LP64)
RESTART_MAGIC = b"NBODYRST"
RESTART_HEADER = struct.Struct("<8sQQfffI")   # magic, n, step, G, dt, min_r2, record bytes


class TrajectoryWriter:
    """Writes the reference's trajectory format.  Like the reference, every call reopens the file."""

    def __init__(self, path: str, masses: np.ndarray, target_steps: int):
        self.path = path
        m = np.ascontiguousarray(masses, dtype="<f4")
        self.n = int(m.shape[0])
        self.passed = 0
        with open(path, "wb") as f:                      # write_header truncates (CheckpointManager.cpp:12)
            f.write(HEADER.pack(self.n, int(target_steps), 0))
            f.write(m.tobytes())                         # write_masses appends (:27-48)
        self.target_steps = int(target_steps)

    def write_step(self, positions: np.ndarray) -> None:
        p = np.ascontiguousarray(positions, dtype="<f4")
        if p.shape != (self.n, 3):
            raise ValueError(f"positions must have shape ({self.n}, 3)")
        with open(self.path, "r+b") as f:
            f.seek(0, os.SEEK_END)
            f.write(p.tobytes())                         # :76-104
            self.passed += 1
            f.seek(0)
            f.write(HEADER.pack(self.n, self.target_steps, self.passed))   # increment_passed_steps (:50-74)


class TrajectoryReader:
    def __init__(self, path: str):
        self.path = path
        with open(path, "rb") as f:
            self.n, self.target_steps, self.passed_steps = HEADER.unpack(f.read(HEADER.size))

    def masses(self) -> np.ndarray:
        with open(self.path, "rb") as f:
            f.seek(HEADER.size)
            return np.frombuffer(f.read(4 * self.n), dtype="<f4").copy()

    def step(self, index: int):
        """(n,3) positions of record `index`, or None past the end (read_step returns 0, :144-177)."""
        with open(self.path, "rb") as f:
            n, _, passed = HEADER.unpack(f.read(HEADER.size))
            if index >= passed:
                return None
            f.seek(HEADER.size + 4 * n + 12 * n * index)
            return np.frombuffer(f.read(12 * n), dtype="<f4").reshape(n, 3).copy()


def write_restart(path: str, particles: np.ndarray, params: SimParams, step: int) -> None:
    """Full state (40-byte records as include/Particle.hpp:53-59) + SimParams + step index."""
    p = np.ascontiguousarray(particles)
    if p.dtype != PARTICLE_DTYPE:
        raise TypeError("particles must be PARTICLE_DTYPE")
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(RESTART_HEADER.pack(RESTART_MAGIC, p.shape[0], int(step), params.G, params.dt, params.min_r2,
                                    PARTICLE_DTYPE.itemsize))
        f.write(p.tobytes())
    os.replace(tmp, path)                                # never leaves a half-written record behind


def read_restart(path: str):
    """Returns (particles, SimParams, step)."""
    with open(path, "rb") as f:
        magic, n, step, G, dt, eps, rec = RESTART_HEADER.unpack(f.read(RESTART_HEADER.size))
        if magic != RESTART_MAGIC or rec != PARTICLE_DTYPE.itemsize:
            raise ValueError(f"{path} is not a restart record")
        data = f.read(n * rec)
        if len(data) != n * rec:
            raise ValueError(f"{path} is truncated")
    return np.frombuffer(data, dtype=PARTICLE_DTYPE).copy(), SimParams(G=G, dt=dt, min_r2=eps), int(step)


def write_ppm(path: str, frame: np.ndarray) -> None:
    """Binary PPM (P6) of an (h, w, 3) uint8 frame, e.g. from CudaNaiveSimulation.render_frame."""
    f = np.ascontiguousarray(frame, dtype=np.uint8)
    with open(path, "wb") as out:
        out.write(b"P6\n%d %d\n255\n" % (f.shape[1], f.shape[0]))
        out.write(f.tobytes())
