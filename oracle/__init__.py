"""ctypes bindings for the CHECKERS: ``liboracle.so`` (this repo's C restatement,
``oracle/nbody_oracle.c``) and ``_ref/libnbody_ref.so`` (the reference's own classes,
compiled unmodified by ``oracle/Makefile``).

TEST INFRASTRUCTURE ONLY.  Importable from ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- never from the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# include/Particle.hpp:53-59 -- 40 bytes, offsets 0/12/24/36
PARTICLE_DTYPE = np.dtype(
    [("position", "<f4", (3,)), ("velocity", "<f4", (3,)), ("acceleration", "<f4", (3,)), ("mass", "<f4")]
)
assert PARTICLE_DTYPE.itemsize == 40


def build(verbose: bool = False) -> None:
    """Run ``make`` in oracle/ (builds liboracle.so; and _ref/ when /root/reference exists)."""
    r = subprocess.run(["make", "-C", HERE], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout, r.stderr)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed")


def _ptr(a: np.ndarray):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """This repo's C restatement (row order, literal pair order, MPI order, FP64 truth, energy)."""

    def __init__(self, path: str | None = None):
        path = path or os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        self.lib = L = C.CDLL(path)
        f, sz, vp, i, d = C.c_float, C.c_size_t, C.c_void_p, C.c_int, C.c_double
        L.oracle_sizeof_particle.restype = i
        L.oracle_naive_accel_pairs_f32.argtypes = [vp, sz, f, f]
        L.oracle_naive_accel_rows_f32.argtypes = [vp, sz, sz, sz, f, f, vp]
        L.oracle_integrate_f32.argtypes = [vp, sz, f]
        L.oracle_naive_step_pairs_f32.argtypes = [vp, sz, f, f, f, i]
        L.oracle_naive_step_rows_f32.argtypes = [vp, sz, f, f, f, i]
        L.oracle_mpi_naive_accel_f32.argtypes = [vp, sz, i, f, f]
        L.oracle_mpi_naive_step_f32.argtypes = [vp, sz, i, f, f, f, i]
        L.oracle_accel_f64.argtypes = [vp, sz, sz, sz, d, d, vp]
        L.oracle_step_f64.argtypes = [vp, vp, vp, vp, sz, d, d, d, i]
        L.oracle_energy_f64.argtypes = [vp, sz, d, d, C.POINTER(d), C.POINTER(d)]
        L.oracle_slice_bounds.argtypes = [sz, i, i, C.POINTER(sz), C.POINTER(sz)]
        assert L.oracle_sizeof_particle() == 40

    # each returns a NEW array; inputs are not modified
    def step_pairs(self, p, G, dt, eps2, steps=1):
        q = np.ascontiguousarray(p).copy()
        self.lib.oracle_naive_step_pairs_f32(_ptr(q), q.shape[0], G, dt, eps2, steps)
        return q

    def step_rows(self, p, G, dt, eps2, steps=1):
        q = np.ascontiguousarray(p).copy()
        self.lib.oracle_naive_step_rows_f32(_ptr(q), q.shape[0], G, dt, eps2, steps)
        return q

    def step_mpi_order(self, p, P, G, dt, eps2, steps=1):
        q = np.ascontiguousarray(p).copy()
        self.lib.oracle_mpi_naive_step_f32(_ptr(q), q.shape[0], P, G, dt, eps2, steps)
        return q

    def accel_rows(self, p, G, eps2, i0=0, i1=None):
        p = np.ascontiguousarray(p)
        n = p.shape[0]
        i1 = n if i1 is None else i1
        out = np.zeros((i1 - i0, 3), np.float32)
        self.lib.oracle_naive_accel_rows_f32(_ptr(p), n, i0, i1, G, eps2, _ptr(out))
        return out

    def accel_f64(self, p, G, eps2, i0=0, i1=None):
        p = np.ascontiguousarray(p)
        n = p.shape[0]
        i1 = n if i1 is None else i1
        out = np.zeros((i1 - i0, 3), np.float64)
        self.lib.oracle_accel_f64(_ptr(p), n, i0, i1, G, eps2, _ptr(out))
        return out

    def step_f64(self, p, G, dt, eps2, steps=1):
        """FP64 trajectory from float inputs; returns (pos, vel, acc) as float64 (n,3) arrays."""
        pos = np.ascontiguousarray(p["position"], dtype=np.float64)
        vel = np.ascontiguousarray(p["velocity"], dtype=np.float64)
        acc = np.zeros_like(pos)
        mass = np.ascontiguousarray(p["mass"], dtype=np.float64)
        self.lib.oracle_step_f64(_ptr(pos), _ptr(vel), _ptr(acc), _ptr(mass), pos.shape[0], G, dt, eps2, steps)
        return pos, vel, acc

    def energy(self, p, G, eps2):
        p = np.ascontiguousarray(p)
        k, u = C.c_double(), C.c_double()
        self.lib.oracle_energy_f64(_ptr(p), p.shape[0], G, eps2, C.byref(k), C.byref(u))
        return k.value, u.value

    def slice_bounds(self, n, P, r):
        b, c = C.c_size_t(), C.c_size_t()
        self.lib.oracle_slice_bounds(n, P, r, C.byref(b), C.byref(c))
        return b.value, c.value


class Reference:
    """The reference's own code (oracle/_ref/libnbody_ref.so).  ``available()`` is False when the
    prebuilt library is absent (e.g. a checkout that never ran oracle/Makefile next to
    /root/reference)."""

    PATH = os.path.join(HERE, "_ref", "libnbody_ref.so")

    @classmethod
    def available(cls) -> bool:
        return os.path.exists(cls.PATH)

    def __init__(self):
        if not self.available():
            raise FileNotFoundError(self.PATH)
        self.lib = L = C.CDLL(self.PATH)
        f, sz, vp, i, d, lg = C.c_float, C.c_size_t, C.c_void_p, C.c_int, C.c_double, C.c_long
        for n in ("ref_sizeof_particle", "ref_offsetof_velocity", "ref_offsetof_acceleration",
                  "ref_offsetof_mass", "ref_sizeof_simparams", "ref_sizeof_output_header"):
            getattr(L, n).restype = i
        L.ref_default_simparams.argtypes = [vp]
        L.ref_seed.argtypes = [C.c_uint]
        L.ref_generate_disk.argtypes = [i, f, f, f, vp, lg, C.POINTER(f)]
        L.ref_generate_disk.restype = lg
        L.ref_generate_plummer.argtypes = [i, f, f, f, vp, lg, C.POINTER(f)]
        L.ref_generate_plummer.restype = lg
        L.ref_generate_clusters.argtypes = [i, i, f, vp, lg]
        L.ref_generate_clusters.restype = lg
        L.ref_setup_solar.argtypes = [i, i, i, f, vp, lg]
        L.ref_setup_solar.restype = lg
        for n in ("ref_naive_steps", "ref_facade_steps"):
            getattr(L, n).argtypes = [vp, sz, f, f, f, i]
            getattr(L, n).restype = d
        L.ref_mpi_naive_steps.argtypes = [vp, sz, f, f, f, i, i]
        L.ref_mpi_naive_steps.restype = d
        L.ref_mpi_headless_loop.argtypes = [vp, sz, f, f, f, i, i]
        L.ref_mpi_headless_loop.restype = d
        L.ref_checkpoint_begin.argtypes = [C.c_char_p, vp, sz, sz]
        L.ref_checkpoint_write_step.argtypes = [C.c_char_p, vp, sz]
        L.ref_checkpoint_read_header.argtypes = [C.c_char_p, vp]
        L.ref_checkpoint_read_masses.argtypes = [C.c_char_p, vp, sz]
        L.ref_checkpoint_read_step.argtypes = [C.c_char_p, vp, sz]
        L.ref_checkpoint_read_step.restype = sz

    def layout(self):
        L = self.lib
        return dict(sizeof_particle=L.ref_sizeof_particle(), velocity=L.ref_offsetof_velocity(),
                    acceleration=L.ref_offsetof_acceleration(), mass=L.ref_offsetof_mass(),
                    sizeof_simparams=L.ref_sizeof_simparams(), sizeof_header=L.ref_sizeof_output_header())

    def default_simparams(self):
        out = np.zeros(3, np.float32)
        self.lib.ref_default_simparams(_ptr(out))
        return tuple(float(x) for x in out)

    # generators -- return (particles, min_r2).  `acceleration` is whatever the generator left
    # there (uninitialised for disk/cluster/Plummer, Simulation.cpp:101,161,208): zeroed here.
    def _gen(self, fn, n, *args, with_eps=True):
        out = np.zeros(n, PARTICLE_DTYPE)
        eps = C.c_float(0.0)
        if with_eps:
            got = fn(*args, _ptr(out), n, C.byref(eps))
        else:
            got = fn(*args, _ptr(out), n)
        assert got == n, (got, n)
        out["acceleration"] = 0.0
        return out, float(eps.value)

    def disk(self, n, seed, radius=300.0, thickness=50.0, G=1.0):
        self.lib.ref_seed(seed)
        return self._gen(self.lib.ref_generate_disk, n, n, radius, thickness, G)

    def plummer(self, n, seed, scale_radius=100.0, total_mass=None, G=1.0):
        self.lib.ref_seed(seed)
        tm = float(n) if total_mass is None else total_mass
        return self._gen(self.lib.ref_generate_plummer, n, n, scale_radius, tm, G)

    def clusters(self, n_clusters, per_cluster, seed, separation=500.0):
        self.lib.ref_seed(seed)
        p, _ = self._gen(self.lib.ref_generate_clusters, n_clusters * per_cluster, n_clusters, per_cluster,
                         separation, with_eps=False)
        return p

    def solar(self, W=2400, H=2400, D=2400, G=1.0):
        p, _ = self._gen(self.lib.ref_setup_solar, 9, W, H, D, G, with_eps=False)
        return p

    def naive_steps(self, p, G, dt, eps2, steps=1, timed=False, facade=False):
        q = np.ascontiguousarray(p).copy()
        fn = self.lib.ref_facade_steps if facade else self.lib.ref_naive_steps
        secs = fn(_ptr(q), q.shape[0], G, dt, eps2, steps)
        return (q, secs) if timed else q

    def mpi_naive_steps(self, p, G, dt, eps2, steps=1, nranks=1, timed=False):
        q = np.ascontiguousarray(p).copy()
        secs = self.lib.ref_mpi_naive_steps(_ptr(q), q.shape[0], G, dt, eps2, steps, nranks)
        return (q, secs) if timed else q

    def mpi_headless_loop(self, p, G, dt, eps2, steps=1, nranks=1):
        """(particles, seconds) with the reference's OWN definition of seconds for an MPI run: the headless
        step loop with its per-step gather to rank 0 (src/apps/headless_main.cpp:344-421, --disable-output)."""
        q = np.ascontiguousarray(p).copy()
        secs = self.lib.ref_mpi_headless_loop(_ptr(q), q.shape[0], G, dt, eps2, steps, nranks)
        return q, secs

    # checkpoint (trajectory) file
    def checkpoint_begin(self, path, p, target_steps):
        p = np.ascontiguousarray(p)
        self.lib.ref_checkpoint_begin(path.encode(), _ptr(p), p.shape[0], target_steps)

    def checkpoint_write_step(self, path, p):
        p = np.ascontiguousarray(p)
        self.lib.ref_checkpoint_write_step(path.encode(), _ptr(p), p.shape[0])

    def checkpoint_read_header(self, path):
        out = np.zeros(3, np.uint64)
        self.lib.ref_checkpoint_read_header(path.encode(), _ptr(out))
        return tuple(int(x) for x in out)

    def checkpoint_read_masses(self, path, n):
        out = np.zeros(n, np.float32)
        self.lib.ref_checkpoint_read_masses(path.encode(), _ptr(out), n)
        return out

    def checkpoint_read_step(self, path, n, step_index):
        out = np.zeros((n, 3), np.float32)
        got = self.lib.ref_checkpoint_read_step(path.encode(), _ptr(out), step_index)
        return out, int(got)
