"""Host-side mirror of the reference's algorithm plug-in for the direct-sum path.

``CudaNaiveSimulation`` has the reference's operator interface --
``computeStep(particles, params)`` advancing the particles one step IN PLACE
(include/SimulationAlgorithm.h:18-22) -- with ``particles`` a NumPy structured array of the
reference's 40-byte record (include/Particle.hpp:53-59) and ``params`` a ``SimParams`` with the
reference's fields and defaults (include/SimulationAlgorithm.h:4-9).  Everything it does goes
through the C ABI of ``libnbody_cuda.so`` (include/nbody_cuda.h); torch is imported only when a
process group is given (one process per GPU; plumbing, not compute).

With a process group the object behaves as ``MpiNaiveSimulation`` does per rank
(src/MpiNaiveSimulation.cpp:10-141): ``particles`` is this rank's contiguous slice, cut by the
rule of src/apps/headless_main.cpp:249-259; the exchange of positions is either direct NVLink
stores from the integrator kernel (``exchange="p2p"``, the product) or an NCCL all-gather issued
through ``torch.distributed`` (``exchange="nccl"``, the baseline it is measured against).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib
from ._lib import NBodyCudaError  # noqa: F401  (re-exported)
from .sharding import SlotLayout, allgather_slices

# include/Particle.hpp:53-59
PARTICLE_DTYPE = np.dtype(
    [("position", "<f4", (3,)), ("velocity", "<f4", (3,)), ("acceleration", "<f4", (3,)), ("mass", "<f4")]
)


@dataclass
class SimParams:
    """include/SimulationAlgorithm.h:4-9 (same names, same defaults)."""

    G: float = 1.0
    dt: float = 0.05
    min_r2: float = 1e-8  # softening^2, added to r^2


def slice_counts(n: int, world: int):
    """counts[r] = n/P + (r < n%P); displs = prefix sum (src/apps/headless_main.cpp:249-259)."""
    lay = SlotLayout(int(n), int(world))
    return lay.counts, lay.displs


def _as_particles(a) -> np.ndarray:
    if not isinstance(a, np.ndarray) or a.dtype != PARTICLE_DTYPE or a.ndim != 1:
        raise TypeError("particles must be a 1-D numpy array of PARTICLE_DTYPE (40-byte particle_t records)")
    if not a.flags["C_CONTIGUOUS"]:
        raise ValueError("particles must be contiguous")
    return a


class _DeviceFloats:
    """Zero-copy view of library-owned device memory for torch (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = {
            "shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2, "strides": None,
        }


class CudaNaiveSimulation:
    """Drop-in for ``NaiveSimulation`` / ``MpiNaiveSimulation`` on one B200 per process.

    mode      "fast" (packed-FP32 kernel, the product), "strict" (the reference's exact operation
              order: bit-identical results) or "fp64" (forces summed in double; optional variant).
    group     a ``torch.distributed`` process group, or None for a single GPU.
    exchange  "p2p" | "nccl" (only with a group).
    """

    def __init__(self, device: int | None = None, mode: str = "fast", group=None, exchange: str = "p2p"):
        self._lib = _lib.load_library()
        if mode not in ("fast", "strict", "fp64"):
            raise ValueError("mode must be 'fast', 'strict' or 'fp64'")
        if exchange not in ("p2p", "nccl"):
            raise ValueError("exchange must be 'p2p' or 'nccl'")
        self.mode = mode
        self.exchange = exchange
        self.group = group
        self.rank, self.world = 0, 1
        self._torch = None
        if group is not None:
            import torch
            import torch.distributed as dist

            self._torch, self._dist = torch, dist
            self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if device is None:
            device = 0
            if group is not None:
                import os

                device = int(os.environ.get("LOCAL_RANK", self.rank))
        self.device = device
        h = C.c_void_p()
        code = self._lib.nbody_cuda_create(C.byref(h), device, {"fast": _lib.MODE_FAST, "strict": _lib.MODE_STRICT, "fp64": _lib.MODE_FP64}[mode])
        _lib.check(self._lib, None, code)
        self._h = h
        self.n_total = None
        self._pos_tensors = {}
        self._ext_stream = None

    # -- plumbing ---------------------------------------------------------------------------
    def _ck(self, code: int) -> None:
        _lib.check(self._lib, self._h, code)

    def _release_peers(self) -> None:
        """COLLECTIVE (every rank of the group): unmap all peer blocks, then meet at a barrier, so that
        no rank frees its exported block while another still has it open (CUDA IPC rule; the order
        include/nbody_cuda.h documents for nbody_cuda_peer_detach_all)."""
        if self.group is not None and self.world > 1 and self._h.value:
            self._ck(self._lib.nbody_cuda_peer_detach_all(self._h))
            self._dist.barrier(group=self.group)

    def close(self) -> None:
        """Frees the handle.  With a process group this is a COLLECTIVE call: every rank must close (the
        ranks unmap one another's position blocks and meet at a barrier before any block is freed)."""
        if getattr(self, "_h", None) is not None and self._h.value:
            self._release_peers()
            self._lib.nbody_cuda_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        # garbage collection cannot be collective: unmap what this rank has open and free, without the barrier
        try:
            if getattr(self, "_h", None) is not None and self._h.value:
                self._lib.nbody_cuda_peer_detach_all(self._h)
                self._lib.nbody_cuda_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _global_count(self, n_local: int) -> int:
        if self.world == 1:
            return n_local
        t = self._torch.tensor([n_local], dtype=self._torch.int64, device=f"cuda:{self.device}")
        self._dist.all_reduce(t, group=self.group)   # MpiNaiveSimulation.cpp:33-35 (counts exchange)
        return int(t.item())

    def configure(self, n_total: int) -> None:
        """Declare the global particle count (collective when a group is set)."""
        if self.n_total == n_total:
            return
        if self.n_total is not None:
            self._release_peers()       # reconfiguring frees the exported block: peers let go of it first
        self._ck(self._lib.nbody_cuda_configure(self._h, n_total, self.rank, self.world))
        self.n_total = n_total
        self._pos_tensors = {}
        if self.world > 1 and self.exchange == "p2p":
            self._attach_peers()

    def slice(self):
        b, c, sb, ss = C.c_size_t(), C.c_size_t(), C.c_size_t(), C.c_size_t()
        self._ck(self._lib.nbody_cuda_slice(self._h, C.byref(b), C.byref(c), C.byref(sb), C.byref(ss)))
        return dict(begin=b.value, count=c.value, slot_begin=sb.value, slot_stride=ss.value)

    def _attach_peers(self) -> None:
        """Swap the peer tokens and map every other rank's position block (CUDA IPC over NVLink).
        If ANY rank cannot (no peer access between the devices, IPC not permitted in this
        container), all ranks agree to use the NCCL all-gather exchange instead and say so."""
        tok = (C.c_ubyte * _lib.PEER_HANDLE_BYTES)()
        self._ck(self._lib.nbody_cuda_peer_export(self._h, tok))
        mine = bytes(tok)
        gathered = [None] * self.world
        self._dist.all_gather_object(gathered, mine, group=self.group)
        ok, why = 1, ""
        for r, t in enumerate(gathered):
            if r == self.rank:
                continue
            buf = (C.c_ubyte * _lib.PEER_HANDLE_BYTES).from_buffer_copy(t)
            try:
                self._ck(self._lib.nbody_cuda_peer_attach(self._h, r, buf))
            except NBodyCudaError as e:
                ok, why = 0, str(e)
                break
        flag = self._torch.tensor([ok], dtype=self._torch.int32, device=f"cuda:{self.device}")
        self._dist.all_reduce(flag, op=self._dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 0:
            import warnings

            warnings.warn("peer-to-peer exchange unavailable" + (f" ({why})" if why else " on another rank") +
                          "; using the NCCL all-gather exchange")
            self.exchange = "nccl"      # mappings that did succeed stay unused until the next configure

    def _allgather_positions(self) -> None:
        """NCCL baseline: in-place all-gather of the current position array on the handle's stream."""
        torch = self._torch
        if self._ext_stream is None:
            self._ext_stream = torch.cuda.ExternalStream(self._lib.nbody_cuda_stream(self._h), device=self.device)
        ptr = self._lib.nbody_cuda_positions(self._h)
        s = self.slice()
        if ptr not in self._pos_tensors:
            n_floats = 4 * s["slot_stride"] * self.world
            self._pos_tensors[ptr] = torch.as_tensor(_DeviceFloats(ptr, n_floats), device=f"cuda:{self.device}")
        full = self._pos_tensors[ptr]
        with torch.cuda.stream(self._ext_stream):
            allgather_slices(full, self.rank, SlotLayout(self.n_total, self.world), self._dist, self.group)

    # -- device-resident API ----------------------------------------------------------------
    def upload(self, particles: np.ndarray, n_total: int | None = None) -> None:
        """Host AoS slice -> device; with a group, also brings in every other rank's positions."""
        p = _as_particles(particles)
        if n_total is None:
            n_total = self._global_count(p.shape[0])
        self.configure(n_total)
        self._ck(self._lib.nbody_cuda_upload_aos(self._h, p.ctypes.data_as(C.c_void_p), p.shape[0]))
        if self.world > 1:
            if self.exchange == "p2p":
                self._ck(self._lib.nbody_cuda_exchange(self._h))
            else:
                self._allgather_positions()

    def generate(self, kind: str, n_total: int, size_a: float, size_b: float, G: float = 1.0, seed: int = 12345) -> float:
        """Synthetic input generated on the device ("plummer": scale radius, total mass (<0: one per
        particle); "disk": radius, thickness); returns the min_r2 the reference's generator would set."""
        self.configure(n_total)
        eps = C.c_float()
        self._ck(self._lib.nbody_cuda_generate(self._h, {"plummer": 0, "disk": 1}[kind], size_a, size_b, G, seed, C.byref(eps)))
        return eps.value

    def step(self, params: SimParams, n_steps: int = 1) -> None:
        """n_steps device-resident steps (asynchronous)."""
        if self.world > 1 and self.exchange == "nccl":
            for _ in range(n_steps):
                self._ck(self._lib.nbody_cuda_step_local(self._h, params.G, params.dt, params.min_r2))
                self._allgather_positions()
        else:
            self._ck(self._lib.nbody_cuda_step(self._h, params.G, params.dt, params.min_r2, n_steps))

    def download(self, particles: np.ndarray) -> None:
        p = _as_particles(particles)
        self._ck(self._lib.nbody_cuda_download_aos(self._h, p.ctypes.data_as(C.c_void_p), p.shape[0]))

    def download_positions(self, out: np.ndarray | None = None) -> np.ndarray:
        """(count, 3) float32 positions of the local slice: one record of the reference's trajectory
        file (src/CheckpointManager.cpp:86-99) without the 40-byte AoS detour.  Pass a reusable
        (ideally page-locked) `out` to avoid a fresh pageable allocation per record."""
        n = self.slice()["count"]
        if out is None:
            out = np.empty((n, 3), np.float32)
        elif out.shape != (n, 3) or out.dtype != np.float32 or not out.flags["C_CONTIGUOUS"]:
            raise ValueError(f"out must be a contiguous float32 array of shape ({n}, 3)")
        self._ck(self._lib.nbody_cuda_download_positions(self._h, out.ctypes.data_as(C.c_void_p), n))
        return out

    def download_positions_async(self, out: np.ndarray) -> None:
        """Start a trajectory record into `out` ((count, 3) float32, page-locked: PinnedArray) and return at
        once; steps queued afterwards overlap the copy.  `out` is valid after wait_records()."""
        n = self.slice()["count"]
        if out.shape != (n, 3) or out.dtype != np.float32 or not out.flags["C_CONTIGUOUS"]:
            raise ValueError(f"out must be a contiguous float32 array of shape ({n}, 3)")
        self._ck(self._lib.nbody_cuda_download_positions_async(self._h, out.ctypes.data_as(C.c_void_p), n))

    def wait_records(self) -> None:
        """Block until every record started with download_positions_async has landed in host memory."""
        self._ck(self._lib.nbody_cuda_download_positions_wait(self._h))

    def render_frame(self, width: int = 800, height: int = 600, camera_distance: float = 800.0,
                     angle_x: float = 0.0, angle_y: float = 0.0) -> np.ndarray:
        """(height, width, 3) uint8 frame of the device positions with the reference viewer's camera
        and sphere markers (src/renderers/GlutRenderer.cpp:98-148), rendered by CUDA kernels."""
        out = np.empty((height, width, 3), np.uint8)
        self._ck(self._lib.nbody_cuda_render_frame(self._h, width, height, camera_distance, angle_x, angle_y,
                                                   out.ctypes.data_as(C.c_void_p)))
        return out

    def synchronize(self) -> None:
        self._ck(self._lib.nbody_cuda_synchronize(self._h))

    # -- the reference interface ------------------------------------------------------------
    def computeStep(self, particles: np.ndarray, params: SimParams, n_total: int | None = None) -> None:
        """One step in place; the host array is the source of truth on entry and on return.

        With a group, `n_total` (the global particle count, the same on every rank) saves the count
        all-reduce that otherwise precedes every call -- the reference re-gathers the counts each step too
        (src/MpiNaiveSimulation.cpp:33-35); a caller whose slices never change size can state the total once.

        Single GPU: NaiveSimulation::computeStep (src/NaiveSimulation.cpp:7-10).
        With a group: MpiNaiveSimulation::computeStep on this rank's slice
        (src/MpiNaiveSimulation.cpp:10-13)."""
        p = _as_particles(particles)
        if self.world == 1:
            self._ck(self._lib.nbody_cuda_compute_step_aos(
                self._h, p.ctypes.data_as(C.c_void_p), p.shape[0], params.G, params.dt, params.min_r2))
            self.n_total = p.shape[0]
            return
        self.upload(p, n_total=n_total)
        self.step(params, 1)
        self.download(p)

    # -- diagnostics ------------------------------------------------------------------------
    def energy(self, params: SimParams):
        """(kinetic, potential) in FP64 of the device state; summed over ranks when sharded."""
        k, u = C.c_double(), C.c_double()
        self._ck(self._lib.nbody_cuda_energy(self._h, params.G, params.min_r2, C.byref(k), C.byref(u)))
        if self.world > 1:
            t = self._torch.tensor([k.value, u.value], dtype=self._torch.float64, device=f"cuda:{self.device}")
            self._dist.all_reduce(t, group=self.group)
            return float(t[0]), float(t[1])
        return k.value, u.value

    def last_step_ms(self) -> float:
        ms = C.c_float()
        self._ck(self._lib.nbody_cuda_last_step_ms(self._h, C.byref(ms)))
        return ms.value

    def profile_force(self, enable: bool = True, reserve: int = 0) -> None:
        """Bracket every force launch with CUDA events; `reserve` pre-creates that many event pairs."""
        self._ck(self._lib.nbody_cuda_profile_force(self._h, max(int(reserve), 1) if enable else 0))

    def force_time(self, reset: bool = True):
        """(total milliseconds, launches) of the force kernel since the last reset."""
        ms, n = C.c_double(), C.c_int()
        self._ck(self._lib.nbody_cuda_force_time(self._h, C.byref(ms), C.byref(n), int(reset)))
        return ms.value, n.value

    def plan(self):
        """Work plan of the fast force kernel: targets per thread, grid, target tiles, source tiles."""
        a, b, c, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self._ck(self._lib.nbody_cuda_plan(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return dict(targets_per_thread=a.value, grid=b.value, target_tiles=c.value, source_tiles=d.value)

    def set_option(self, name: str, value: int) -> None:
        """"targets_per_thread" (0 auto, 1/2/4/7), "barrier_timeout_ms" or "resident_kernel" (0 off, 1 small
        systems, 2 always) -- nbody_cuda_set_option."""
        opt = {"targets_per_thread": _lib.OPT_TARGETS_PER_THREAD, "barrier_timeout_ms": _lib.OPT_BARRIER_TIMEOUT_MS,
               "resident_kernel": _lib.OPT_RESIDENT_KERNEL}[name]
        self._ck(self._lib.nbody_cuda_set_option(self._h, opt, int(value)))

    def launch_count(self) -> int:
        return int(self._lib.nbody_cuda_launch_count(self._h))

    def measure_fma_peak(self, packed: bool) -> float:
        v = C.c_double()
        self._ck(self._lib.nbody_cuda_measure_fma_peak(self._h, int(packed), C.byref(v)))
        return v.value


class PinnedArray:
    """A page-locked NumPy array of the given shape and dtype (nbody_cuda_host_alloc), e.g. the
    (count, 3) float32 landing buffer of an asynchronous trajectory record."""

    def __init__(self, shape, dtype=np.float32):
        self._lib = _lib.load_library()
        self._ptr = C.c_void_p()
        dt = np.dtype(dtype)
        count = int(np.prod(shape))
        nbytes = max(1, count) * dt.itemsize
        _lib.check(self._lib, None, self._lib.nbody_cuda_host_alloc(C.byref(self._ptr), nbytes))
        self._raw = (C.c_ubyte * nbytes).from_address(self._ptr.value)
        self.array = np.frombuffer(self._raw, dtype=dt, count=count).reshape(shape)

    def close(self) -> None:
        if self._ptr is not None and self._ptr.value:
            self.array = None
            self._raw = None
            self._lib.nbody_cuda_host_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PinnedParticles:
    """A page-locked PARTICLE_DTYPE array of n records (nbody_cuda_host_alloc) for full-rate,
    asynchronous PCIe copies.  ``.array`` is valid until ``close()`` (or garbage collection)."""

    def __init__(self, n: int):
        self._lib = _lib.load_library()
        self._ptr = C.c_void_p()
        nbytes = max(1, n) * PARTICLE_DTYPE.itemsize
        _lib.check(self._lib, None, self._lib.nbody_cuda_host_alloc(C.byref(self._ptr), nbytes))
        self._raw = (C.c_ubyte * nbytes).from_address(self._ptr.value)
        self.array = np.frombuffer(self._raw, dtype=PARTICLE_DTYPE, count=n)

    def close(self) -> None:
        if self._ptr is not None and self._ptr.value:
            self.array = None
            self._raw = None
            self._lib.nbody_cuda_host_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
