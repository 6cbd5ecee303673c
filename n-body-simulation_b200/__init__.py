"""B200-native direct-sum N-body hot path (drop-in for the reference's NaiveSimulation /
MpiNaiveSimulation ``computeStep``).  The product is ``csrc/libnbody_cuda.so`` behind the C ABI
in ``include/nbody_cuda.h``; this package is the thin host-side mirror of the reference's
plug-in interface on top of it.  Importable as ``nbody_b200`` (see ``nbody_b200.py`` at the
repository root: the directory name has a hyphen)."""
from ._lib import LIB_PATH, NBodyCudaError, SIGNATURES, build_library, load_library
from .simulation import PARTICLE_DTYPE, CudaNaiveSimulation, PinnedArray, PinnedParticles, SimParams, slice_counts
from .sharding import SlotLayout, allgather_slices
from . import datasets, trajectory
from .trajectory import TrajectoryReader, TrajectoryWriter, read_restart, write_ppm, write_restart

__all__ = [
    "CudaNaiveSimulation", "SimParams", "PARTICLE_DTYPE", "PinnedParticles", "PinnedArray", "slice_counts", "datasets", "trajectory", "TrajectoryWriter", "TrajectoryReader", "write_restart", "read_restart", "write_ppm", "SlotLayout", "allgather_slices",
    "NBodyCudaError", "build_library", "load_library", "LIB_PATH", "SIGNATURES",
]
