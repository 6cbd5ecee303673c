"""ctypes binding of ``libnbody_cuda.so`` -- every symbol ``include/nbody_cuda.h`` declares.

The library is the product; this module only loads it.  There is no fallback: if the shared
object is missing, or no CUDA device is usable, the error is raised to the caller.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libnbody_cuda.so")
HEADER = os.path.join(os.path.dirname(HERE), "include", "nbody_cuda.h")

NBODY_OK = 0
MODE_FAST = 0
MODE_STRICT = 1
MODE_FP64 = 2
PEER_HANDLE_BYTES = 128
OPT_TARGETS_PER_THREAD = 2
OPT_BARRIER_TIMEOUT_MS = 3
OPT_RESIDENT_KERNEL = 4


class NBodyCudaError(RuntimeError):
    """A C-ABI call returned a negative code; ``code`` holds it, the text is the library's message."""

    def __init__(self, code: int, message: str):
        super().__init__(f"nbody_cuda error {code}: {message}")
        self.code = code


def build_library(force: bool = False) -> str:
    """Compile csrc/ for sm_100a with nvcc (cross-compiles without a GPU)."""
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    r = subprocess.run(["make", "-C", CSRC], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libnbody_cuda.so failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


# name -> (restype, argtypes); one entry per declaration in include/nbody_cuda.h
_vp, _sz, _i, _f, _d = C.c_void_p, C.c_size_t, C.c_int, C.c_float, C.c_double
SIGNATURES = {
    "nbody_cuda_particle_bytes": (_sz, []),
    "nbody_cuda_version": (C.c_char_p, []),
    "nbody_cuda_device_count": (_i, []),
    "nbody_cuda_create": (_i, [C.POINTER(_vp), _i, _i]),
    "nbody_cuda_destroy": (None, [_vp]),
    "nbody_cuda_last_error": (C.c_char_p, [_vp]),
    "nbody_cuda_configure": (_i, [_vp, _sz, _i, _i]),
    "nbody_cuda_slice": (_i, [_vp, C.POINTER(_sz), C.POINTER(_sz), C.POINTER(_sz), C.POINTER(_sz)]),
    "nbody_cuda_upload_aos": (_i, [_vp, _vp, _sz]),
    "nbody_cuda_download_aos": (_i, [_vp, _vp, _sz]),
    "nbody_cuda_download_positions": (_i, [_vp, _vp, _sz]),
    "nbody_cuda_download_positions_async": (_i, [_vp, _vp, _sz]),
    "nbody_cuda_download_positions_wait": (_i, [_vp]),
    "nbody_cuda_generate": (_i, [_vp, _i, _f, _f, _f, C.c_ulonglong, C.POINTER(_f)]),
    "nbody_cuda_render_frame": (_i, [_vp, _i, _i, _f, _f, _f, _vp]),
    "nbody_cuda_step": (_i, [_vp, _f, _f, _f, _i]),
    "nbody_cuda_step_local": (_i, [_vp, _f, _f, _f]),
    "nbody_cuda_compute_step_aos": (_i, [_vp, _vp, _sz, _f, _f, _f]),
    "nbody_cuda_synchronize": (_i, [_vp]),
    "nbody_cuda_last_step_ms": (_i, [_vp, C.POINTER(_f)]),
    "nbody_cuda_profile_force": (_i, [_vp, _i]),
    "nbody_cuda_force_time": (_i, [_vp, C.POINTER(_d), C.POINTER(_i), _i]),
    "nbody_cuda_plan": (_i, [_vp, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)]),
    "nbody_cuda_plan_describe": (_i, [_sz, _sz, _i, _i, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), _sz,
                                      C.POINTER(_i), C.POINTER(_i), _sz]),
    "nbody_cuda_set_option": (_i, [_vp, _i, C.c_longlong]),
    "nbody_cuda_launch_count": (C.c_longlong, [_vp]),
    "nbody_cuda_energy": (_i, [_vp, _f, _f, C.POINTER(_d), C.POINTER(_d)]),
    "nbody_cuda_stream": (_vp, [_vp]),
    "nbody_cuda_positions": (_vp, [_vp]),
    "nbody_cuda_peer_export": (_i, [_vp, _vp]),
    "nbody_cuda_peer_attach": (_i, [_vp, _i, _vp]),
    "nbody_cuda_peer_attach_local": (_i, [_vp, _i, _vp]),
    "nbody_cuda_peer_detach_all": (_i, [_vp]),
    "nbody_cuda_exchange": (_i, [_vp]),
    "nbody_cuda_measure_fma_peak": (_i, [_vp, _i, C.POINTER(_d)]),
    "nbody_cuda_accel_noalloc": (_i, [_vp, _vp, _sz, _f, _f, _vp, C.c_char_p, _sz]),
    "nbody_cuda_host_alloc": (_i, [C.POINTER(_vp), _sz]),
    "nbody_cuda_host_free": (_i, [_vp]),
}

_lib = None


def load_library() -> C.CDLL:
    """Load the in-tree ``libnbody_cuda.so``.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C n-body-simulation_b200/csrc`).  There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(lib: C.CDLL, handle, code: int) -> None:
    if code != NBODY_OK:
        msg = lib.nbody_cuda_last_error(handle)
        raise NBodyCudaError(code, msg.decode() if msg else "")
