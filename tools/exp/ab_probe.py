"""A/B probe (experiment, not product): times resident steps and the force kernel alone through the raw C ABI
of whichever libnbody_cuda build is given, so two builds can be compared in one GPU session.
usage: ab_probe.py <lib.so> <sizes,comma> [general_loop]"""
import ctypes as C, sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import nbody_b200 as nb

lib = C.CDLL(sys.argv[1])
sizes = [int(x) for x in sys.argv[2].split(",")]
loop = int(sys.argv[3]) if len(sys.argv) > 3 else None
vp, f = C.c_void_p, C.c_float
lib.nbody_cuda_create.argtypes = [C.POINTER(vp), C.c_int, C.c_int]
lib.nbody_cuda_configure.argtypes = [vp, C.c_size_t, C.c_int, C.c_int]
lib.nbody_cuda_upload_aos.argtypes = [vp, vp, C.c_size_t]
lib.nbody_cuda_step.argtypes = [vp, f, f, f, C.c_int]
lib.nbody_cuda_synchronize.argtypes = [vp]
lib.nbody_cuda_last_step_ms.argtypes = [vp, C.POINTER(f)]
lib.nbody_cuda_profile_force.argtypes = [vp, C.c_int]
lib.nbody_cuda_force_time.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_int]
lib.nbody_cuda_last_error.argtypes = [vp]; lib.nbody_cuda_last_error.restype = C.c_char_p
h = vp()
assert lib.nbody_cuda_create(C.byref(h), 0, 0) == 0
if loop is not None and hasattr(lib, "nbody_cuda_set_option"):
    lib.nbody_cuda_set_option.argtypes = [vp, C.c_int, C.c_longlong]
    assert lib.nbody_cuda_set_option(h, 1, loop) == 0
PEAK = 148 * 128 * 2 * 1.965e9 / 20
def ck(rc):
    assert rc == 0, lib.nbody_cuda_last_error(h)
for n in sizes:
    for gen in ("plummer", "disk"):
        p, eps = (nb.datasets.galaxy_disk if gen == "disk" else nb.datasets.plummer_sphere)(n, seed=1)
        ck(lib.nbody_cuda_configure(h, n, 0, 1)); ck(lib.nbody_cuda_upload_aos(h, p.ctypes.data_as(vp), n))
        steps = max(4, min(2000, int(2e12 / (float(n) * n)))); steps += steps % 2
        ck(lib.nbody_cuda_step(h, 1.0, 0.5, eps, steps)); ck(lib.nbody_cuda_synchronize(h))
        best = 1e30
        for rep in range(3):
            ck(lib.nbody_cuda_step(h, 1.0, 0.5, eps, steps)); ms = f(); ck(lib.nbody_cuda_last_step_ms(h, C.byref(ms))); best = min(best, ms.value)
        # force kernel alone (plain launches, events around each)
        ck(lib.nbody_cuda_profile_force(h, 64)); k = min(steps, 20)
        ck(lib.nbody_cuda_force_time(h, None, None, 1))
        ck(lib.nbody_cuda_step(h, 1.0, 0.5, eps, k)); tot, cnt = C.c_double(), C.c_int()
        ck(lib.nbody_cuda_force_time(h, C.byref(tot), C.byref(cnt), 1)); ck(lib.nbody_cuda_profile_force(h, 0))
        fr = float(n) * n * cnt.value / (tot.value * 1e-3)
        r = float(n) * n * steps / (best * 1e-3)
        print(json.dumps({"lib": os.path.basename(sys.argv[1]), "n": n, "gen": gen, "loop": loop, "us_per_step": best / steps * 1e3,
                          "step_pct": 100 * r / PEAK, "force_us": tot.value / cnt.value * 1e3, "force_pct": 100 * fr / PEAK}), flush=True)
