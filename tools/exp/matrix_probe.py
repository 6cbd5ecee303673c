"""Experiment: times the resident-step rate of many builds of the library (tools/exp/r1lib/m/*.so) in one process.
usage: matrix_probe.py <n> <steps> <lib.so>..."""
import ctypes as C, sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import nbody_b200 as nb
n, steps = int(sys.argv[1]), int(sys.argv[2])
vp, f = C.c_void_p, C.c_float
data = {g: (nb.datasets.galaxy_disk if g == "disk" else nb.datasets.plummer_sphere)(n, seed=1) for g in ("plummer", "disk")}
PEAK = 148 * 128 * 2 * 1.965e9 / 20
for path in sys.argv[3:]:
    lib = C.CDLL(path)
    lib.nbody_cuda_create.argtypes = [C.POINTER(vp), C.c_int, C.c_int]
    lib.nbody_cuda_configure.argtypes = [vp, C.c_size_t, C.c_int, C.c_int]
    lib.nbody_cuda_upload_aos.argtypes = [vp, vp, C.c_size_t]
    lib.nbody_cuda_step.argtypes = [vp, f, f, f, C.c_int]
    lib.nbody_cuda_synchronize.argtypes = [vp]
    lib.nbody_cuda_last_step_ms.argtypes = [vp, C.POINTER(f)]
    lib.nbody_cuda_destroy.argtypes = [vp]
    h = vp(); assert lib.nbody_cuda_create(C.byref(h), 0, 0) == 0
    out = {"lib": os.path.basename(path), "n": n}
    for g, (p, eps) in data.items():
        assert lib.nbody_cuda_configure(h, n, 0, 1) == 0 and lib.nbody_cuda_upload_aos(h, p.ctypes.data_as(vp), n) == 0
        assert lib.nbody_cuda_step(h, 1.0, 0.5, eps, 4) == 0 and lib.nbody_cuda_synchronize(h) == 0
        best = 1e30
        for rep in range(3):
            assert lib.nbody_cuda_step(h, 1.0, 0.5, eps, steps) == 0
            ms = f(); lib.nbody_cuda_last_step_ms(h, C.byref(ms)); best = min(best, ms.value)
        out[g] = round(100 * float(n) * n * steps / (best * 1e-3) / PEAK, 2)
    lib.nbody_cuda_destroy(h)
    print(json.dumps(out), flush=True)
