"""Small end-to-end case for compute-sanitizer (one tool per gpurun call): every kernel of the
library on awkward sizes, both modes, trajectory download, energy, the allocation-free entry point."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb

for mode in ("fast", "strict"):
    sim = nb.CudaNaiveSimulation(mode=mode)
    for n in (1, 257, 1793, 5003):
        p, eps = nb.datasets.plummer_sphere(n, seed=n)
        if n == 5003:
            p["mass"][::2] = 2.0
        params = nb.SimParams(1.0, 0.5, eps)
        sim.computeStep(p, params)
        sim.upload(p)
        sim.step(params, 5)
        sim.download(p)
        sim.download_positions()
        sim.energy(params)
    sim.close()
# slice handles (multi-rank layout) on one device
lib = nb.load_library()
for r in range(3):
    s = nb.CudaNaiveSimulation(mode="fast")
    s.rank, s.world = r, 3
    s._ck(lib.nbody_cuda_configure(s._h, 4001, r, 3))
    counts, displs = nb.slice_counts(4001, 3)
    p, eps = nb.datasets.plummer_sphere(4001, seed=1)
    sl = p[displs[r]:displs[r] + counts[r]].copy()
    s._ck(lib.nbody_cuda_upload_aos(s._h, sl.ctypes.data_as(C.c_void_p), counts[r]))
    s._ck(lib.nbody_cuda_step_local(s._h, 1.0, 0.5, eps))
    s.download(sl)
    s.close()
print("sanitize case finished")
