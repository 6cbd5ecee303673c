"""The reference's own benchmark points (data/seq-naive.csv, data/mpi-naive.csv: N = 10k..100k, 100 steps):
wall time of 100 device-resident steps here, beside the published seconds."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb
published = {10000: (17.12, 2.42), 20000: (68.59, None), 50000: (455.1, 59.28), 100000: (1889.1, 237.0)}   # seq, mpi np=8
sim = nb.CudaNaiveSimulation(mode="fast")
for n, (seq_s, mpi_s) in published.items():
    for gen in ("disk", "plummer"):
        p, eps = (nb.datasets.galaxy_disk if gen == "disk" else nb.datasets.plummer_sphere)(n, seed=1)
        params = nb.SimParams(1.0, 0.5, eps)
        sim.upload(p); sim.step(params, 10); sim.synchronize()
        t0 = time.perf_counter()
        sim.upload(p); sim.step(params, 100); sim.download(p)
        wall = time.perf_counter() - t0
        ms = sim.last_step_ms()
        print(f"n={n} {gen}: 100 steps {wall*1e3:.1f} ms wall ({ms:.1f} ms device), {n*n*100/(ms*1e-3):.3e} int/s, plan {sim.plan()}; "
              f"published seq {seq_s} s, mpi np=8 {mpi_s} s", flush=True)
