"""Times the O(N) pieces around the hot path at BASELINE configs[4] size (N = 4,194,304) on one GPU:
boundary transposes + PCIe, trajectory record download, device generators, frame renderer, energy (N=262144)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nbody_b200 as nb

def timed(fn, reps=5):
    fn(); best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); best = min(best, time.perf_counter() - t0)
    return best

n = 4194304
sim = nb.CudaNaiveSimulation(mode="fast")
pin = nb.PinnedParticles(n)
p = pin.array
p["position"] = np.random.default_rng(1).standard_normal((n, 3)).astype(np.float32) * 100
p["mass"] = 1.0
pageable = p.copy()
def up_pinned(): sim.upload(p); sim.synchronize()
def up_pageable(): sim.upload(pageable); sim.synchronize()
t = timed(up_pinned);   print(f"upload   pinned   {n*40/1e6:.0f} MB: {t*1e3:.2f} ms  {n*40/t/1e9:.1f} GB/s")
t = timed(up_pageable); print(f"upload   pageable {n*40/1e6:.0f} MB: {t*1e3:.2f} ms  {n*40/t/1e9:.1f} GB/s")
t = timed(lambda: sim.download(p));        print(f"download pinned   {n*40/1e6:.0f} MB: {t*1e3:.2f} ms  {n*40/t/1e9:.1f} GB/s")
t = timed(lambda: sim.download(pageable)); print(f"download pageable {n*40/1e6:.0f} MB: {t*1e3:.2f} ms  {n*40/t/1e9:.1f} GB/s")
t = timed(lambda: sim.download_positions()); print(f"positions record {n*12/1e6:.0f} MB: {t*1e3:.2f} ms  {n*12/t/1e9:.1f} GB/s")
def gen(): sim.generate("plummer", n, 100.0, -1.0, 1.0, 3); sim.synchronize()
t = timed(gen); print(f"generate plummer on device, {n} particles: {t*1e3:.2f} ms")
t0 = time.perf_counter(); nb.datasets.plummer_sphere(n, seed=3); print(f"generate plummer numpy (host): {(time.perf_counter()-t0)*1e3:.0f} ms")
t = timed(lambda: sim.render_frame(800, 600, 800.0, 20.0, 30.0)); print(f"render 800x600 frame of {n} particles: {t*1e3:.2f} ms")
sim.generate("plummer", 262144, 100.0, -1.0, 1.0, 3)
t = timed(lambda: sim.energy(nb.SimParams(1.0, 0.5, 1.0)), reps=2); print(f"energy FP64, N=262144: {t*1e3:.1f} ms  ({262144**2/2/t:.3e} pairs/s)")
