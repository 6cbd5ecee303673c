"""Small-N step rate on the device-resident path (CUDA-graph replay of step pairs)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb
sim = nb.CudaNaiveSimulation(mode="fast")
for n in (1024, 4096, 16384, 65536):
    p, eps = nb.datasets.plummer_sphere(n, seed=1)
    sim.upload(p); params = nb.SimParams(1.0, 0.5, eps)
    sim.step(params, 10); sim.synchronize()
    for steps in (3, 400):        # 3: plain launches; 400: graph replay
        sim.step(params, steps); ms = sim.last_step_ms()
        print(f"n={n} steps={steps}: {ms/steps*1e3:.1f} us/step  {n*n*steps/(ms*1e-3):.3e} int/s", flush=True)
