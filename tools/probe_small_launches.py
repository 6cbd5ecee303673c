"""A few plain (non-graph) steps at small N, for an ncu launch list: which kernel the step time goes to."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb
sim = nb.CudaNaiveSimulation(mode="fast")
for n in (4096, 10000, 20000):
    p, eps = nb.datasets.plummer_sphere(n, seed=1)
    params = nb.SimParams(1.0, 0.5, eps)
    sim.upload(p); sim.step(params, 1); sim.synchronize()
    for _ in range(3):
        sim.step(params, 3)          # < 4 steps: plain launches
    sim.synchronize()
    print(n, sim.plan(), sim.last_step_ms() / 3 * 1e3, "us/step (plain launches)")
