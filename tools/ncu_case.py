"""One small, fixed sequence of force-kernel launches for an ncu capture.
usage: ncu_case.py <n> <plummer|disk> ; launches: step 1 enqueues both force loops (the mass flag is not yet known to the
host; one returns at once), then 2 steps with only the loop that applies -> capture with -k regex:force_fast_kernel -s 2 -c 1."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb
n, gen = int(sys.argv[1]), sys.argv[2]
p, eps = (nb.datasets.galaxy_disk if gen == "disk" else nb.datasets.plummer_sphere)(n, seed=12345)
sim = nb.CudaNaiveSimulation(mode="fast")
sim.set_option("resident_kernel", 0)
sim.upload(p)
params = nb.SimParams(1.0, 0.5, eps)
sim.step(params, 1); sim.synchronize()
sim.step(params, 2); sim.synchronize()
print(n, gen, sim.plan(), sim.last_step_ms() / 2, "ms/step")
sim.close()
