"""Quick on-GPU probe: FMA-pipe micro-benchmark and force-kernel throughput at a few sizes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb

sim = nb.CudaNaiveSimulation(mode="fast")
for packed in (False, True):
    print("fma peak packed=%d: %.2f TFLOP/s" % (packed, sim.measure_fma_peak(packed) / 1e12), flush=True)
sizes = [int(a) for a in sys.argv[1:]] or [4096, 16384, 65536, 262144, 1048576]
for n in sizes:
    p, eps = nb.datasets.plummer_sphere(n, seed=1)
    sim.upload(p)
    params = nb.SimParams(1.0, 0.5, eps)
    steps = max(2, min(50, int(3e12 / n / n)))
    sim.step(params, 3); sim.synchronize()
    sim.profile_force(True)
    sim.step(params, steps)
    ms = sim.last_step_ms()
    fms, fl = sim.force_time()
    sim.profile_force(False)
    rate = n * n * steps / (ms * 1e-3)
    frate = n * n * fl / (fms * 1e-3)
    print(sim.plan(), end=" ")
    print(f"fast n={n} steps={steps} {ms/steps:.3f} ms/step {rate:.3e} int/s  force-only {frate:.3e} int/s "
          f"= {frate*20/1e12:.1f} TFLOP/s(20-flop)", flush=True)
st = nb.CudaNaiveSimulation(mode="strict")
for n in [65536]:
    p, eps = nb.datasets.plummer_sphere(n, seed=1)
    st.upload(p); params = nb.SimParams(1.0, 0.5, eps)
    st.step(params, 1); st.synchronize()
    st.step(params, 3); ms = st.last_step_ms()
    print(f"strict n={n} {ms/3:.3f} ms/step {n*n*3/(ms*1e-3):.3e} int/s", flush=True)
