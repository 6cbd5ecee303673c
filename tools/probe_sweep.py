"""Force-kernel sweep on one GPU: sizes x input kind x forced targets per thread.
Device time of k resident steps (CUDA-graph replay), int/s and % of the FP32 FMA peak."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb

PEAK = 148 * 128 * 2 * 1.965e9 / 20     # interactions/s at 20 flop
sizes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [4096, 10000, 20000, 50000, 65536, 100000, 262144, 1048576]
widths = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0]
sim = nb.CudaNaiveSimulation(mode="fast")
for n in sizes:
    for gen in ("plummer", "disk"):
        p, eps = (nb.datasets.galaxy_disk if gen == "disk" else nb.datasets.plummer_sphere)(n, seed=1)
        params = nb.SimParams(1.0, 0.5, eps)
        for loop in (0,):
            for I in widths:
                sim.set_option("targets_per_thread", 0)
                sim.upload(p)
                sim.set_option("targets_per_thread", I)
                steps = max(4, min(2000, int(2e12 / (float(n) * n))))
                steps += steps % 2
                sim.step(params, steps); sim.synchronize()       # warm-up (clock, graph)
                best = 1e30
                for rep in range(3):
                    sim.step(params, steps); best = min(best, sim.last_step_ms())
                rate = float(n) * n * steps / (best * 1e-3)
                print(json.dumps({"n": n, "gen": gen, "loop": "unit" if gen == "plummer" else "general", "forced_I": I,
                                  "plan": sim.plan(), "steps": steps, "us_per_step": best / steps * 1e3, "int_per_s": rate,
                                  "pct_of_peak": 100 * rate / PEAK}), flush=True)
