"""BASELINE.json configs[1] in full: N = 65,536, 100 steps, strict mode against the CPU restatement of
the reference's sequential algorithm (oracle rows, OpenMP) -- bit for bit, every field -- for both
inputs the reference's apps use (galaxy disk: `--random` in the sequential build; Plummer: the MPI
build).  One-off verification (the CPU side needs ~4.3e11 pair terms per input); the -m gpu tests
cover the same at 3 steps.  Writes profiles/<out>.json."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb
from oracle import Oracle, Reference

n, steps = 65536, int(sys.argv[1]) if len(sys.argv) > 1 else 100
oracle = Oracle()
ref = Reference() if Reference.available() else None
out = {"n": n, "steps": steps, "cases": []}
for kind in ("disk", "plummer"):
    if ref is not None:
        p, eps = (ref.disk if kind == "disk" else ref.plummer)(n, 12345)
        src = "reference generator, seed 12345"
    else:
        p, eps = (nb.datasets.galaxy_disk if kind == "disk" else nb.datasets.plummer_sphere)(n, seed=12345)
        src = "numpy generator, seed 12345"
    params = nb.SimParams(1.0, 0.5, eps)
    sim = nb.CudaNaiveSimulation(mode="strict")
    q = p.copy()
    t0 = time.time()
    sim.upload(q); sim.step(params, steps); sim.download(q)
    t_gpu = time.time() - t0
    fast = nb.CudaNaiveSimulation(mode="fast")
    f = p.copy()
    fast.upload(f); fast.step(params, steps); fast.download(f)
    t0 = time.time()
    want = oracle.step_rows(p, 1.0, 0.5, eps, steps)
    t_cpu = time.time() - t0
    same = {k: bool(np.ascontiguousarray(q[k]).tobytes() == np.ascontiguousarray(want[k]).tobytes())
            for k in ("position", "velocity", "acceleration", "mass")}
    d = np.linalg.norm(f["position"].astype(np.float64) - want["position"], axis=1) / np.linalg.norm(want["position"].astype(np.float64), axis=1)
    case = dict(input=kind, source=src, min_r2=eps, strict_bit_identical=same, gpu_strict_seconds=t_gpu, cpu_oracle_seconds=t_cpu,
                fast_vs_reference_position_rel_err=dict(max=float(d.max()), median=float(np.median(d))))
    print(case, flush=True)
    out["cases"].append(case)
    sim.close(); fast.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "verify_config2.json"), "w"), indent=1)
assert all(all(c["strict_bit_identical"].values()) for c in out["cases"])
print("config 2 verified: strict mode bit-identical over", steps, "steps")
