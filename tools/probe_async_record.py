"""Trajectory records overlapped with stepping: device time per step with no record, with a
synchronous record every step (the reference's default, src/apps/headless_main.cpp:402-411) and with asynchronous
records; and the record bandwidth at config-5 size."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nbody_b200 as nb

def run(n, steps, mode):
    p, eps = nb.datasets.plummer_sphere(n, seed=1)
    params = nb.SimParams(1.0, 0.5, eps)
    sim = nb.CudaNaiveSimulation(mode="fast")
    sim.upload(p); sim.step(params, 2); sim.synchronize()
    bufs = [nb.PinnedArray((n, 3)) for _ in range(2)]
    sim.download_positions_async(bufs[0].array); sim.wait_records()       # creates the copy stream and staging
    t0 = time.perf_counter()
    for k in range(steps):
        sim.step(params, 1)
        if mode == "sync":
            sim.download_positions(out=bufs[k & 1].array)
        elif mode == "async":
            sim.download_positions_async(bufs[k & 1].array)
    if mode == "async":
        sim.wait_records()
    sim.synchronize()
    wall = time.perf_counter() - t0
    for b in bufs: b.close()
    sim.close()
    return wall / steps

for n, steps in ((65536, 200), (262144, 40)):
    r = {m: run(n, steps, m) for m in ("none", "sync", "async")}
    print(json.dumps({"n": n, "steps": steps, "ms_per_step": {k: v * 1e3 for k, v in r.items()},
                      "record_overhead_pct": {k: 100 * (r[k] / r["none"] - 1) for k in ("sync", "async")}}), flush=True)

# record bandwidth at config-5 size (no stepping): 12 B x 4,194,304 per record
n = 4194304
sim = nb.CudaNaiveSimulation(mode="fast")
sim.generate("plummer", n, 100.0, -1.0)
sim.synchronize()
bufs = [nb.PinnedArray((n, 3)) for _ in range(2)]
for mode in ("sync", "async"):
    for rep in range(2):
        t0 = time.perf_counter()
        for k in range(8):
            if mode == "sync": sim.download_positions(out=bufs[k & 1].array)
            else: sim.download_positions_async(bufs[k & 1].array)
        sim.wait_records(); sim.synchronize()
        dt = time.perf_counter() - t0
    print(json.dumps({"n": n, "mode": mode, "records": 8, "GB_per_s": 8 * n * 12 / dt / 1e9}), flush=True)
