"""Worker for tests/test_gpu_multi.py: one rank per GPU (torchrun).  Checks the sharded path --
i-slices per the reference's rule, positions republished each step -- against a single-GPU run
of the same system on the rank's own device.  Exit code 0 = all checks passed."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402


def rel_err(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    d = np.linalg.norm(b, axis=1)
    return np.linalg.norm(a - b, axis=1) / np.where(d == 0, 1.0, d)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    ok = True
    for n in (10007, 65536):                       # n % world != 0 and a config size
        p, eps = nb.datasets.plummer_sphere(n, seed=3)
        p["mass"] = np.random.default_rng(5).uniform(0.5, 2.0, n).astype(np.float32)
        params = nb.SimParams(1.0, 0.5, eps)
        counts, displs = nb.slice_counts(n, world)
        lo, cnt = displs[rank], counts[rank]
        steps = 4
        for mode in ("strict", "fast"):
            single = nb.CudaNaiveSimulation(device=local, mode=mode)
            want = p.copy()
            single.upload(want); single.step(params, steps); single.download(want)
            single.close()
            for exchange in ("p2p", "nccl"):
                sim = nb.CudaNaiveSimulation(device=local, mode=mode, group=dist.group.WORLD, exchange=exchange)
                # device-resident
                got = p[lo:lo + cnt].copy()
                sim.upload(got, n_total=n); sim.step(params, steps); sim.download(got)
                # reference-call semantics (MpiNaiveSimulation::computeStep on the local slice)
                got2 = p[lo:lo + cnt].copy()
                for _ in range(steps):
                    sim.computeStep(got2, params)
                k, u = sim.energy(params)
                sim.close()
                ref = want[lo:lo + cnt]
                if mode == "strict":
                    good = got.tobytes() == ref.tobytes() and got2.tobytes() == ref.tobytes()
                    detail = "bitwise"
                else:
                    e = max(rel_err(got["position"], ref["position"]).max(), rel_err(got2["position"], ref["position"]).max())
                    ea = rel_err(got["acceleration"], ref["acceleration"]).max()
                    good = e < 1e-5 and ea < 1e-4 and got.tobytes() == got2.tobytes()
                    detail = f"pos {e:.2e} acc {ea:.2e}"
                print(f"rank {rank} n={n} mode={mode} exchange={exchange}: {'ok' if good else 'FAIL'} ({detail}) E={k + u:.6e}", flush=True)
                ok = ok and good
    t = torch.tensor([0 if ok else 1], device=f"cuda:{local}")
    dist.all_reduce(t)
    dist.destroy_process_group()
    sys.exit(int(t.item() != 0))


if __name__ == "__main__":
    main()
