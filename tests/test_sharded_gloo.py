"""CPU, world_size 2 and 3 over gloo: the host-side logic of the sharded path -- the reference's
slice rule, the padded slot layout, and the in-place all-gather that republishes positions --
driven with the oracle's row-order force as the compute stand-in (tests may use the oracle; the
product never does).  The sharded trajectory must equal the unsharded one bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, steps, out_dir):
    sys.path.insert(0, ROOT)
    import nbody_b200 as nb
    from oracle import Oracle

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    oracle = Oracle()
    p, eps = nb.datasets.plummer_sphere(n, seed=17)
    p["mass"] = np.random.default_rng(3).uniform(0.5, 2.0, n).astype(np.float32)
    G, dt = np.float32(1.0), np.float32(0.5)
    lay = nb.SlotLayout(n, world)
    counts, displs = nb.slice_counts(n, world)
    lo, cnt = displs[rank], counts[rank]
    local = p[lo:lo + cnt].copy()

    # replicated slot array {x, y, z, m}; everything starts as zero-mass padding
    full = torch.zeros(lay.n_slots * 4, dtype=torch.float32)
    full.view(-1, 4)[:, :3] = 1.0e18
    seg = full.view(-1, 4)[rank * lay.slot_stride: rank * lay.slot_stride + cnt]

    def publish():
        seg[:, :3] = torch.from_numpy(np.ascontiguousarray(local["position"]))
        seg[:, 3] = torch.from_numpy(np.ascontiguousarray(local["mass"]))
        nb.allgather_slices(full, rank, lay, dist)

    def global_view():
        g = np.zeros(n, nb.PARTICLE_DTYPE)
        slots = full.view(-1, 4).numpy()
        for r in range(world):
            s = slots[r * lay.slot_stride: r * lay.slot_stride + counts[r]]
            g["position"][displs[r]:displs[r] + counts[r]] = s[:, :3]
            g["mass"][displs[r]:displs[r] + counts[r]] = s[:, 3]
        return g

    publish()
    for _ in range(steps):
        acc = oracle.accel_rows(global_view(), float(G), eps, lo, lo + cnt)
        local["acceleration"] = acc
        local["velocity"] = local["velocity"] + acc * dt          # float32: product and sum rounded separately
        local["position"] = local["position"] + local["velocity"] * dt
        publish()
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), local)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 1001), (3, 700)])
def test_sharded_trajectory_equals_unsharded(tmp_path, oracle, nb, world, n):
    steps = 3
    port = 29600 + world
    mp.spawn(_worker, args=(world, port, n, steps, str(tmp_path)), nprocs=world, join=True)
    p, eps = nb.datasets.plummer_sphere(n, seed=17)
    p["mass"] = np.random.default_rng(3).uniform(0.5, 2.0, n).astype(np.float32)
    want = oracle.step_rows(p, 1.0, 0.5, eps, steps)
    got = np.concatenate([np.load(tmp_path / f"rank{r}.npy") for r in range(world)])
    assert got.tobytes() == want.tobytes()
