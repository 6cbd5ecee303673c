"""BASELINE.json configs[4]: N = 4,194,304 Plummer on 8 B200 -- device-resident run with trajectory
records written through the reference's UNMODIFIED CheckpointManager (oracle/_ref), read back and
compared bit for bit, and the FP64 energy tracked at every record.

    torchrun --nproc-per-node 8 tools/run_config5.py [--particles 4194304 --steps 1000 --every 100 --mode fast|fp64]

This is a measurement/validation script (it may use oracle/_ref as the checker and as the
reference's checkpoint writer); the product path is nbody_b200.CudaNaiveSimulation.
Weak-scaling series (N^2/P constant): N(P) = 4,194,304 * sqrt(P/8) -> use --particles with P ranks.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402


def all_positions(sim, lib, n, world):
    """(n,3) float32 positions in global order from this rank's replicated device array."""
    from nbody_b200.simulation import _DeviceFloats

    lay = nb.SlotLayout(n, world)
    sim.synchronize()
    t = torch.as_tensor(_DeviceFloats(lib.nbody_cuda_positions(sim._h), lay.n_slots * 4), device=f"cuda:{sim.device}")
    slots = t.view(-1, 4).cpu().numpy()
    out = np.empty((n, 3), np.float32)
    for r, (c, d) in enumerate(zip(lay.counts, lay.displs)):
        out[d:d + c] = slots[r * lay.slot_stride: r * lay.slot_stride + c, :3]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", type=int, default=4194304, help="not --n: torchrun would read that as one of its own options")
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--every", type=int, default=100)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "config5.json"))
    ap.add_argument("--traj", default="/tmp/config5_simulation_output.bin")
    ap.add_argument("--mode", default="fast", choices=["fast", "strict", "fp64"],
                    help="fp64: forces summed in double -- the yardstick the fast run's energy drift is compared with")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
        group = dist.group.WORLD
    n = args.particles
    p, eps = nb.datasets.plummer_sphere(n, 100.0, float(n), seed=12345)
    params = nb.SimParams(1.0, 0.5, eps)
    counts, displs = nb.slice_counts(n, world)
    lo, cnt = displs[rank], counts[rank]
    sim = nb.CudaNaiveSimulation(device=local, mode=args.mode, group=group, exchange="p2p")
    sim.set_option("barrier_timeout_ms", 120000)
    lib = nb.load_library()
    sim.upload(p[lo:lo + cnt].copy(), n_total=n)

    ref = None
    if rank == 0:
        from oracle import Reference

        ref = Reference() if Reference.available() else None
        if ref is not None:
            ref.checkpoint_begin(args.traj, p, args.steps)      # header + masses (CheckpointManager.cpp:10-48)
    records, energies, step_ms = [], [], []
    k0, u0 = sim.energy(params)
    energies.append((0, k0, u0))
    if rank == 0:
        print(f"step 0: E = {k0 + u0:.10e}", flush=True)
    done = 0
    t_start = time.time()
    while done < args.steps:
        chunk = min(args.every, args.steps - done)
        sim.step(params, chunk)
        sim.synchronize()
        step_ms.append(sim.last_step_ms() / chunk)
        done += chunk
        k, u = sim.energy(params)
        energies.append((done, k, u))
        if rank == 0:
            pos = all_positions(sim, lib, n, world)
            if ref is not None:
                aos = np.zeros(n, nb.PARTICLE_DTYPE)
                aos["position"] = pos
                aos["mass"] = p["mass"]
                ref.checkpoint_write_step(args.traj, aos)        # CheckpointManager.cpp:76-107
                back, got = ref.checkpoint_read_step(args.traj, n, len(records))   # :144-177
                ok = got == n and back.tobytes() == pos.tobytes()
            else:
                ok = None
            records.append((done, ok))
            print(f"step {done}: E = {k + u:.10e}  dE/|E0| = {(k + u - k0 - u0) / abs(k0 + u0):+.3e}  "
                  f"{step_ms[-1]:.2f} ms/step  checkpoint round trip {'ok' if ok else ok}", flush=True)
    wall = time.time() - t_start
    if rank == 0:
        rate = float(n) * n / (np.mean(step_ms) * 1e-3)
        res = dict(mode=args.mode, n=n, gpus=world, steps=args.steps, every=args.every, min_r2=eps, ms_per_step=float(np.mean(step_ms)),
                   interactions_per_s=rate, wall_s=wall, energies=energies, records=records,
                   drift=(energies[-1][1] + energies[-1][2] - k0 - u0) / abs(k0 + u0),
                   trajectory_bytes=os.path.getsize(args.traj) if ref is not None else None,
                   checkpoint_writer="reference CheckpointManager (oracle/_ref)" if ref is not None else "unavailable")
        os.makedirs(os.path.dirname(args.out), exist_ok=True)
        json.dump(res, open(args.out, "w"), indent=1)
        print(json.dumps({k: v for k, v in res.items() if k not in ("energies", "records")}))
        if ref is not None:
            hdr = ref.checkpoint_read_header(args.traj)
            assert hdr == (n, args.steps, len(records)), hdr
            assert all(ok for _, ok in records)
            os.unlink(args.traj)
    sim.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
