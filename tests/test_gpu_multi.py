"""GPU: the sharded path on >= 2 GPUs (one process per GPU over NCCL / NVLink peer stores),
and its single-GPU emulation (several slice handles on one device, gathered by hand)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, rel_err

pytestmark = pytest.mark.gpu


def n_devices(nb):
    return nb.load_library().nbody_cuda_device_count()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_matches_single_gpu(nb, world):
    if n_devices(nb) < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29500 + world), os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    print(r.stdout[-4000:], r.stderr[-2000:])
    assert r.returncode == 0


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("mode", ["strict", "fast"])
def test_slices_emulated_on_one_gpu(nb, oracle, world, mode):
    """All `world` slice handles live on cuda:0; after every local step their position slices are
    gathered with plain device copies (what the all-gather does).  Strict mode must equal the
    unsharded reference bit for bit: each target's sources are swept in global index order whatever
    the slicing (unlike MpiNaiveSimulation, whose sum order depends on the rank count)."""
    import ctypes as C

    import torch

    n, steps = 5003, 3
    p, eps = nb.datasets.plummer_sphere(n, seed=9)
    p["mass"] = np.random.default_rng(1).uniform(0.5, 2.0, n).astype(np.float32)
    params = nb.SimParams(1.0, 0.5, eps)
    want = oracle.step_rows(p, 1.0, 0.5, eps, steps)
    counts, displs = nb.slice_counts(n, world)
    lib = nb.load_library()
    sims = []
    for r in range(world):
        s = nb.CudaNaiveSimulation(device=0, mode=mode)
        s.rank, s.world = r, world                 # slice handle without a process group
        s._ck(lib.nbody_cuda_configure(s._h, n, r, world))
        s.n_total = n
        sl = p[displs[r]:displs[r] + counts[r]].copy()
        s._ck(lib.nbody_cuda_upload_aos(s._h, sl.ctypes.data_as(C.c_void_p), counts[r]))
        sims.append(s)
    lay = nb.SlotLayout(n, world)
    per = 4 * lay.slot_stride

    def views():
        from nbody_b200.simulation import _DeviceFloats

        out = []
        for s in sims:
            s.synchronize()
            out.append(torch.as_tensor(_DeviceFloats(lib.nbody_cuda_positions(s._h), per * world), device="cuda:0"))
        return out

    def gather():
        v = views()
        for r in range(world):
            for q in range(world):
                if q != r:
                    v[q][r * per:(r + 1) * per].copy_(v[r][r * per:(r + 1) * per])
        torch.cuda.synchronize()

    gather()
    for _ in range(steps):
        for s in sims:
            s._ck(lib.nbody_cuda_step_local(s._h, params.G, params.dt, params.min_r2))
        gather()
    got = p.copy()
    for r, s in enumerate(sims):
        sl = got[displs[r]:displs[r] + counts[r]].copy()
        s.download(sl)
        got[displs[r]:displs[r] + counts[r]] = sl
        s.close()
    if mode == "strict":
        assert got.tobytes() == want.tobytes()
    else:
        assert rel_err(got["position"], want["position"]).max() < 1e-5
        assert rel_err(got["acceleration"], want["acceleration"]).max() < 5e-5
