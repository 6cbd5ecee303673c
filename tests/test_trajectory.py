"""CPU: the trajectory file written by this repo's TrajectoryWriter is byte-identical to the one
the reference's CheckpointManager writes (src/CheckpointManager.cpp), both readers agree, and
the restart sidecar round-trips."""
import os

import numpy as np
import pytest


def states(nb, n, k):
    p, _ = nb.datasets.plummer_sphere(n, seed=5)
    out = []
    for s in range(k):
        q = p.copy()
        q["position"] += np.float32(0.25 * (s + 1))
        out.append(q)
    return p, out


def test_writer_matches_reference_checkpoint_manager(nb, reference, tmp_path):
    n, k = 321, 4
    p, st = states(nb, n, k)
    ref_file, our_file = str(tmp_path / "ref.bin"), str(tmp_path / "ours.bin")
    reference.checkpoint_begin(ref_file, p, 9)
    w = nb.TrajectoryWriter(our_file, p["mass"], 9)
    for q in st:
        reference.checkpoint_write_step(ref_file, q)
        w.write_step(q["position"])
    assert open(ref_file, "rb").read() == open(our_file, "rb").read()
    # and each reader reads the other's file
    r = nb.TrajectoryReader(ref_file)
    assert (r.n, r.target_steps, r.passed_steps) == (n, 9, k) == reference.checkpoint_read_header(our_file)
    assert r.masses().tobytes() == np.ascontiguousarray(p["mass"]).tobytes()
    for i, q in enumerate(st):
        assert r.step(i).tobytes() == np.ascontiguousarray(q["position"]).tobytes()
        pos, got = reference.checkpoint_read_step(our_file, n, i)
        assert got == n and pos.tobytes() == np.ascontiguousarray(q["position"]).tobytes()
    assert r.step(k) is None and reference.checkpoint_read_step(our_file, n, k)[1] == 0


def test_file_size_law(nb, tmp_path):
    n, k = 100, 3
    p, st = states(nb, n, k)
    path = str(tmp_path / "t.bin")
    w = nb.TrajectoryWriter(path, p["mass"], 5)
    for q in st:
        w.write_step(q["position"])
    assert os.path.getsize(path) == 24 + 4 * n + 12 * n * k          # SURVEY s4 test 7
    with pytest.raises(ValueError):
        w.write_step(np.zeros((n + 1, 3), np.float32))


def test_restart_round_trip(nb, tmp_path):
    p, _ = nb.datasets.galaxy_disk(777, seed=9)
    p["acceleration"] = 1.5
    path = str(tmp_path / "run.restart")
    nb.write_restart(path, p, nb.SimParams(0.5, 0.125, 3.0), 42)
    q, params, step = nb.read_restart(path)
    assert q.tobytes() == p.tobytes() and step == 42 and (params.G, params.dt, params.min_r2) == (0.5, 0.125, 3.0)
    open(path, "r+b").truncate(100)
    with pytest.raises(ValueError):
        nb.read_restart(path)
