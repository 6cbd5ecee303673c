"""GPU: `nbody_headless_cuda` -- the reference's unmodified NBodySimulation facade, generators,
CheckpointManager and PerformanceLogger with CudaNaiveSimulation plugged in.
In strict mode its trajectory file must be byte-identical to the one the reference's own sequential
algorithm and checkpoint writer produce from the same seeded input."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

APP = os.path.join(ROOT, "n-body-simulation_b200", "host", "bin", "nbody_headless_cuda")


def run_app(cwd, *args):
    r = subprocess.run([APP, *args], cwd=cwd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout


@pytest.fixture(scope="module")
def app():
    if not os.path.exists(APP):
        pytest.skip("nbody_headless_cuda not built (needs the reference tree at build time)")
    return APP


def test_strict_trajectory_file_is_byte_identical_to_the_reference(app, reference, tmp_path):
    n, steps, seed = 4096, 20, 12345
    out = run_app(tmp_path, "--random", str(n), "--steps", str(steps), "--seed", str(seed), "--mode", "strict",
                  "--checkpoint", "cuda.bin", "--out", "cuda.csv")
    assert "CudaNaiveSimulation (strict)" in out and "Simulation completed in" in out
    # the same run through the reference's own algorithm and writer
    p, eps = reference.disk(n, seed)
    ref_file = str(tmp_path / "ref.bin")
    reference.checkpoint_begin(ref_file, p, steps)
    q = p
    for _ in range(steps):
        q = reference.naive_steps(q, 1.0, 0.5, eps, 1)
        reference.checkpoint_write_step(ref_file, q)
    a, b = open(tmp_path / "cuda.bin", "rb").read(), open(ref_file, "rb").read()
    assert len(a) == 24 + 4 * n + 12 * n * steps
    assert a == b
    line = open(tmp_path / "cuda.csv").read().strip().split(",")      # PerformanceLogger.hpp:31
    assert line[0] == str(n) and line[1] == str(steps) and float(line[2]) > 0 and line[3] == "1"


def test_resident_mode_with_sparse_checkpoints(app, reference, tmp_path):
    n, steps, seed = 3000, 20, 7
    run_app(tmp_path, "--plummer", str(n), "--steps", str(steps), "--seed", str(seed), "--mode", "strict", "--resident",
            "--checkpoint-every", "5", "--checkpoint", "res.bin", "--out", "res.csv")
    p, eps = reference.plummer(n, seed)
    assert reference.checkpoint_read_header(str(tmp_path / "res.bin")) == (n, steps, 4)
    for k in range(4):
        want = reference.naive_steps(p, 1.0, 0.5, eps, 5 * (k + 1))
        pos, got = reference.checkpoint_read_step(str(tmp_path / "res.bin"), n, k)
        assert got == n and pos.tobytes() == np.ascontiguousarray(want["position"]).tobytes()


def test_fast_mode_run_reports_energy_and_rate(app, tmp_path):
    out = run_app(tmp_path, "--plummer", "16384", "--steps", "10", "--seed", "1", "--disable-output", "--resident",
                  "--energy-every", "5", "--out", "fast.csv")
    e = [float(x) for x in re.findall(r"E = ([-+0-9.e]+)", out)]
    assert len(e) == 3 and abs(e[2] - e[0]) / abs(e[0]) < 1e-2
    assert re.search(r"pairwise interactions/s", out) and not os.path.exists(tmp_path / "simulation_output.bin")


def test_two_gpus_same_trajectory_as_one(app, nb, tmp_path):
    if nb.load_library().nbody_cuda_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    common = ["--plummer", "20000", "--steps", "8", "--seed", "3", "--mode", "strict", "--resident", "--checkpoint-every", "4"]
    run_app(tmp_path, *common, "--checkpoint", "one.bin", "--out", "one.csv")
    run_app(tmp_path, *common, "--gpus", "2", "--checkpoint", "two.bin", "--out", "two.csv")
    assert open(tmp_path / "one.bin", "rb").read() == open(tmp_path / "two.bin", "rb").read()
    assert open(tmp_path / "two.csv").read().strip().endswith(",2")


def test_frames_are_written(app, tmp_path):
    run_app(tmp_path, "--random", "5000", "--steps", "4", "--seed", "2", "--disable-output", "--resident",
            "--frame-every", "2", "--frame-size", "160x120", "--camera", "900,20,30", "--out", "f.csv")
    for step in (2, 4):
        raw = open(tmp_path / f"frame_{step:06d}.ppm", "rb").read()
        head = b"P6\n160 120\n255\n"
        assert raw.startswith(head) and len(raw) == len(head) + 160 * 120 * 3
        body = np.frombuffer(raw[len(head):], np.uint8).reshape(120, 160, 3)
        assert (body.sum(axis=2) > 0).mean() > 0.01 and body[..., 2].max() == 0      # orange discs on black
