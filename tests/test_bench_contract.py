"""bench.py contract: the JSON line of both arms carries the keys the driver reads.  The reference
arm runs on CPU; our arm needs a GPU (and must fail loudly without one)."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT, have_gpu

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "cpu_baseline"}


def run_bench(*args, timeout=600):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                       timeout=timeout, cwd=ROOT)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    return r, (json.loads(lines[-1]) if lines else None)


def test_reference_arm_line():
    r, d = run_bench("--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-n", "2048")
    assert r.returncode == 0 and d is not None, r.stderr[-2000:]
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["metric"] == "pairwise interactions/sec" and d["unit"] == "interactions/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["vs_baseline"] is None and d["config"]["workload"].startswith("c4")
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert set(d["config"]) == {"workload", "n", "generator", "generator_source", "G", "dt", "min_r2", "l2"}
    own = d["reference_own_clock"]          # the reference's own definition of seconds: step loop + per-step gather
    assert own is None or (own["value"] > 0 and own["ms_per_step"] > 0)


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, env=env, cwd=ROOT)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_our_arm_fails_loudly_without_a_gpu():
    if have_gpu():
        pytest.skip("a GPU is present")
    r, d = run_bench("--steps", "1", "--warmup", "0", "--particles", "1024", timeout=300)
    assert r.returncode != 0 and d is None


@pytest.mark.gpu
def test_our_arm_line():
    r, d = run_bench("--particles", "32768", "--steps", "4", "--warmup", "3")
    assert r.returncode == 0 and d is not None, r.stderr[-2000:]
    assert BASE_KEYS | {"clocks", "roofline"} <= set(d)
    assert d["n_gpus"] == 1 and d["steps"] == 4 and d["warmup"] == 3 and d["dtype"] == "f32" and d["data"] == "synthetic"
    assert d["gpu_launches"] >= 2 * d["steps"] and d["value"] > 1e11
    rf = d["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(rf) and 0.1 < rf["frac"] < 1.0
    assert abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 32768 * 40 == e["d2h_bytes_per_step"] and e["value"] > 1e11
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] == 1 and cb["value"] > 0
    assert "sm_mhz" in d["clocks"] and "reasons" in d["clocks"]
    # the general (unequal-mass) loop is measured in the same run, by the same per-launch method
    g = d["roofline_general"]
    assert g["launches_timed"] >= 1 and 0.1 < g["frac"] < 1.0 and abs(g["frac"] - g["achieved"] / g["peak"]) < 1e-9
    # both arms describe the workload with the same dictionary
    r2, ref = run_bench("--impl", "reference", "--particles", "32768", "--steps", "1", "--warmup", "0", "--ref-n", "2048")
    assert r2.returncode == 0 and ref["config"] == d["config"]
    assert ref["reference_own_clock"]["ms_per_step"] > 0
