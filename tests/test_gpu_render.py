"""GPU: the headless frame renderer (nbody_cuda_render_frame) -- the reference viewer's camera and
sphere markers computed by CUDA kernels, no graphics API -- against a NumPy
restatement of the same arithmetic.  Pixel coverage may differ on disc edges where float
rounding decides; the bar is < 0.3 % of pixels and identical colours elsewhere to +-1."""
import numpy as np
import pytest

from render_ref import render_reference

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dist,ax,ay", [(800.0, 0.0, 0.0), (400.0, 25.0, -40.0), (60.0, 10.0, 5.0)])
def test_frame_matches_numpy_restatement(nb, dist, ax, ay):
    w, h = 200, 150
    rng = np.random.default_rng(5)
    p = np.zeros(400, nb.PARTICLE_DTYPE)
    p["position"] = rng.uniform(-150, 150, (400, 3)).astype(np.float32)
    p["position"][:20] = rng.uniform(-20, 20, (20, 3)).astype(np.float32)      # some close to the camera axis
    p["mass"] = 1.0
    sim = nb.CudaNaiveSimulation(mode="fast")
    sim.upload(p)
    got = sim.render_frame(w, h, dist, ax, ay)
    sim.close()
    want, owner = render_reference(p["position"], w, h, dist, ax, ay)
    assert got.shape == (h, w, 3) and got.dtype == np.uint8
    covered_got, covered_want = got.sum(axis=2) > 0, want.sum(axis=2) > 0
    assert covered_want.sum() > 50                                            # the scene is visible at all
    mismatch = (covered_got != covered_want).mean()
    both = covered_got & covered_want
    diff = np.abs(got.astype(int) - want.astype(int))[both]
    print("coverage", covered_want.mean(), "mismatch", mismatch, "max colour diff", diff.max() if diff.size else 0)
    assert mismatch < 0.003
    assert (diff.max(axis=1) <= 1).mean() > 0.995                             # a different winner on a few ties only
    assert np.all(got[..., 2] == 0) and np.all(got[..., 0] >= got[..., 1])    # orange: B = 0, G = R/2


def test_empty_and_clipped_scenes_are_black(nb, tmp_path):
    p = np.zeros(10, nb.PARTICLE_DTYPE)
    p["position"][:, 2] = 5000.0            # behind the camera at distance 800
    sim = nb.CudaNaiveSimulation(mode="fast")
    sim.upload(p)
    assert sim.render_frame(64, 48).sum() == 0
    with pytest.raises(nb.NBodyCudaError):
        sim.render_frame(0, 48)
    p["position"][:, 2] = 0.0
    sim.upload(p)
    frame = sim.render_frame(64, 48)
    assert frame[24, 32, 0] > 100                                            # a sphere at the origin lights the centre pixel
    nb.write_ppm(str(tmp_path / "f.ppm"), frame)
    raw = open(tmp_path / "f.ppm", "rb").read()
    assert raw.startswith(b"P6\n64 48\n255\n") and len(raw) == len(b"P6\n64 48\n255\n") + 64 * 48 * 3
    sim.close()
