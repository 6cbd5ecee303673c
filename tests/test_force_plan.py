"""Invariants of the fast force kernel's work plan, checked on the CPU through the C ABI
(nbody_cuda_plan_describe is pure host arithmetic -- the same inline functions the kernels use).

What the kernels rely on (csrc/nbody_kernels.cuh, "Work plan of the fast force kernel"):
  * the CTAs' granule ranges tile [0, tiles * n_gran) exactly, in order, and none is empty;
  * the closed-form owner of a granule (used by the finishing kernel to find which scratch slots to
    add) is the CTA whose range contains it;
  * the target tiles cover every local target once, with fewer than 256 padded rows in all;
  * the scratch size covers slot (c + t) of every segment.
"""
import ctypes as C

import numpy as np
import pytest


def describe(nb, n_local, n_slots, sms, I=0):
    lib = nb.load_library()
    header = (C.c_longlong * 12)()
    rc = lib.nbody_cuda_plan_describe(n_local, n_slots, sms, I, header, None, 0, None, None, 0)
    assert rc == 0, (n_local, n_slots, sms, I)
    h = list(header)
    grid, tiles = h[5], h[11]
    first = (C.c_longlong * (grid + 1))()
    of, ol = (C.c_int * tiles)(), (C.c_int * tiles)()
    rc = lib.nbody_cuda_plan_describe(n_local, n_slots, sms, I, header, first, grid + 1, of, ol, tiles)
    assert rc == 0
    keys = ["I_main", "n_main", "n_tail", "tail0", "tail1", "grid", "n_gran", "cost_total", "ctas_per_sm",
            "scratch", "gran", "tiles"]
    return dict(zip(keys, h)), np.array(first[:]), np.array(of[:]), np.array(ol[:])


SHAPES = [
    # (n_local, n_slots): whole systems (n_slots = n rounded up to 256) and slices of sharded runs
    (1, 256), (2, 256), (255, 256), (256, 256), (257, 512), (1000, 1024), (2049, 2304), (4096, 4096),
    (10000, 10240), (20000, 20224), (50000, 50176), (65536, 65536), (100000, 100096), (262144, 262144),
    (1048576, 1048576), (131072, 1048576), (524288, 1048576), (350, 1024), (8192, 65536), (524288, 4194304),
    (333, 768), (1792, 1792), (1793, 2048), (3 * 1792 + 6 * 256, 7168),
]


@pytest.mark.parametrize("sms", [148, 132, 1, 7])
@pytest.mark.parametrize("forced", [0, 7, 4, 2, 1])
def test_plan_invariants(nb, sms, forced):
    for n_local, n_slots in SHAPES:
        h, first, of, ol = describe(nb, n_local, n_slots, sms, forced)
        n_gran, tiles, grid = h["n_gran"], h["tiles"], h["grid"]
        assert n_gran * h["gran"] == n_slots
        assert tiles == h["n_main"] + h["n_tail"] and 0 <= h["n_tail"] <= 2
        # ranges: ordered, non-empty, cover everything
        assert first[0] == 0 and first[-1] == tiles * n_gran
        assert (np.diff(first) > 0).all(), (n_local, n_slots, sms, forced, h)
        assert 1 <= grid <= sms * h["ctas_per_sm"]
        # targets: main tiles then narrower tails, padded rows < 256
        widths = [h["I_main"]] * h["n_main"] + [h["tail0"], h["tail1"]][: h["n_tail"]]
        assert all(w in (1, 2, 4, 7) for w in widths)
        assert all(w < h["I_main"] for w in widths[h["n_main"]:])
        covered = 256 * sum(widths)
        assert n_local <= covered < n_local + 256, (n_local, widths)
        # owner formula == the range that contains the granule (first and last granule of every tile, and a sample)
        def owner_by_search(g):
            return int(np.searchsorted(first, g, side="right") - 1)
        for t in range(tiles):
            assert of[t] == owner_by_search(t * n_gran), (n_local, n_slots, sms, forced, t)
            assert ol[t] == owner_by_search((t + 1) * n_gran - 1)
            assert of[t] <= ol[t]
        # scratch covers slot (c + t) for every segment
        assert h["scratch"] >= (grid + tiles) * 256 * h["I_main"]
        assert max(ol[t] + t for t in range(tiles)) < grid + tiles


def test_plan_balance_at_the_reference_benchmark_sizes(nb):
    """Cost per CTA is within one granule of the mean, so the busiest SM is at most a few percent over
    the average even at N = 10 000 (round 1: 2 units against a mean of 1.62)."""
    for n in (10000, 20000, 50000, 100000, 65536, 1048576):
        n_slots = (n + 255) // 256 * 256
        h, first, _, _ = describe(nb, n, n_slots, 148)
        def cost(I):                                        # plan_granule_cost() of csrc/nbody_kernels.cuh
            return 73 if I == 7 else 10 * I + 7

        cost_main, n_main_gran = cost(h["I_main"]), h["n_main"] * h["n_gran"]
        tail_cost = [cost(h["tail0"]), cost(h["tail1"])]

        def cost_of(g0, g1):
            total = 0
            a, b = g0, g1
            m = min(b, n_main_gran) - min(a, n_main_gran)
            total += m * cost_main
            for k in range(h["n_tail"]):
                lo, hi = n_main_gran + k * h["n_gran"], n_main_gran + (k + 1) * h["n_gran"]
                total += max(0, min(b, hi) - max(a, lo)) * tail_cost[k]
            return total

        costs = np.array([cost_of(first[c], first[c + 1]) for c in range(h["grid"])])
        assert costs.sum() == h["cost_total"]
        assert costs.max() <= costs.mean() + cost_main
        assert costs.max() / costs.mean() < 1.08, (n, costs.max(), costs.mean())


def test_plan_describe_rejects_bad_shapes(nb):
    lib = nb.load_library()
    header = (C.c_longlong * 12)()
    assert lib.nbody_cuda_plan_describe(0, 256, 148, 0, header, None, 0, None, None, 0) < 0
    assert lib.nbody_cuda_plan_describe(10, 300, 148, 0, header, None, 0, None, None, 0) < 0
    assert lib.nbody_cuda_plan_describe(600, 512, 148, 0, header, None, 0, None, None, 0) < 0
    assert lib.nbody_cuda_plan_describe(10, 256, 148, 0, None, None, 0, None, None, 0) < 0
