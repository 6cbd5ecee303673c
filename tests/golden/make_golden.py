"""Regenerates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/libnbody_ref.so, i.e.
the reference's NaiveSimulation / MpiNaiveSimulation / generators compiled unmodified by
oracle/Makefile).  Run in a container that has /root/reference:

    make -C oracle && python tests/golden/make_golden.py

The reference has no golden vectors of its own; these are outputs of its code on
seeded inputs from its own generators (Random.hpp engine seeded through getRandomEngine()).
Headless parameters: G = 1, dt = 0.5 (src/apps/headless_main.cpp:33-36), min_r2 as set by the
generator (src/Simulation.cpp:131,195) or 2.0 where the generator sets none.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import Reference  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
G, DT = 1.0, 0.5


def main():
    ref = Reference()
    cases = {}

    p, eps = ref.disk(256, 12345)
    cases["disk_n256"] = dict(initial=p, min_r2=eps, steps=[1, 10])
    p, eps = ref.plummer(512, 7)
    cases["plummer_n512"] = dict(initial=p, min_r2=eps, steps=[1, 10])
    p = ref.clusters(4, 64, 99)
    cases["clusters_4x64"] = dict(initial=p, min_r2=2.0, steps=[1, 5])
    p = ref.solar()
    cases["solar_9"] = dict(initial=p, min_r2=2.0, steps=[1, 20])
    p, eps = ref.disk(300, 4242)          # not a multiple of any tile size
    cases["disk_n300"] = dict(initial=p, min_r2=eps, steps=[3])

    for name, c in cases.items():
        out = dict(initial=c["initial"], G=np.float32(G), dt=np.float32(DT), min_r2=np.float32(c["min_r2"]))
        for s in c["steps"]:
            out[f"naive_step{s}"] = ref.naive_steps(c["initial"], G, DT, c["min_r2"], s)
        # MpiNaiveSimulation through the thread-rank shim: summation order depends on the rank count
        for nranks in (1, 3, 4):
            out[f"mpi_np{nranks}_step2"] = ref.mpi_naive_steps(c["initial"], G, DT, c["min_r2"], 2, nranks=nranks)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print(name, c["initial"].shape[0], "particles, min_r2 =", c["min_r2"])

    # layout facts the C ABI relies on (include/Particle.hpp:53-59, SimulationAlgorithm.h:4-9, SimulationOutput.h:8-13)
    lay = ref.layout()
    lay["default_simparams"] = ref.default_simparams()
    np.savez(os.path.join(OUT, "layout.npz"), **{k: np.array(v) for k, v in lay.items()})
    print(lay)


if __name__ == "__main__":
    main()
