"""Runs one small case in THIS process (the library reads its A/B environment switches once per process) and prints a
digest of the result: tests/test_variants_gpu.py compares the digests of the default build with every switch set."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import waterlily_jl_b200 as wl
    case = sys.argv[1]
    if case == "sphere":
        # 12 x 8 K-MOM columns (TMA inside, cp.async at the rim); three 128-cell tiles in x, so that level 1 has all three
        # tile classes (general near the body, constants in the middle tile, constants + boundary at the ends)
        sim = wl.Simulation((384, 48, 32), (1.0, 0.0, 0.0), 12.0, nu=12.0 / 500,
                            body=wl.AutoBody(kind="sphere", center=(60, 23, 15), radius=6.0))
    else:
        N = 64
        k = np.float32(2 * np.pi / N)

        def u0(i, x):
            if i == 1:
                return -np.sin(k * x[0]) * np.cos(k * x[1]) * np.cos(k * x[2])
            if i == 2:
                return np.cos(k * x[0]) * np.sin(k * x[1]) * np.cos(k * x[2])
            return 0 * x[0] * x[1] * x[2]
        sim = wl.Simulation((N, N, N), (0.0, 0.0, 0.0), float(N), perdir=(1, 2, 3), nu=float(1 / (k * 1600)), U=1.0, u0=u0)
    for _ in range(4):
        sim.sim_step(remeasure=False)
    h = hashlib.sha256()
    for a in (sim.flow.u, sim.flow.p, np.asarray(sim.flow.Δt, np.float32), np.asarray(sim.pois.n, np.int32)):
        h.update(np.ascontiguousarray(a).tobytes())
    print("DIGEST", case, h.hexdigest(), list(sim.pois.n), sim.tile_classes(0), flush=True)
    sim.close()


if __name__ == "__main__":
    main()
