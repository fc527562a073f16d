#!/usr/bin/env python
"""Generates the golden fixtures of this directory from the CPU oracle (oracle/wl_oracle.cpp).

The reference stores no field dumps and cannot run here (Julia is absent from this image and from the
GPU box), so these vectors come from the oracle -- which is pinned on the reference's known-answer tests
(tests/test_oracle_*.py) and agrees bit for bit with a second independent restatement
(tests/test_oracle_restatement2.py).  They freeze that state: the oracle must keep reproducing them
(tests/test_golden.py, CPU) and the CUDA path must match them through the C ABI (tests/test_parity_gpu.py).

usage: python tests/golden/make_golden.py        (rewrites tests/golden/*.npz)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import wl_oracle as wo  # noqa: E402


def tgv3d(i, x, k):
    if i == 1:
        return -np.sin(k * x[0]) * np.cos(k * x[1]) * np.cos(k * x[2])
    if i == 2:
        return np.cos(k * x[0]) * np.sin(k * x[1]) * np.cos(k * x[2])
    return 0 * x[0] * x[1] * x[2]


def cases():
    """name -> (constructor kwargs, number of sim_step!, remeasure).  Shared with the tests."""
    m = 64
    R = m / 8
    out = {}
    # BASELINE config A: README circle 96x64, Re=100, 100 steps (README.md:41-58)
    out["A_circle_96x64"] = (dict(dims=(96, 64), uBC=(1.0, 0.0), L=2 * R, nu=2 * R / 100,
                                  body=dict(kind="circle", center=(m / 2 - 1, m / 2 - 1), radius=R)), 100, False)
    # config B reduced to 16^3: periodic TGV
    N = 16
    k = np.float32(2 * np.pi / N)
    out["B_tgv_16"] = (dict(dims=(N, N, N), uBC=(0.0, 0.0, 0.0), L=N, perdir=(1, 2, 3), nu=float(1 / (k * 1600)), U=1.0,
                            u0=("tgv", float(k))), 10, False)
    # config C reduced to 32x16x16: static sphere, convective exit
    m = 16
    R = m / 8
    out["C_sphere_32x16x16_exit"] = (dict(dims=(2 * m, m, m), uBC=(1.0, 0.0, 0.0), L=2 * R, nu=2 * R / 3700, exitBC=True,
                                          body=dict(kind="sphere", center=(m / 2 - 1,) * 3, radius=R)), 8, False)
    # config D reduced to 32x32x8: oscillating cylinder, periodic z, remeasure every step
    R = 4.0
    out["D_cylinder_32x32x8"] = (dict(dims=(32, 32, 8), uBC=(0.0, 0.0, 0.0), L=2 * R, U=1.0, nu=2 * R / 1000, perdir=(3,),
                                      body=dict(kind="cylinder", axis=2, center=(16, 16, 0), radius=R, amp=(R, 0, 0),
                                                omega=2 * np.pi / (8 * R))), 6, True)
    return out


def widened_cases():
    """Rows added after the first fixtures: name -> (runner(leaf ctor, Simulation ctor) -> sim, has body).
    Frozen for the oracle here; the CUDA path is compared with the live oracle in tests/test_parity_gpu.py
    (`test_csg_rigid_bodies`) until these fixtures have been run against it on the box."""
    def rotating_annulus_with_capsule(leaf, Sim):
        ring = leaf("circle", (14.0, 12.0), 0.3, radius=6.0, omega=0.05) - leaf("circle", (14.0, 12.0), 0.3, radius=3.0, omega=0.05)
        body = ring + leaf("capsule", (22.0, 6.0), -0.4, radius=1.5, minor=3.0, axis=0, V=(0.0, 0.1))
        sim = Sim((32, 24), (1.0, 0.0), 8.0, body=body, nu=0.05)
        for n in range(6):
            th = np.float32(0.3) + np.float32(0.05) * np.float32(sum(sim.flow_dts()))
            sim.body = sim.body.setmap(theta=float(th))
            sim.sim_step(remeasure=True)
        return sim
    return {"E_csg_annulus_capsule_32x24": rotating_annulus_with_capsule}


def build(mod, make_body, kw):
    """Instantiate `mod.Simulation`-like object (oracle or CUDA mirror) from a case description."""
    kw = dict(kw)
    dims, uBC, L = kw.pop("dims"), kw.pop("uBC"), kw.pop("L")
    body = kw.pop("body", None)
    u0 = kw.pop("u0", None)
    if u0 is not None:
        kk = np.float32(u0[1])
        kw["u0"] = lambda i, x: tgv3d(i, x, kk)
    return dims, uBC, L, (make_body(**body) if body else None), kw


def main():
    for name, (kw, nsteps, remeasure) in cases().items():
        dims, uBC, L, body, rest = build(wo, wo.make_body, kw)
        sim = wo.OracleSim(dims, uBC, L, body=body, **rest)
        for _ in range(nsteps):
            sim.sim_step(remeasure=remeasure)
        force = sim.pressure_force() if body is not None else np.zeros(len(dims))
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, u=sim.field("u"), p=sim.field("p"), dt=np.asarray(sim.dts, np.float32),
                            n=np.asarray(sim.n, np.int16), force=np.asarray(force, np.float64))
        print(name, os.path.getsize(path), "bytes", "n =", sim.n[-4:], "dt =", sim.dts[-1])
    for name, run in widened_cases().items():
        sim = run(wo.CsgBody.leaf, _OracleWithDts)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, u=sim.field("u"), p=sim.field("p"), dt=np.asarray(sim.dts, np.float32),
                            n=np.asarray(sim.n, np.int16), force=np.asarray(sim.pressure_force(), np.float64),
                            mu0=sim.field("mu0"), V=sim.field("V"))
        print(name, os.path.getsize(path), "bytes", "n =", sim.n[-4:], "dt =", sim.dts[-1])


class _OracleWithDts(wo.OracleSim):
    def flow_dts(self):
        return [float(x) for x in self.dts]


if __name__ == "__main__":
    main()
