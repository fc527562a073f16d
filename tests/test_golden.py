"""The oracle keeps reproducing the committed golden fixtures (tests/golden/*.npz, made by
tests/golden/make_golden.py).  OpenMP reductions may round a Float64 sum differently with another thread
count, so fields are compared to 1e-6 (absolute, velocities are O(1)) and iteration counts exactly."""
import os
import sys

import numpy as np
import pytest

from oracle import wl_oracle as wo

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden as mg  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def compare(name, u, p, dts, n, force, utol=1e-6, ptol=1e-5):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    assert list(n) == list(g["n"]), (list(n), list(g["n"]))
    assert np.allclose(np.asarray(dts, np.float32), g["dt"], rtol=2e-6)
    assert float(np.abs(u - g["u"]).max()) <= utol
    assert float(np.abs(p - g["p"]).max()) <= ptol * max(1e-30, float(np.abs(g["p"]).max()))
    assert np.allclose(force, g["force"], rtol=1e-4, atol=1e-4 * max(1e-30, float(np.abs(g["force"]).max())))


@pytest.mark.parametrize("name", sorted(mg.cases()))
def test_oracle_reproduces_golden(name):
    kw, nsteps, remeasure = mg.cases()[name]
    dims, uBC, L, body, rest = mg.build(wo, wo.make_body, kw)
    sim = wo.OracleSim(dims, uBC, L, body=body, **rest)
    for _ in range(nsteps):
        sim.sim_step(remeasure=remeasure)
    force = sim.pressure_force() if body is not None else np.zeros(len(dims))
    compare(name, sim.field("u"), sim.field("p"), sim.dts, sim.n, force)


@pytest.mark.parametrize("name", sorted(mg.widened_cases()))
def test_oracle_reproduces_widened_golden(name):
    sim = mg.widened_cases()[name](wo.CsgBody.leaf, mg._OracleWithDts)
    compare(name, sim.field("u"), sim.field("p"), sim.dts, sim.n, sim.pressure_force())
    g = np.load(os.path.join(GOLD, name + ".npz"))
    assert float(np.abs(sim.field("mu0") - g["mu0"]).max()) <= 2e-6 and float(np.abs(sim.field("V") - g["V"]).max()) <= 2e-6
