"""Checkpoint / restart through the device path, after test/test_ioext.jl:5-62 (vtk restart, save!/load! of a
simulation, save!/load! of a MeanFlow).  Runs last (file name) so the parity tests proper come first."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wl():
    import __graft_entry__ as ge
    ge.build()
    import waterlily_jl_b200 as wl
    return wl


def sphere_sim(wl, D, L=16):
    dims = (L, L, L)[:D]
    body = wl.AutoBody("sphere", center=(L / 2,) * D, radius=L / 4)
    return wl.Simulation(dims, (1.0,) + (0.0,) * (D - 1), L / 2, body=body, nu=0.05)


@pytest.mark.parametrize("D", [2, 3])
def test_vtk_restart(wl, D, tmp_path, monkeypatch):
    from waterlily_jl_b200 import io as wio
    monkeypatch.chdir(tmp_path)
    sim = sphere_sim(wl, D)
    wr = wio.VTKWriter(f"test_vtk_reader_{D}", dir="TEST_DIR")
    sim.sim_step(1.0)
    wr.save(sim)
    wr.close()
    restart = sphere_sim(wl, D)
    wio.load_pvd(restart, f"test_vtk_reader_{D}.pvd")
    assert np.array_equal(sim.flow.p, restart.flow.p)
    assert np.array_equal(sim.flow.u, restart.flow.u)
    assert np.array_equal(sim.flow.mu0, restart.flow.mu0)
    assert sim.flow.Δt[-1] == restart.flow.Δt[-1]
    assert abs(sim.sim_time() - restart.sim_time()) < 1e-3
    sim.sim_step(remeasure=False)                       # the restarted run continues like the original
    restart.sim_step(remeasure=False)
    assert np.allclose(sim.flow.u, restart.flow.u, rtol=0, atol=1e-6)
    assert np.allclose(sim.flow.Δt[-1], restart.flow.Δt[-1], rtol=1e-6)


def test_npz_checkpoint(wl, tmp_path):
    from waterlily_jl_b200 import io as wio
    sim1, sim2 = sphere_sim(wl, 3), sphere_sim(wl, 3)
    sim1.sim_step(remeasure=False)
    sim1.sim_step(remeasure=False)
    wio.save("sim1_sphere.ckpt", sim1, dir=str(tmp_path))
    wio.load(sim2, fname="sim1_sphere.ckpt", dir=str(tmp_path))
    assert np.array_equal(sim1.flow.p, sim2.flow.p) and np.array_equal(sim1.flow.u, sim2.flow.u)
    assert [float(x) for x in sim1.flow.Δt] == [float(x) for x in sim2.flow.Δt]
    sim1.sim_step(remeasure=False)
    sim2.sim_step(remeasure=False)
    assert np.allclose(sim1.flow.u, sim2.flow.u, rtol=0, atol=1e-6)
    assert sim1.sim_time() == sim2.sim_time()


def test_meanflow_checkpoint(wl, tmp_path):
    from waterlily_jl_b200 import io as wio
    sim = sphere_sim(wl, 3)
    m1, m2 = wl.MeanFlow(sim, uu_stats=True), wl.MeanFlow(sim, uu_stats=True)
    for _ in range(2):
        sim.sim_step(remeasure=False)
        m1.update()
    P, U, UU = m1.P, m1.U, m1.UU
    wio.save_meanflow("meanflow.ckpt", m1, dir=str(tmp_path))
    m1.reset()                                          # m1 and m2 share the simulation's device arrays
    assert float(np.abs(m2.U).max()) == 0.0
    wio.load_meanflow(m2, fname="meanflow.ckpt", dir=str(tmp_path))
    assert np.array_equal(m2.P, P) and np.array_equal(m2.U, U) and np.array_equal(m2.UU, UU)
    assert len(m2.t) == 3
