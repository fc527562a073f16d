"""Function uBC(i,x,t) / g(i,x,t) through the device path (wl_set_bc_tables, wl_set_accel_field): CUDA vs oracle.
First run on a B200 in round 2 (profiles/r2_c01_pytest_fnbc_ungated.log: 5 passed)."""
import numpy as np
import pytest

from oracle import wl_oracle as wo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wl():
    import __graft_entry__ as ge
    ge.build()
    import waterlily_jl_b200 as wl
    return wl


def profile(L):
    return lambda i, x, t: (np.float32(4.0) * (((x[1] + np.float32(0.5)) / np.float32(2 * L)) - ((x[1] + np.float32(0.5)) / np.float32(2 * L)) ** 2)
                            if i == 1 else 0 * x[0])


@pytest.mark.parametrize("exitBC", [False, True])
def test_boundary_layer_profile_tables(wl, exitBC):
    """make_bl_flow (test/helper.jl:18-21, test/test_flow.jl:134-140): parabolic inflow as a Function uBC through
    wl_set_bc_tables, CUDA vs oracle over 20 steps."""
    L = 32
    prof = profile(L)
    sim = wl.Simulation((L, L), (0.0, 0.0), L, nu=0.001, U=1, exitBC=exitBC)
    sim.set_bc_function(prof)
    sim.apply_ic(lambda i, x: prof(i, x, np.float32(0)))
    ref = wo.OracleSim((L, L), lambda i, x, t: float(prof(i, x, t)), L, nu=0.001, U=1, exitBC=exitBC)
    assert np.array_equal(sim.flow.u, ref.field("u"))
    for _ in range(20):
        sim.sim_step()
        ref.sim_step()
    assert sim.pois.n == ref.n
    assert np.allclose(sim.flow.Δt, ref.dts, rtol=2e-6)
    assert float(np.abs(sim.flow.u - ref.field("u")).max()) <= 1e-6
    u = sim.flow.u
    assert np.allclose(u[0][:, 0], u[0][:, -1], rtol=float(np.sqrt(np.finfo(np.float32).eps)), atol=0) or exitBC


def test_bc_tables_3d_kernel(wl):
    """One BC! with a space-varying uBC on random data, 3D with a periodic direction: CUDA table kernel vs oracle BC_fn."""
    from test_oracle_fnbc import bc_fn
    dims = (8, 6, 4)
    ub = lambda i, x, t: np.float32(0.3 * i) + np.float32(0.1) * x[i % 3] * x[(i + 1) % 3]
    for perdir, exitBC in (((), False), ((3,), True), ((1, 2), False)):
        sim = wl.Simulation(dims, (0.0, 0.0, 0.0), 4.0, U=1, perdir=perdir, exitBC=exitBC)
        sim.set_bc_function(ub)
        a = np.random.default_rng(3).standard_normal(sim.field_shape("u")).astype(np.float32)
        sim.upload("u", a)
        sim.debug_op(0)                                   # BC!(u, uBC, exitBC, perdir)
        b = a.copy()
        bc_fn(b, lambda i, x, t: float(ub(i, x, t)), saveexit=exitBC, perdir=perdir)
        assert np.array_equal(sim.flow.u, b), (perdir, exitBC)
        sim.set_bc_function(None)


def test_constant_jerk_through_g(wl):
    """test/test_flow.jl:120-131 in Float32: g(i,x,t) = (jerk·t, 0) as a sampled acceleration field, CUDA vs oracle."""
    N, jerk = 8, 4.0
    g = lambda i, x, t: (t * np.float32(jerk) + 0 * x[0]) if i == 1 else 0 * x[0]
    kw = dict(nu=0.001, dt=0.001, perdir=(1,))
    sim = wl.Simulation((N, N), (float(np.sqrt(N)), 0.0), N, **kw)
    sim.set_accel_function(g)
    ref = wo.OracleSim((N, N), (float(np.sqrt(N)), 0.0), N, g=lambda i, x, t: float(g(i, x, t)), **kw)
    for _ in range(25):
        sim.sim_step()
        ref.sim_step()
    assert sim.pois.n == ref.n and np.allclose(sim.flow.Δt, ref.dts, rtol=2e-6)
    assert float(np.abs(sim.flow.u - ref.field("u")).max()) <= 2e-6 * np.sqrt(N)


def test_rotating_frame_fields_3d(wl):
    """Space- and time-varying g in 3D through the fused K-MOM's field slot (and the band patch's): CUDA vs oracle."""
    dims = (16, 12, 8)
    g = lambda i, x, t: np.float32(0.01) * x[i % 3] - np.float32(0.02) * t * x[(i + 1) % 3]
    kw = dict(nu=0.01, perdir=(3,))
    body = dict(kind="sphere", center=(6.0, 6.0, 4.0), radius=2.5)
    sim = wl.Simulation(dims, (1.0, 0.0, 0.0), 5.0, body=wl.AutoBody(**body), **kw)
    sim.set_accel_function(g)
    ref = wo.OracleSim(dims, (1.0, 0.0, 0.0), 5.0, body=wo.make_body(**body), g=lambda i, x, t: float(g(i, x, t)), **kw)
    for _ in range(4):
        sim.sim_step(remeasure=False)
        ref.sim_step(remeasure=False)
    assert sim.pois.n == ref.n and np.allclose(sim.flow.Δt, ref.dts, rtol=2e-6)
    assert float(np.abs(sim.flow.u - ref.field("u")).max()) <= 2e-6
    sim.set_accel_function(None)                       # fields cleared: the uniform term alone again
    sim.sim_step(remeasure=False)
