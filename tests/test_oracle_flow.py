"""Pins the oracle's momentum step / body / force path against the reference's known answers
(test/test_flow.jl, test/test_bodies.jl, test/test_metrics.jl, test/test_simulation.jl)."""
import math

import numpy as np
import pytest

from oracle import wl_oracle as wo


def L2(a):
    sl = tuple(slice(1, -1) for _ in a.shape)
    return float(np.sum(a[sl].astype(np.float64) ** 2))


def test_schemes():
    lib = wo.load(np.float64)
    q, c, v = 0, 1, 2
    # test_flow.jl:4-9
    assert lib.wlo_scheme(v, 1, 0, 1) == 0 and lib.wlo_scheme(v, 1, 2, 1) == 2
    assert lib.wlo_scheme(v, 1, 2, 3) == 2.5 and lib.wlo_scheme(v, 3, 2, 1) == 1.5
    assert lib.wlo_scheme(c, 1, 0, 1) == 0.5 and lib.wlo_scheme(c, 1, 2, -1) == 0.5
    # quick(u,c,d) = median((5c+2d-u)/6, c, median(10c-9u,c,d))  Flow.jl:4
    assert lib.wlo_scheme(q, 0.0, 0.5, 2.0) == pytest.approx(np.median([(2.5 + 4) / 6, 0.5, np.median([5.0, 0.5, 2.0])]))
    assert lib.wlo_scheme(q, 2.0, 0.5, 0.0) == pytest.approx(np.median([(2.5 - 2) / 6, 0.5, np.median([5 - 18.0, 0.5, 0.0])]))


def test_boundary_fluxes_1d_equivalents():
    """ϕuL/ϕuR/ϕuP (test_flow.jl:18-41) exercised through conv_diff! on a flow uniform in y:
    with ν=0 and u_y=0 the x-momentum RHS at the first/last interior cell is the difference of the
    boundary-face and inner-face fluxes named by the reference's helper tests."""
    sim = wo.OracleSim((8, 8), (1.0, 0.0), nu=0.0)
    u = sim.field("u")
    prof = np.array([1.0, 1.0, 1.5, 2.0, 1.0, 0.5, 2.0, 3.0, 3.0, 3.0], np.float32)
    u[0][...] = prof[None, :]
    u[1][...] = 0
    sim.lib.wlo_op_conv_diff(sim.h, 0)
    r = sim.field("f")[0][4]            # a y-row in the middle; Julia r[:,5,1]
    q = lambda a, b, c: sim.lib.wlo_scheme(0, a, b, c)
    ph = lambda i: (prof[i] + prof[i - 1]) / 2   # ϕ(1,I,u), advecting velocity at the lower face of cell i
    # cell 1 (Julia 2): +ϕuL(face 1) - ϕu(face 2); positive flux everywhere
    lower = ph(1) * ph(1)                              # inlet, positive flux -> CD   (test_flow.jl:18)
    inner2 = ph(2) * q(prof[0], prof[1], prof[2])
    assert r[1] == pytest.approx(lower - inner2, rel=1e-6)
    # last interior cell 8 (Julia 9): +ϕu(face 8) - ϕuR(face 9); outlet positive flux -> QUICK (test_flow.jl:22)
    inner8 = ph(8) * q(prof[6], prof[7], prof[8])
    upper = ph(9) * q(prof[7], prof[8], prof[9])
    assert r[8] == pytest.approx(inner8 - upper, rel=1e-6)


def test_impulsive_box_flow():
    # test_flow.jl:76-84
    U = (2 / 3, -1 / 3)
    sim = wo.OracleSim((16, 16), U)
    sim.mom_step()
    u = sim.field("u")
    assert L2(u[0] - np.float32(U[0])) < 2e-5
    assert L2(u[1] - np.float32(U[1])) < 1e-5


def test_scheme_selection():
    # test_flow.jl:87-98
    def mk(s):
        sim = wo.OracleSim((16, 16), (1.0, 0.0), 16, U=1, dtype=np.float64, perdir=(1, 2), scheme=s,
                           u0=lambda i, x: np.sin(np.pi * x[0] / 8) if i == 1 else 0 * x[0])
        sim.sim_step()
        return sim.field("u").copy()
    assert np.max(np.abs(mk("quick") - mk("cds"))) > 1e-6


def tgv2d(i, x, t, k, nu):
    X, Y = x[0] * k, x[1] * k
    e = math.exp(-2 * k * k * nu * t)
    return (-np.sin(X) * np.cos(Y) * e) if i == 1 else (np.cos(X) * np.sin(Y) * e)


def test_periodic_tgv_2d():
    # test_flow.jl:100-109 with fixture test/helper.jl:4-15 (Re=1e8: ∂ₜuBC forcing is O(1e-10), omitted)
    L = 64; k = np.float32(2 * np.pi / L); nu = np.float32(1 / (k * 1e8))
    sim = wo.OracleSim((L, L), (0.0, 0.0), L, U=1, nu=nu, perdir=(1, 2),
                       u0=lambda i, x: tgv2d(i, x, 0.0, k, nu))
    sim.sim_step(np.pi / 100)
    t = sim.time()
    u = sim.field("u")
    for i in (1, 2):
        ue = np.broadcast_to(tgv2d(i, sim.loc(i), t, k, nu), u[i - 1].shape)
        assert L2(u[i - 1] - ue) < 1e-4


def test_increasing_body_force():
    # test_flow.jl:111-122 with fixture test/helper.jl:26-33 (Float64, periodic x, g = jerk*t)
    N = 8; jerk = 4.0; U0 = math.sqrt(N)
    sim = wo.OracleSim((N, N), (U0, 0.0), N, nu=0.001, dt=0.001, perdir=(1,), dtype=np.float64)
    sim.sim_step(1.0, dUdt=lambda t: (jerk * t, 0.0))
    u = sim.field("u")
    uF = U0 + 0.5 * jerk * sim.time() ** 2
    assert L2(u[0] - uF) < 1e-4 and L2(u[1]) < 1e-4


def test_body_kernels():
    lib = wo.load(np.float64)
    # test_bodies.jl:2-5
    assert lib.wlo_mu0(3.0, 6.0) == lib.wlo_mu0(0.5, 1.0)
    assert lib.wlo_mu0(0.0, 1.0) == 0.5
    assert lib.wlo_mu0(np.finfo(np.float64).eps - 1, 1.0) == 0
    assert lib.wlo_mu1(0.0, 2.0) == pytest.approx(2 * (1 / 4 - 1 / np.pi ** 2), rel=1e-15)


def test_hydrostatic_pressure_force():
    # test_metrics.jl:33-40 : p = y, disc radius N/4 at N/2 -> force/(πR²) ≈ [0,1]
    N = 32
    sim = wo.OracleSim((N - 2, N - 2), (0.0, 0.0), dtype=np.float64, build_poisson=False,
                       body=wo.make_body("circle", (N / 2, N / 2), N // 4))
    p = sim.field("p")
    p[...] = 0
    p[1:-1, 1:-1] = np.broadcast_to(sim.loc(0)[1], p.shape)[1:-1, 1:-1]
    force = sim.pressure_force(0.0)
    assert np.sum(np.abs(force / (np.pi * (N / 4) ** 2) - np.array([0, 1]))) < 2e-3


def test_circle_in_accelerating_flow():
    # test_flow.jl:161-173 : uBC(t) = (t,0) -> U(t₁) in BC!, dU/dt = (1,0) in accelerate!
    radius, H = 32, 16
    sim = wo.OracleSim((2 * H * radius,) * 2, (0.0, 0.0), radius, U=1,
                       body=wo.make_body("circle", (H * radius,) * 2, radius))
    kw = dict(uBC_t=lambda t: (t, 0.0), dUdt=lambda t: (1.0, 0.0))
    sim.sim_step(**kw)
    force = sim.pressure_force() / (np.pi * sim.L ** 2)
    assert np.allclose(force, [-1, 0], atol=0.04), force
    u = sim.field("u")
    assert u.max() / u[0][1, 1] > 1.91
    for _ in range(3):
        sim.sim_step(**kw)
    assert all(n <= 2 for n in sim.n), sim.n


@pytest.mark.parametrize("exitBC", [True, False])
def test_moving_bodies(exitBC):
    # test_simulation.jl:20-33
    radius = 8; nu = radius / 250; nm = (4 * radius, 4 * radius)
    sim = wo.OracleSim(nm, (1, 0), radius, nu=nu, exitBC=exitBC,
                       body=wo.make_body("circle", (2 * radius,) * 2, radius, vel=(1, 0)))
    sim.sim_step()
    u = sim.field("u")
    assert np.allclose(u[0].T[:, radius - 1], 1)          # sim.flow.u[:,radius,1] ≈ 1
    sim = wo.OracleSim(nm, (0, 0), radius, U=1, nu=nu, exitBC=exitBC,
                       body=wo.make_body("circle", (2 * radius,) * 2, radius, acc=(2, 0)))
    sim.sim_step()
    assert len(sim.n) == 2 and all(n < 5 for n in sim.n)
    assert sim.field("u").max() > sim.field("V").max() > 0


def test_sim_time_stop():
    # test_simulation.jl:15-19
    radius = 8
    sim = wo.OracleSim((4 * radius,) * 2, (1, 0), radius, nu=radius / 250,
                       body=wo.make_body("circle", (2 * radius,) * 2, radius))
    assert sim.sim_time() == 0
    sim.sim_step(0.1, remeasure=False)
    assert sim.sim_time() >= 0.1 > float(np.sum(sim.dts[:-2])) * sim.U / sim.L


def test_metrics_moments_and_viscous_known_answers():
    """test/test_metrics.jl:41-64: zero velocity gives zero viscous force and moment; a hydrostatic pressure field has no
    moment about the disc/sphere centre; a uniform-gradient or pure-shear velocity gives no net viscous force on a
    closed surface."""
    N = 32
    for D in (2, 3):
        dims = (N - 2,) * D
        c = (N / 2,) * D
        sim = wo.OracleSim(dims, (0.0,) * D, dtype=np.float64, nu=1.0, build_poisson=False,
                           body=wo.make_body("sphere", c, N // 4))
        u, p = sim.field("u"), sim.field("p")
        u[...] = 0
        assert np.all(sim.viscous_force(0.0) == 0) and np.all(sim.moment(c, True, 0.0) == 0)
        p[...] = np.broadcast_to(sim.loc(0)[1], p.shape)                        # apply!(x->x[2], p)
        m = sim.moment(c, False, 0.0)
        assert np.all(np.abs(m) < 1e-9 * np.pi * (N / 4) ** D), m
        for fld in (lambda i: sim.loc(i + 1)[i], lambda i: sim.loc(i + 1)[(i + 1) % D]):   # x[i] ; x[i%D+1]
            for i in range(D):
                u[i][...] = np.broadcast_to(fld(i), u[i].shape)
            f = sim.viscous_force(0.0)
            assert np.all(np.abs(f) < 1e-9 * N ** (D - 1)), f


def test_meanflow_of_a_steady_boundary_layer():
    """test/test_metrics.jl:66-92: the parabolic-inflow boundary-layer flow (a Function uBC) is steady, so the running
    averages equal the instantaneous fields to √eps; uu(); time bookkeeping; reset!."""
    L = 32
    prof = lambda i, x, t: np.float32(4.0 * (((x[1] + 0.5) / (2 * L)) - ((x[1] + 0.5) / (2 * L)) ** 2)) if i == 1 else 0.0
    sim = wo.OracleSim((L, L), prof, L, nu=0.001, U=1)
    mf = wo.MeanFlow(sim, uu_stats=True)
    for t in np.arange(0, 10.0001, 0.1):
        sim.sim_step(t)
        mf.update()
    tol = float(np.sqrt(np.finfo(np.float32).eps))
    u, p = sim.field("u"), sim.field("p")
    assert np.allclose(u, mf.U, rtol=0, atol=tol) and np.allclose(p, mf.P, rtol=0, atol=tol)
    for i in range(2):
        for j in range(2):
            assert np.allclose(u[i] * u[j], mf.UU[j][i], rtol=0, atol=tol)
            assert np.allclose(mf.UU[j][i] - mf.U[i] * mf.U[j], mf.uu()[j][i], rtol=0, atol=tol)
    assert np.float32(sim.time()) == np.float32(mf.time())
