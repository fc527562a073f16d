"""The C++ oracle against a second, independent restatement (oracle/restate2.py: plain Python written
from the Julia source with Julia's own 1-based index arithmetic) on tiny grids, bit for bit.
SURVEY §8c "residual risk": two restatements that share no code and no index convention must agree
on every cell, ghosts included, for every branch of conv_diff!, BDIM!, BC! and CFL."""
import ctypes as C

import numpy as np
import pytest

from oracle import restate2 as r2
from oracle import wl_oracle as wo


def _rand(sim, names, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    for nm in names:
        a = sim.field(nm)
        a[...] = (scale * rng.standard_normal(a.shape)).astype(np.float32)


CASES = [((6, 5), ()), ((6, 5), (1,)), ((6, 5), (2,)), ((6, 5), (1, 2)),
         ((5, 4, 4), ()), ((5, 4, 4), (3,)), ((4, 5, 4), (1, 2, 3)), ((4, 4, 5), (2,))]


@pytest.mark.parametrize("dims,perdir", CASES)
@pytest.mark.parametrize("scheme", ["quick", "vanLeer", "cds"])
def test_conv_diff_two_restatements(dims, perdir, scheme):
    """conv_diff! (Flow.jl:38-62): r on every cell (upper ghost rows included) and the Φ left in σ."""
    sim = wo.OracleSim(dims, (1.0,) + (0.0,) * (len(dims) - 1), nu=0.07, perdir=perdir, scheme=scheme, build_poisson=False)
    _rand(sim, ["u", "sigma"], seed=5)
    u = sim.field("u").copy()
    sig = sim.field("sigma").copy()
    r = np.full_like(u, 7.0)
    r2.conv_diff(r.T, u.T, sig.T, r2.SCHEME[scheme], 0.07, perdir)
    sim.lib.wlo_op_conv_diff(sim.h, 0)
    assert np.array_equal(sim.field("f"), r)
    assert np.array_equal(sim.field("sigma"), sig)


@pytest.mark.parametrize("dims", [(6, 5), (5, 4, 4)])
def test_bdim_two_restatements(dims):
    """BDIM! (Flow.jl:176-180): f on ALL indices, then u += μddn + V + μ₀·f inside."""
    sim = wo.OracleSim(dims, (1.0,) + (0.0,) * (len(dims) - 1), dt=0.25, build_poisson=False)
    _rand(sim, ["u", "u0", "f", "V", "mu0", "mu1"], seed=9)
    a = {nm: sim.field(nm).copy() for nm in ("u", "u0", "f", "V", "mu0", "mu1")}
    r2.BDIM(a["u"].T, a["u0"].T, a["f"].T, a["V"].T, a["mu0"].T, a["mu1"].T, 0.25)
    sim.lib.wlo_op_bdim(sim.h)
    assert np.array_equal(sim.field("f"), a["f"])
    assert np.array_equal(sim.field("u"), a["u"])


@pytest.mark.parametrize("shape,perdir,saveexit", [((6, 7), (), False), ((6, 7), (2,), True), ((6, 7), (1, 2), False),
                                                   ((5, 6, 4), (), True), ((5, 6, 4), (3,), False), ((5, 6, 4), (1, 3), True)])
def test_BC_two_restatements(shape, perdir, saveexit):
    """BC! (core.jl:200-219) incl. corners, whose values depend on the (i,j) pass order."""
    lib = wo.load(np.float32)
    D = len(shape)
    U = (1.0, 0.5, -0.25)[:D]
    a = np.random.default_rng(2).random((D,) + shape[::-1]).astype(np.float32)
    b = a.copy()
    ng = (C.c_int * 3)(*(list(shape) + [1] * (3 - D)))
    per = (C.c_int * 3)(*[1 if d + 1 in perdir else 0 for d in range(3)])
    lib.wlo_BC(a.ctypes.data, D, ng, (lib.real * 3)(*(list(U) + [0.0] * (3 - D))), int(saveexit), per)
    r2.BC(b.T, U, saveexit, perdir)
    assert np.array_equal(a, b)


@pytest.mark.parametrize("dims", [(6, 5), (5, 4, 4)])
def test_cfl_two_restatements(dims):
    """CFL (Flow.jl:234-244): the maximum runs over the whole σ, stale ghost values included."""
    sim = wo.OracleSim(dims, (1.0,) + (0.0,) * (len(dims) - 1), nu=0.03, build_poisson=False)
    _rand(sim, ["u"], seed=4)
    sig = sim.field("sigma")
    sig[...] = 0
    sig[(-1,) * len(dims)] = 50.0          # a stale Φ in the upper ghost corner binds the maximum
    u, s2 = sim.field("u").copy(), sig.copy()
    dt = r2.CFL(u.T, s2.T, 0.03)
    assert np.float32(sim.cfl()) == dt
    assert np.array_equal(sim.field("sigma"), s2)
    sig[(-1,) * len(dims)] = 0.0
    s2[(-1,) * len(dims)] = 0.0
    assert np.float32(sim.cfl()) == r2.CFL(u.T, s2.T, 0.03)


# ------------------------------------------------------------------------------------------ pressure side
def _levels(sim):
    """Copies of every level's arrays in Julia order (what restate2 works on)."""
    per = tuple(d + 1 for d in range(sim.D) if sim.cfg.perdir[d])
    return [dict({k: sim.field(k, l).copy().T for k in ("x", "r", "eps", "z", "L", "D", "iD")}, perdir=per)
            for l in range(sim.nlevels)]


def _same(sim, lev, names, l):
    for k in names:
        assert np.array_equal(sim.field(k, l), lev[l][k].T), (k, l)


MG_CASES = [((12, 8), (), dict(kind="sphere", center=(5, 4), radius=2.5)),
            ((8, 12), (1,), None),
            ((8, 8, 4), (3,), dict(kind="sphere", center=(4, 4, 2), radius=2.0)),
            ((8, 4, 6), (1, 2, 3), None)]


@pytest.mark.parametrize("dims,perdir,body", MG_CASES)
def test_poisson_ops_two_restatements(dims, perdir, body):
    """Level construction (restrictL!, set_diag!) and every operator of the V-cycle, one at a time and as a
    whole `Vcycle!`, on a variable-coefficient hierarchy incl. semi-coarsened and odd-sized levels."""
    b = wo.make_body(**body) if body else None
    sim = wo.OracleSim(dims, (0.0,) * len(dims), perdir=perdir, body=b)
    rng = np.random.default_rng(21)
    inner = tuple(slice(1, -1) for _ in dims)
    for nm in ("x", "z"):
        a = sim.field(nm, 0)
        a[...] = 0
        a[inner] = rng.standard_normal(a[inner].shape).astype(np.float32)
    lev = _levels(sim)
    # hierarchy: coarse L from the fine L, D and iD on every level
    for l in range(1, sim.nlevels):
        c = r2.coarsen_mask(lev[l - 1]["x"], lev[l]["x"])
        a = np.zeros_like(lev[l]["L"])
        r2.restrictL(a, lev[l - 1]["L"], c, lev[l]["perdir"])
        assert np.array_equal(a, lev[l]["L"]), l
    for l in range(sim.nlevels):
        p = dict(lev[l], D=np.zeros_like(lev[l]["D"]), iD=np.zeros_like(lev[l]["iD"]))
        r2.set_diag(p)
        assert np.array_equal(p["D"], lev[l]["D"]) and np.array_equal(p["iD"], lev[l]["iD"]), l
    # single operators on level 1
    sim.lib.wlo_op_residual(sim.h, 0); r2.residual(lev[0]); _same(sim, lev, ("r", "x"), 0)
    sim.lib.wlo_op_jacobi(sim.h, 0); r2.jacobi(lev[0]); _same(sim, lev, ("r", "x", "eps"), 0)
    sim.lib.wlo_op_smooth(sim.h, 0, 0.9); r2.gauss_seidel_rb(lev[0], 4, 0.9); _same(sim, lev, ("r", "x", "eps"), 0)
    c = r2.coarsen_mask(lev[0]["x"], lev[1]["x"])
    sim.lib.wlo_op_restrict(sim.h, 0); r2.restrict(lev[1]["r"], lev[0]["r"], c); _same(sim, lev, ("r",), 1)
    a = sim.field("x", 1)
    a[inner] = rng.standard_normal(a[inner].shape).astype(np.float32)
    lev[1]["x"][...] = a.T
    sim.lib.wlo_op_prolongate(sim.h, 0); r2.prolongate(lev[0]["eps"], lev[1]["x"], c); _same(sim, lev, ("eps",), 0)
    # one whole V-cycle
    sim.lib.wlo_op_vcycle(sim.h, 0.95); r2.vcycle(lev, 1, 0.95)
    for l in range(sim.nlevels):
        _same(sim, lev, ("x", "r", "eps"), l)


# ------------------------------------------------------------------------------------------ whole steps
STEP_CASES = [((16, 12), (), True, "quick", dict(kind="sphere", center=(6, 6), radius=2.5)),
              ((16, 12), (2,), False, "vanLeer", dict(kind="sphere", center=(6, 6), radius=2.5)),
              ((12, 8, 4), (3,), True, "quick", dict(kind="sphere", center=(5, 4, 2), radius=2.0)),
              ((8, 8, 8), (1, 2, 3), False, "cds", None)]


@pytest.mark.parametrize("dims,perdir,exitBC,scheme,body", STEP_CASES)
def test_mom_step_two_restatements(dims, perdir, exitBC, scheme, body):
    """Three whole `mom_step!`s (Flow.jl:156-167: predictor, exitBC!, mom_project! with the multigrid
    `solver!` loop and its ω adaptation, corrector, CFL) by both restatements from the same constructed state:
    identical u, p, Δt history and iteration log."""
    D = len(dims)
    U = (1.0,) + (0.0,) * (D - 1)
    b = wo.make_body(**body) if body else None
    k = np.float32(2 * np.pi / dims[0])
    u0 = lambda i, x: (0.2 * np.sin(k * x[0]) * np.cos(k * x[1]) + (1.0 if i == 1 else 0.0)) * np.ones(np.broadcast(*x).shape, np.float32)
    sim = wo.OracleSim(dims, U, 4.0, nu=0.05, perdir=perdir, exitBC=exitBC, scheme=scheme, body=b, u0=u0)
    fl = {nm: sim.field(nm).copy().T for nm in ("u", "u0", "f", "p", "sigma", "V", "mu0", "mu1")}
    fl.update(dt=[np.float32(sim.dts[-1])], uBC=U, exitBC=exitBC, perdir=perdir, scheme=scheme, nu=0.05)
    lev = _levels(sim)
    lev[0].update(x=fl["p"], L=fl["mu0"], z=fl["sigma"])          # the aliases of WaterLily.jl:97
    n_log = []
    for _ in range(3):
        sim.sim_step(remeasure=False)
        r2.mom_step(fl, lev, n_log)
    assert n_log == list(sim.n), (n_log, sim.n)
    assert [float(x) for x in fl["dt"]] == [float(x) for x in sim.dts]
    assert np.array_equal(sim.field("u"), fl["u"].T)
    assert np.array_equal(sim.field("p"), fl["p"].T)


# ------------------------------------------------------------------------------------------ body side
@pytest.mark.parametrize("dims,center,R,vel,t,exitBC,perdir", [
    ((16, 12), (6.3, 5.7), 3.2, (0.0, 0.0), 0.0, False, ()),
    ((16, 12), (5.0, 6.0), 2.5, (0.5, -0.25), 1.5, True, (2,)),
    ((10, 8, 8), (4.4, 4.0, 3.6), 2.3, (0.25, 0.0, 0.125), 2.0, False, (3,))])
def test_measure_and_force_two_restatements(dims, center, R, vel, t, exitBC, perdir):
    """measure! (Body.jl:28-51: band logic, sign fix, μ₀/μ₁ kernels, the two BC!) and pressure_force
    (Metrics.jl:116-133) for a translating sphere, by both restatements."""
    D = len(dims)
    b = wo.make_body(kind="sphere", center=center, radius=R, vel=vel)
    sim = wo.OracleSim(dims, (1.0,) + (0.0,) * (D - 1), perdir=perdir, exitBC=exitBC, body=b, build_poisson=False)
    sim.lib.wlo_measure(sim.h, C.byref(sim.body), float(t), 1.0)
    fl = {nm: np.zeros_like(sim.field(nm)).T for nm in ("V", "mu0", "mu1", "sigma")}
    fl["sigma"][...] = sim.field("sigma").T * 0 + 7          # interior is overwritten, ghosts are not compared
    r2.measure_flow(fl, center, R, vel, t, 1.0, exitBC, perdir)
    inner = tuple(slice(1, -1) for _ in dims)
    assert np.array_equal(sim.field("sigma")[inner], fl["sigma"].T[inner])
    for nm in ("V", "mu0", "mu1"):
        a, c = sim.field(nm), fl[nm].T
        assert np.allclose(a, c, rtol=0, atol=2e-7), nm       # libm sin/cos may differ in the last bit
        assert (a != c).mean() < 0.01, nm
    p = sim.field("p")
    p[...] = np.random.default_rng(8).standard_normal(p.shape).astype(np.float32)
    f_ref = sim.pressure_force(t)
    f_py = r2.pressure_force(p.T, center, R, vel, t)
    assert np.allclose(f_ref, f_py, rtol=1e-6, atol=1e-6)
    u = sim.field("u")
    u[...] = np.random.default_rng(9).standard_normal(u.shape).astype(np.float32)
    sim.cfg.nu = 0.07
    # viscous_force (Metrics.jl:140-152): strain-rate tensor from the staggered differences
    sim2 = wo.OracleSim(dims, (1.0,) + (0.0,) * (D - 1), perdir=perdir, exitBC=exitBC, body=b, nu=0.07, build_poisson=False)
    sim2.field("u")[...] = u
    v_ref = sim2.viscous_force(t)
    v_py = r2.viscous_force(u.T, 0.07, center, R, vel, t)
    assert np.abs(v_ref).max() > 1e-3 and np.allclose(v_ref, v_py, rtol=1e-6, atol=1e-6), (v_ref, v_py)
    # pressure_moment / viscous_moment (Metrics.jl:168-192) about an off-centre point
    x0 = tuple(c + 0.75 for c in center)
    sim2.field("p")[...] = p
    for visc in (False, True):
        m_ref = sim2.moment(x0, visc, t)
        m_py = r2.force_moment(p.T, u.T, 0.07, x0, center, R, vel, t, visc)
        assert np.abs(m_ref).max() > 1e-3 and np.allclose(m_ref, m_py, rtol=1e-6, atol=1e-5), (visc, m_ref, m_py)


def test_meanflow_two_restatements():
    """MeanFlow update! (Metrics.jl:239-249): ε schedule and the three running averages, incl. the u⊗u tensor."""
    sim = wo.OracleSim((16, 12), (1.0, 0.0), 4.0, nu=0.05, body=wo.make_body(kind="sphere", center=(6, 6), radius=2.5))
    mf = wo.MeanFlow(sim, uu_stats=True)
    P, U, UU = (np.zeros_like(a).T for a in (mf.P, mf.U, mf.UU))
    t = [np.float32(sim.time())]
    for n in range(4):
        sim.sim_step(remeasure=False)
        mf.update()
        dt = np.float32(sim.time()) - t[-1]
        eps = np.float32(1) if len(t) == 1 else dt / (dt + (t[-1] - t[0]) + np.finfo(np.float32).eps)
        r2.mean_update(P, U, UU, sim.field("p").T, sim.field("u").T, eps)
        t.append(t[-1] + dt)
    assert np.array_equal(mf.P, P.T) and np.array_equal(mf.U, U.T) and np.array_equal(mf.UU, UU.T)
    assert np.abs(mf.uu()).max() > 0 and [float(x) for x in mf.t] == [float(x) for x in t]


# ------------------------------------------------------------------------------------------ Function uBC / g
@pytest.mark.parametrize("shape,perdir,saveexit", [((7, 6), (), False), ((7, 6), (2,), True), ((5, 6, 4), (3,), False),
                                                   ((5, 6, 4), (), True)])
def test_bc_and_accelerate_functions_two_restatements(shape, perdir, saveexit):
    """BC!(a,uBC::Function,...) (core.jl:201-219) and accelerate!(r,t,g,U) (Flow.jl:69-73) by both restatements."""
    from test_oracle_fnbc import accel_fn, bc_fn
    D = len(shape)
    rng = np.random.default_rng(5)
    ub = lambda i, x, t: np.float32(0.3 * i) + np.float32(0.1) * x[i % D] * x[(i + 1) % D] + t * x[0]
    a = rng.standard_normal((D,) + shape[::-1]).astype(np.float32)
    b = np.ascontiguousarray(a.T).copy()
    bc_fn(a, ub, saveexit=saveexit, perdir=perdir, t=0.75)
    r2.BC_fn(b, ub, saveexit, perdir, 0.75)
    assert np.array_equal(a, b.T)
    g = lambda i, x, t: np.float32(i) * x[0] - t
    dU = lambda i, x, t: x[D - 1] * np.float32(0.5)
    for gg, dd in ((g, None), (None, dU), (g, dU)):
        a = rng.standard_normal((D,) + shape[::-1]).astype(np.float32)
        b = np.ascontiguousarray(a.T).copy()
        accel_fn(a, 0.5, gg, dd)
        r2.accelerate_fn(b, 0.5, gg, dd)
        assert np.array_equal(a, b.T)
