"""Parity at the sizes that are TIMED: the CUDA path through the C ABI against the CPU
oracle on BASELINE's own grids, 20 steps, `u`, `p`, Δt history, `pois.n`, `pressure_force`; plus the host-measured
(closure) body route -- `wl_set_body_coeffs`, `wl_set_force_weights` -- and time-dependent Function boundary data.
Tolerances as in test_parity_gpu.py (north_star: 1e-5 relative, identical iteration counts)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import wl_oracle as wo
from test_parity_gpu import check_state, pair, step_both, tgv3d


@pytest.fixture(scope="module")
def wl():
    import __graft_entry__ as ge
    ge.build()
    import waterlily_jl_b200 as wl
    return wl


def tgv_pair(wl, dims):
    N = dims[0]
    k = np.float32(2 * np.pi / N)
    kw = dict(perdir=(1, 2, 3), nu=float(1 / (k * 1600)), U=1.0, u0=lambda i, x: tgv3d(i, x, k))
    return pair(wl, dims, (0.0, 0.0, 0.0), float(N), **kw)


def test_config_B_tgv128_20_steps_vs_oracle(wl):
    """BASELINE configs[1] at FULL size: 3D TGV 128³ periodic Float32, 20 steps."""
    sim, ref = tgv_pair(wl, (128, 128, 128))
    step_both(sim, ref, 20)
    r = check_state(sim, ref)
    print("config B 128^3 x20:", r, "n =", sim.pois.n[-4:])


def test_config_C_sphere512_20_steps_vs_oracle(wl):
    """BASELINE configs[2] at FULL size: sphere Re=3700 on 512x256x256, 20 steps, incl. pressure_force (§8d row C).
    This is the grid bench.py times by default: wave-aware march length, 64-plane chunks, ~10^5-block reductions."""
    m = 256
    R = m / 8
    body = dict(kind="sphere", center=(m / 2 - 1, m / 2 - 1, m / 2 - 1), radius=R)
    sim, ref = pair(wl, (2 * m, m, m), (1.0, 0.0, 0.0), 2 * R, body=body, nu=2 * R / 3700)
    step_both(sim, ref, 20, remeasure=False)
    r = check_state(sim, ref)
    f_gpu, f_ref = sim.pressure_force(), ref.pressure_force()
    assert np.allclose(f_gpu, f_ref, rtol=1e-5, atol=1e-5 * np.abs(f_ref).max()), (f_gpu, f_ref)
    print("config C 512x256x256 x20:", r, "n =", sim.pois.n[-4:], "force", f_gpu)


def test_config_E_slice_1024x1024x16_vs_oracle(wl):
    """A reduced-z slice of BASELINE configs[4] (TGV 1024³): the long-row paths (1026-float rows, 32 column tiles,
    multigrid levels semi-coarsened once z stops dividing), 3 steps."""
    N = 1024
    k = np.float32(2 * np.pi / N)
    kw = dict(perdir=(1, 2, 3), nu=float(1 / (k * 1600)), U=1.0, u0=lambda i, x: tgv3d(i, x, k))
    sim, ref = pair(wl, (N, N, 16), (0.0, 0.0, 0.0), float(N), **kw)
    step_both(sim, ref, 3)
    check_state(sim, ref)


# ---------------------------------------------------------------- host-measured (closure) bodies
def ellipsoid(c, r):
    """Pseudo-sdf of an ellipsoid, sdf(ξ) = (|ξ/r| - 1)·min(r), with its analytic gradient; `measure(::AutoBody)`
    divides both by |∇sdf| (src/AutoBody.jl:29-37).  An off-library shape: no wl_body kind describes it."""
    c = [np.float32(v) for v in c]
    r = [np.float32(v) for v in r]
    rm = np.float32(min(r))

    def sdf(x, t):
        q = np.sqrt(sum(((a - ci) / ri) ** 2 for a, ci, ri in zip(x, c, r)))
        return ((q - np.float32(1)) * rm).astype(np.float32)

    def grad(x, t):
        q = np.sqrt(sum(((a - ci) / ri) ** 2 for a, ci, ri in zip(x, c, r)))
        with np.errstate(invalid="ignore", divide="ignore"):
            return [(rm * (a - ci) / (ri * ri) / q).astype(np.float32) for a, ci, ri in zip(x, c, r)]
    return sdf, grad


def measure_loop(ref, sdf, grad, eps=1.0):
    """measure! (src/Body.jl:28-48) cell by cell, written from the Julia source with the oracle's own μ₀/μ₁ kernels:
    the independent check of the mirror's vectorised `host_measure`."""
    T = np.float32
    D = ref.D
    n = [d + 2 for d in ref.dims]
    mu0 = np.ones(ref.field("mu0").shape, T)
    mu1 = np.zeros(ref.field("mu1").shape, T)
    V = np.zeros(ref.field("V").shape, T)
    sig = np.zeros(ref.field("sigma").shape, T)
    d2 = T((2 + eps) ** 2)

    def meas(x, fastd2):                                # measure(::AutoBody) src/AutoBody.jl:29-37, identity map
        xa = [np.array([v], T) for v in x]
        d = T(sdf(xa, 0)[0])
        if d * d > fastd2:
            return d, np.zeros(D, T)
        g = np.array([gg[0] for gg in grad(xa, 0)], T)
        if np.isnan(g).any():
            return d, np.zeros(D, T)
        m = np.sqrt(np.sum(g * g), dtype=T)
        return T(d / m), (g / m).astype(T)
    for I in np.ndindex(*[v - 2 for v in n]):
        I = tuple(v + 1 for v in I)                      # 0-based interior index, x fastest LAST in this tuple order
        a = I[::-1]                                      # array index (z, y, x)
        xc = [T(I[d]) - T(0.5) for d in range(D)]
        d = T(sdf([np.array([v], T) for v in xc], 0)[0])  # measure_sdf!: the raw sdf
        sig[a] = d
        if d * d < d2:
            for i in range(D):
                xi = list(xc)
                xi[i] = xi[i] - T(0.5)
                di, ni = meas(xi, d2)
                di = di if abs(di) <= 0.5 else T(np.copysign(di, d))
                mu0[(i,) + a] = ref.lib.wlo_mu0(di, eps)
                for j in range(D):
                    mu1[(j, i) + a] = T(ref.lib.wlo_mu1(di, eps)) * ni[j]
        elif d < 0:
            for i in range(D):
                mu0[(i,) + a] = 0
    return mu0, mu1, V, sig


def test_closure_body_coeffs_and_force_weights(wl):
    """The route of an arbitrary `AutoBody(sdf,map)` closure (src/Body.jl:28-51, src/AutoBody.jl:29-37,
    src/Metrics.jl:110-133): host measure! -> wl_set_body_coeffs -> device steps; host nds -> wl_set_force_weights ->
    wl_pressure_force_weights.  An ellipsoid (no library kind), CUDA vs oracle stepping on the same coefficients."""
    import ctypes as C
    dims = (32, 24, 16)
    sdf, grad = ellipsoid((12.3, 11.6, 8.2), (6.0, 4.0, 3.0))
    body = wl.ClosureBody.from_sdf(sdf, grad)
    kw = dict(nu=0.05, perdir=(3,), exitBC=True)
    sim = wl.Simulation(dims, (1.0, 0.0, 0.0), 8.0, body=body, **kw)
    ref = wo.OracleSim(dims, (1.0, 0.0, 0.0), 8.0, build_poisson=False, **kw)
    # (1) the mirror's vectorised host measure! against a cell-by-cell restatement
    mu0, mu1, V, sig = wl.host_measure(body, dims, 0.0, 1.0)
    l0, l1, lV, ls = measure_loop(ref, sdf, grad)
    assert float(np.abs(ls - sig).max()) <= 2e-6 and float(np.abs(l0 - mu0).max()) <= 2e-6
    assert float(np.abs(l1 - mu1).max()) <= 2e-6 and np.count_nonzero(mu1) > 1000
    # (2) the oracle steps on the SAME coefficients: fill its fields, then the two BC! of Body.jl:49-50, then the hierarchy
    for nm, a in (("mu0", mu0), ("mu1", mu1), ("V", V), ("sigma", sig)):
        ref.field(nm)[...] = a
    ng = (C.c_int * 3)(*[d + 2 for d in dims])
    per = (C.c_int * 3)(0, 0, 1)
    zero = (C.c_float * 3)(0, 0, 0)
    ref.lib.wlo_BC(ref.field("mu0").ctypes.data, 3, ng, zero, 0, per)
    ref.lib.wlo_BC(ref.field("V").ctypes.data, 3, ng, zero, 1, per)
    assert ref.lib.wlo_build_poisson(ref.h) > 0
    assert np.array_equal(sim.flow.mu0, ref.field("mu0")) and np.array_equal(sim.flow.mu1, ref.field("mu1"))
    assert np.array_equal(sim.pois.levels[1].L, ref.field("L", 1))
    step_both(sim, ref, 6, remeasure=False)
    check_state(sim, ref)
    # (3) pressure_force: Σ p·nds over the interior, Float32 products, Float64 sum (src/Metrics.jl:126-133)
    f_gpu = sim.pressure_force()
    p = ref.field("p")
    idx, w = wl.host_nds(body, dims, sim.time())
    df = (p.ravel()[idx][:, None] * w).astype(np.float32)
    f_ref = df.astype(np.float64).sum(axis=0)
    assert np.abs(f_ref).max() > 1e-2
    assert np.allclose(f_gpu, f_ref, rtol=1e-6, atol=1e-6 * np.abs(f_ref).max()), (f_gpu, f_ref)
    # the weights are nds itself: a library sphere measured both ways gives the same force
    sph = dict(kind="sphere", center=(12.0, 11.0, 8.0), radius=4.0)
    a = wl.Simulation(dims, (1.0, 0.0, 0.0), 8.0, body=wl.AutoBody(**sph), **kw)
    c = np.float32(sph["center"])
    rr = lambda x: np.sqrt(sum((xx - cc) ** 2 for xx, cc in zip(x, c)))
    b = wl.Simulation(dims, (1.0, 0.0, 0.0), 8.0, **kw,
                      body=wl.ClosureBody.from_sdf(lambda x, t: (rr(x) - np.float32(4.0)).astype(np.float32),
                                                   lambda x, t: [((xx - cc) / rr(x)).astype(np.float32) for xx, cc in zip(x, c)]))
    for _ in range(3):
        a.sim_step(remeasure=False)
        b.sim_step(remeasure=False)
    fa, fb = a.pressure_force(), b.pressure_force()
    assert np.allclose(fa, fb, rtol=2e-4, atol=2e-4 * np.abs(fa).max()), (fa, fb)


def test_moving_closure_body_remeasure(wl):
    """A translating closure body remeasured every step (host measure! + upload + update! per step) follows the
    library's device `measure!` of the same sphere: V, μ₀, μ₁ agree to rounding, so do the fields after 3 steps."""
    dims = (32, 24, 16)
    c0, Vb, R = np.float32([10.3, 11.7, 8.2]), np.float32([0.5, 0.0, 0.0]), np.float32(4.1)   # no cell sits on the band edge
    lib_body = wl.AutoBody("sphere", tuple(c0), float(R), vel=tuple(Vb))

    def measure(x, t, fastd2=np.inf):
        c = [cc + vv * np.float32(t) for cc, vv in zip(c0, Vb)]
        q = np.sqrt(sum((xx - cc) ** 2 for xx, cc in zip(x, c)))
        d = (q - R).astype(np.float32)
        near = d * d <= np.float32(fastd2)                          # measure(::AutoBody) returns n = V = 0 beyond fastd²
        return d, [np.where(near, (xx - cc) / q, 0).astype(np.float32) for xx, cc in zip(x, c)], [np.where(near, vv, 0).astype(np.float32) for vv in Vb]
    a = wl.Simulation(dims, (0.0, 0.0, 0.0), 8.0, body=lib_body, nu=0.05, U=1.0)
    b = wl.Simulation(dims, (0.0, 0.0, 0.0), 8.0, body=wl.ClosureBody(measure), nu=0.05, U=1.0)
    for _ in range(3):
        a.sim_step()
        b.sim_step()
    assert a.pois.n == b.pois.n
    assert float(np.abs(a.flow.V - b.flow.V).max()) <= 1e-6
    assert float(np.abs(a.flow.u - b.flow.u).max()) <= 2e-5


# ---------------------------------------------------------------- time-dependent Function uBC through the constructor
def test_time_dependent_function_ubc(wl):
    """`Simulation(dims, uBC::Function, L; U, ...)`: the boundary tables are re-sampled at t₁ = sum(Δt) before every
    mom_step (src/Flow.jl:194,209,231 use t₁ in all four BC!), ∂ₜuBC enters accelerate! (Flow.jl:72-73); CUDA vs the
    oracle's function-valued path over 12 steps, 2D and 3D."""
    T = np.float32
    for dims, perdir in (((32, 24), ()), ((16, 12, 8), (3,))):
        D = len(dims)
        ubc = lambda i, x, t: (T(1) + T(0.2) * t + T(0.01) * x[1]) if i == 1 else T(0) * x[0] * t
        dub = lambda i, x, t: (T(0.2) + T(0) * x[1] * t) if i == 1 else T(0) * x[0] * t
        kw = dict(nu=0.02, U=1.0, perdir=perdir, dt=0.1)
        sim = wl.Simulation(dims, ubc, 8.0, dUBCdt=dub, **kw)
        ref = wo.OracleSim(dims, lambda i, x, t: float(ubc(i, x, t)), 8.0, dUdt=lambda i, x, t: float(dub(i, x, t)), **kw)
        assert np.array_equal(sim.flow.u, ref.field("u"))
        for _ in range(12):
            sim.sim_step()
            ref.sim_step()
        assert sim.pois.n == ref.n, (sim.pois.n, ref.n)
        assert np.allclose(sim.flow.Δt, ref.dts, rtol=2e-6)
        assert float(np.abs(sim.flow.u - ref.field("u")).max()) <= 2e-6 * 4
        assert float(sim.flow.u[0].max()) > 1.2          # the inflow did grow with t
        with pytest.raises(wl.WaterLilyB200Error):
            sim.sim_step(uBC_t=lambda t: (1.0,) * D)
