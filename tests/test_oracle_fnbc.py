"""Function-valued `uBC(i,x,t)` and `g(i,x,t)` in the oracle (the oracle comes first), pinned on the
reference's known answers: test/test_core.jl:37-45,56-69 (BC! with a function), test/test_flow.jl:53-75
(accelerate!), :120-131 (constant jerk through g), :134-140 (boundary-layer profile), :142-159 (rotating frame),
:161-173 (circle in accelerating flow)."""
import ctypes as C

import numpy as np
import pytest

from oracle import wl_oracle as wo


def _cb(lib, fn, D, T):
    return lib.FN(lambda i, x, t: float(fn(i, [T(x[k]) for k in range(D)], T(t))))


def _ng(shape):
    dims = list(shape[::-1])
    return (C.c_int * 3)(*(dims + [1] * (3 - len(dims))))


def bc_fn(a, fn, saveexit=False, perdir=(), t=0.0):
    """BC!(a, uBC::Function, saveexit, perdir, t) on a numpy array shaped (D, [nz,] ny, nx)."""
    lib = wo.load(a.dtype.type)
    D = a.shape[0]
    per = (C.c_int * 3)(*[1 if (d + 1) in perdir else 0 for d in range(3)])
    lib.wlo_BC_fn(a.ctypes.data, D, _ng(a.shape[1:]), _cb(lib, fn, D, a.dtype.type), int(saveexit), per, float(t))


def accel_fn(r, t, g=None, dUdt=None):
    lib = wo.load(r.dtype.type)
    D = r.shape[0]
    f = lambda h: lib.FN() if h is None else _cb(lib, h, D, r.dtype.type)
    lib.wlo_accelerate_fn(r.ctypes.data, D, _ng(r.shape[1:]), float(t), f(g), f(dUdt))


def test_bc_function_known_answers():
    """test/test_core.jl:37-45 and :56-69.  numpy index [c, j, i] is Julia [i+1, j+1, c+1]."""
    rng = np.random.default_rng(0)
    T = np.float64
    u, v = rng.random((2, 6, 6)), rng.random((2, 6, 6))
    lib = wo.load(T)
    per0 = (C.c_int * 3)(0, 0, 0)
    U = (lib.real * 3)(1.0, 0.5, 0.0)
    lib.wlo_BC(u.ctypes.data, 2, _ng(u.shape[1:]), U, 0, per0)
    Ubc = lambda i, x, t: 1.0 if i == 1 else 0.5
    bc_fn(v, Ubc)
    assert np.array_equal(v[0][:, [0, 1, -1]], u[0][:, [0, 1, -1]]) and np.array_equal(v[1][[0, 1, -1], :], u[1][[0, 1, -1], :])
    v[0][:, -1] = 3
    bc_fn(v, Ubc, saveexit=True)
    assert np.all(v[0][:, -1] == 3)
    # non-uniform
    v[...] = 0
    bc_fn(v, lambda i, x, t: x[1] if i == 1 else x[0])
    assert np.allclose(v[0][1:-1, 0], v[0][1:-1, -1]) and np.allclose(v[1][0, 1:-1], v[1][-1, 1:-1])
    u3 = np.zeros((3, 8, 8, 8))
    bc_fn(u3, lambda i, x, t: np.cos(2 * np.pi * x[0] / 8) if i == 1 else np.sin(2 * np.pi * x[1] / 8) if i == 2 else np.tan(np.pi * x[2] / 16))
    pi = np.pi
    assert np.allclose(u3[0][:, :, 0], np.cos(-pi / 4)) and np.allclose(u3[0][:, :, 1], 1.0) and np.allclose(u3[0][:, :, -1], np.cos(6 * pi / 4), atol=1e-12)
    assert np.allclose(u3[1][:, 0, :], np.sin(-pi / 4)) and np.allclose(u3[1][:, 1, :], 0.0) and np.allclose(u3[1][:, -1, :], np.sin(6 * pi / 4))
    assert np.allclose(u3[2][0], np.tan(-pi / 16)) and np.allclose(u3[2][1], 0.0) and np.all(u3[2][-1] - np.tan(6 * pi / 16) < 1e-6)


def test_accelerate_known_answers():
    """test/test_flow.jl:53-75."""
    a = np.zeros((2, 4, 4))
    accel_fn(a, 1.0, None, None)
    assert np.all(a == 0)
    accel_fn(a, 1.0, lambda i, x, t: t if i == 1 else 2 * t)
    assert np.all(a[0] == 1) and np.all(a[1] == 2)
    accel_fn(a, 1.0, None, lambda i, x, t: -1.0 if i == 1 else -2.0)          # ∂ₜ of U = -t, -2t
    assert np.all(a == 0)
    accel_fn(a, 1.0, lambda i, x, t: t if i == 1 else 2 * t, lambda i, x, t: -1.0 if i == 1 else -2.0)
    assert np.all(a == 0)
    b = np.zeros((2, 4, 4))
    accel_fn(b, 0.0, lambda i, x, t: 1.0)
    assert np.all(b == 1)
    accel_fn(b, 1.0, lambda i, x, t: 0.0, lambda i, x, t: 1.0)                # ∂ₜ of U = t
    assert np.all(b == 2)
    a[...] = 0
    b[...] = 1
    accel_fn(a, 0.0, None, lambda i, x, t: x[i - 1] + 1.0)                    # ∂ₜ of U = t(xᵢ+1)
    accel_fn(b, 0.0, lambda i, x, t: x[i - 1])
    assert np.array_equal(a, b)
    accel_fn(b, 1.0, lambda i, x, t: x[i - 1] + 1.0)
    accel_fn(a, 1.0, None, lambda i, x, t: x[i - 1] + 1.0)
    assert np.array_equal(a, b)


def L2(a):
    return float(np.sum(np.asarray(a, np.float64)[..., 1:-1, 1:-1] ** 2))


def test_constant_jerk_through_g():
    """test/test_flow.jl:120-131 with use_g=true (Float64, periodic in x): uₓ = U + ½·jerk·t²."""
    N, jerk = 8, 4.0
    sim = wo.OracleSim((N, N), (np.sqrt(N), 0.0), N, nu=0.001, dt=0.001, perdir=(1,), dtype=np.float64,
                       g=lambda i, x, t: t * jerk if i == 1 else 0.0)
    sim.sim_step(1.0)
    u = sim.field("u")
    uF = np.sqrt(N) + 0.5 * jerk * sim.time() ** 2
    assert L2(u[0] - uF) < 1e-4 and L2(u[1]) < 1e-4


def test_boundary_layer_profile_is_kept():
    """test/test_flow.jl:134-140: parabolic inflow as a space-varying uBC; u[1,:,1] ≈ u[end,:,1] after sim_step!(sim,10)."""
    L = 32
    prof = lambda i, x, t: np.float32(4.0 * (((x[1] + 0.5) / (2 * L)) - ((x[1] + 0.5) / (2 * L)) ** 2)) if i == 1 else 0.0
    sim = wo.OracleSim((L, L), prof, L, nu=0.001, U=1)
    assert np.abs(sim.field("u")[0]).max() > 0.5
    sim.sim_step(10)
    u = sim.field("u")
    assert np.allclose(u[0][:, 0], u[0][:, -1], rtol=np.sqrt(np.finfo(np.float32).eps), atol=0)


def test_rotating_reference_frame():
    """test/test_flow.jl:142-159 (Float64): solid-body rotation seen from the rotating frame needs no pressure."""
    L = 4
    N, w = 2 * L, 1.0 / L
    x0 = (float(L), float(L))

    def vel(i, x, t):
        s, c = np.sin(w * t), np.cos(w * t)
        y = [w * (x[0] - x0[0]), w * (x[1] - x0[1])]
        return s * y[0] + c * y[1] if i == 1 else -c * y[0] + s * y[1]

    def dvel(i, x, t):                                     # ∂ₜ velocity
        s, c = np.sin(w * t), np.cos(w * t)
        y = [w * (x[0] - x0[0]), w * (x[1] - x0[1])]
        return w * (c * y[0] - s * y[1]) if i == 1 else w * (s * y[0] + c * y[1])

    g = lambda i, x, t: (2 * w * vel(2, x, t) if i == 1 else -2 * w * vel(1, x, t)) + w ** 2 * (x[i - 1] - x0[i - 1])
    sim = wo.OracleSim((N, N), vel, N, U=1, dtype=np.float64, g=g, dUdt=dvel)
    sim.sim_step()
    assert L2(sim.field("p")) < 3e-3


def test_function_path_equals_uniform_path():
    """The accelerating-circle case of test/test_flow.jl:161-173 (already pinned at full size through the uniform
    time-dependent path in test_oracle_flow.py) run through the Function path at a small size: uBC(i,x,t) = (t, 0)
    must reproduce the uniform path bit for bit, Δt history and Poisson iteration counts included."""
    R, H = 8, 4
    mk = lambda: wo.make_body("circle", center=(H * R, H * R), radius=R)
    a = wo.OracleSim((2 * H * R, 2 * H * R), lambda i, x, t: t if i == 1 else 0.0, R, U=1, body=mk(),
                     dUdt=lambda i, x, t: 1.0 if i == 1 else 0.0)
    b = wo.OracleSim((2 * H * R, 2 * H * R), (0.0, 0.0), R, U=1, body=mk())
    for _ in range(4):
        a.sim_step()
        b.sim_step(uBC_t=lambda t: (t, 0.0), dUdt=lambda t: (1.0, 0.0))
    assert a.n == b.n and [float(x) for x in a.dts] == [float(x) for x in b.dts]
    assert np.array_equal(a.field("u"), b.field("u")) and np.array_equal(a.field("p"), b.field("p"))
    assert np.abs(a.field("u")).max() > 0.5
