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
