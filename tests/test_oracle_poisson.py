"""Pins the oracle's Poisson / MultiLevelPoisson against the reference's known answers
(test/test_poisson.jl).  Julia a[i,j] == numpy a[j-1,i-1]."""
import ctypes as C

import numpy as np
import pytest

from oracle import wl_oracle as wo


def L2(a):   # WaterLily.L₂: sum of squares over inside(a)  (Poisson.jl:188)
    sl = tuple(slice(1, -1) for _ in a.shape)
    return float(np.sum(a[sl].astype(np.float64) ** 2))


def poisson_setup(N, multilevel, dtype=np.float32):
    """`Poisson_setup` test/test_poisson.jl:1-12 (N includes ghosts)."""
    D = len(N)
    sim = wo.OracleSim(tuple(n - 2 for n in N), (0.0,) * D, dtype=dtype, build_poisson=False)
    nl = sim.lib.wlo_build_poisson(sim.h)
    if multilevel and nl < 0:
        raise AssertionError("MultiLevelPoisson requires size=a2ⁿ, where n>2")
    x = sim.field("p")
    soln = np.zeros_like(x)
    soln[...] = (np.arange(N[0], dtype=dtype) - 1).reshape((1,) * (D - 1) + (N[0],))
    sim.lib.wlo_mult(sim.h, soln.ctypes.data)
    n = sim.lib.wlo_solver(sim.h) if multilevel else sim.lib.wlo_poisson_solver(sim.h, 1000)
    x -= x[(1,) * D]
    return L2(x - soln) / L2(soln), sim, n


def test_poisson_diag_golden():
    # test_poisson.jl:16-19
    err, sim, n = poisson_setup((5, 5), False)
    D = sim.field("D", 0).T
    iD = sim.field("iD", 0).T
    gold = np.array([[0, 0, 0, 0, 0], [0, -2, -3, -2, 0], [0, -3, -4, -3, 0], [0, -2, -3, -2, 0], [0, 0, 0, 0, 0]], np.float32)
    assert np.array_equal(D, gold)
    with np.errstate(divide="ignore"):
        assert np.allclose(iD, np.where(gold == 0, 0, 1 / gold))
    assert err < 1e-5


def test_poisson_pcg():
    # test_poisson.jl:20-26
    err, sim, n = poisson_setup((66, 66), False)
    assert err < 5e-6 and n < 340
    assert sim.lib.wlo_op_Linf(sim.h, 0) < 2e-3
    err, sim, n = poisson_setup((18, 18, 18), False)
    assert err < 1e-6 and n < 40


def test_mlp_size_assertion():
    # test_poisson.jl:42
    with pytest.raises(AssertionError):
        poisson_setup((17, 83), True)


def test_level_shapes():
    """coarsen_mask / divisible / level sizes (MultiLevelPoisson.jl:29,36,52-54)."""
    def shapes(N):
        sim = wo.OracleSim(N, (0.0,) * len(N))
        return [sim.field("x", l).shape[::-1] for l in range(sim.nlevels)]
    assert shapes((96, 64)) == [(98, 66), (50, 34), (26, 18), (14, 10), (8, 6), (5, 4)]
    assert shapes((16, 16, 16)) == [(18,) * 3, (10,) * 3, (6,) * 3, (4,) * 3]
    # semi-coarsening masks test_poisson.jl:45-47: (18,18,6)->all, (18,18,4)->no z, (18,17,6)->no y
    assert shapes((16, 16, 4))[1] == (10, 10, 4)
    assert shapes((16, 16, 2))[1] == (10, 10, 4)
    assert shapes((16, 15, 4))[1] == (10, 17, 4)


def test_mlp_coarse_diag_golden():
    # test_poisson.jl:54-60
    err, sim, n = poisson_setup((10, 10), True)
    assert np.array_equal(sim.field("D", 2).T, np.array([[0, 0, 0, 0], [0, -2, -2, 0], [0, -2, -2, 0], [0, 0, 0, 0]], np.float32))
    assert err < 1e-5
    L = sim.field("L", 0)
    L[0].T[4:6, :] = 0          # pois.levels[1].L[5:6,:,1] .= 0
    sim.lib.wlo_update_poisson(sim.h)
    assert np.array_equal(sim.field("D", 2).T, np.array([[0, 0, 0, 0], [0, -1, -1, 0], [0, -1, -1, 0], [0, 0, 0, 0]], np.float32))


def test_mlp_iteration_counts():
    # test_poisson.jl:62-70
    err, sim, n = poisson_setup((66, 66), True)
    assert err < 1e-6 and n <= 4
    assert sim.lib.wlo_op_Linf(sim.h, 0) < 2e-3
    err, sim, n = poisson_setup((18, 18, 18), True)
    assert err < 1e-6 and n <= 3


def test_semicoarsening_sims():
    # test_poisson.jl:73-82
    H = 16; R = H // 4
    sim = wo.OracleSim((8 * H, H), (1, 0), R, nu=R / 100, body=wo.make_body("circle", (4 * H, H // 2), R))
    for _ in range(4):
        sim.sim_step(remeasure=False)
    assert all(n <= 10 for n in sim.n), sim.n
    H = 8; R = H // 4
    sim = wo.OracleSim((8 * H, H, H), (1, 0, 0), R, nu=R / 100, body=wo.make_body("sphere", (4 * H, H // 2, H // 2), R))
    for _ in range(4):
        sim.sim_step(remeasure=False)
    assert all(n <= 12 for n in sim.n), sim.n
    assert not np.isnan(sim.field("u")).any()
