"""Pins the oracle's boundary conditions against the reference's own assertions
(test/test_core.jl:19-69).  Julia a[i,j,c] == numpy a[c-1,j-1,i-1]."""
import ctypes as C

import numpy as np
import pytest

from oracle import wl_oracle as wo


def _BC(lib, a, U, saveexit=False, perdir=()):
    D = a.shape[0]
    ng = (C.c_int * 3)(*(list(a.shape[1:][::-1]) + [1] * (3 - D)))
    per = (C.c_int * 3)(*[1 if d + 1 in perdir else 0 for d in range(3)])
    Ua = (lib.real * 3)(*(list(U) + [0.0] * (3 - D)))
    lib.wlo_BC(a.ctypes.data, D, ng, Ua, int(saveexit), per)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_BC_2d(dtype):
    lib = wo.load(dtype)
    rng = np.random.default_rng(0)
    U = (1.0, 0.5)
    u = rng.random((2, 6, 6)).astype(dtype)   # Julia (6,6,2)
    _BC(lib, u, U)
    ux, uy = u[0].T, u[1].T                   # ux[i,j] Julia order
    # test_core.jl:24-27
    assert np.all(ux[0, :] == U[0]) and np.all(ux[1, :] == U[0]) and np.all(ux[-1, :] == U[0])
    assert np.all(ux[2:-1, 0] == ux[2:-1, 1]) and np.all(ux[2:-1, -1] == ux[2:-1, -2])
    assert np.all(uy[:, 0] == U[1]) and np.all(uy[:, 1] == U[1]) and np.all(uy[:, -1] == U[1])
    assert np.all(uy[0, 2:-1] == uy[1, 2:-1]) and np.all(uy[-1, 2:-1] == uy[-2, 2:-1])
    # saveexit test_core.jl:29-31
    ux[-1, :] = 3
    _BC(lib, u, U, True)
    assert np.all(u[0].T[-1, :] == 3)
    # exitBC! conservative exit test_core.jl:33-34
    ng = (C.c_int * 3)(6, 6, 1)
    lib.wlo_exitBC(u.ctypes.data, u.ctypes.data, 2, ng, 0.0)
    assert np.all(u[0].T[-1, 1:-1] == U[0])
    # periodic in y + saveexit test_core.jl:47-48
    _BC(lib, u, U, True, (2,))
    ux = u[0].T
    assert np.all(ux[:, 0:2] == ux[:, -2:])
    # perBC! test_core.jl:49-50
    s = rng.random((6, 6)).astype(dtype)
    per = (C.c_int * 3)(1, 1, 0)
    lib.wlo_perBC(s.ctypes.data, 2, ng, per)
    sj = s.T
    assert np.all(sj[0, 1:-1] == sj[-2, 1:-1]) and np.all(sj[1:-1, 0] == sj[1:-1, -2])
    # x-periodic, saveexit has no effect test_core.jl:52-55
    u = rng.random((2, 6, 6)).astype(dtype)
    _BC(lib, u, U, True, (1,))
    ux, uy = u[0].T, u[1].T
    assert np.all(ux[0:2, :] == ux[-2:, :]) and np.all(uy[0:2, :] == uy[-2:, :])
    assert np.all(uy[:, 0] == U[1]) and np.all(uy[:, 1] == U[1]) and np.all(uy[:, -1] == U[1])


def test_BC_3d_planes():
    """Constant-tuple analogue of test_core.jl:62-69: Dirichlet planes 1,2,end for the normal comp."""
    lib = wo.load(np.float32)
    U = (0.25, -0.5, 2.0)
    u = np.random.default_rng(1).random((3, 8, 8, 8)).astype(np.float32)
    _BC(lib, u, U)
    for c in range(3):
        a = np.moveaxis(u[c], 2 - c, 0)       # bring Julia dim c+1 first
        assert np.all(a[0] == np.float32(U[c])) and np.all(a[1] == np.float32(U[c])) and np.all(a[-1] == np.float32(U[c]))


def test_index_ranges():
    """inside / inside_u / slice extents (core.jl:47-59,188) via their observable effect:
    conv_diff! writes r on 2:N (other dims) and never on the lower ghost."""
    sim = wo.OracleSim((8, 8), (1.0, 0.0), nu=0.1)
    rng = np.random.default_rng(3)
    sim.field("u")[...] = rng.standard_normal(sim.field("u").shape).astype(np.float32)
    sim.lib.wlo_op_conv_diff(sim.h, 0)
    f = sim.field("f")
    for c in range(2):
        fj = f[c].T
        assert np.all(fj[0, :] == 0) and np.all(fj[:, 0] == 0)      # lower ghosts untouched
        assert np.any(fj[-1, 1:-1] != 0) and np.any(fj[1:-1, -1] != 0)  # upper ghost rows ARE written
