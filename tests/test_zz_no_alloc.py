"""wl_cfl_noalloc (include/wl_b200.h): the allocation-free CFL entry point, against the existing wl_cfl, through the C
ABI, two calls in a row (the second finds the first one's leftovers in the shared static scratch array).  Collected
last on purpose (pytest -x): a failure here must not hide any other test.  torch only provides device memory."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wl():
    import __graft_entry__ as ge
    ge.build()
    import waterlily_jl_b200 as wl
    return wl


def dt_noalloc(wl, torch, u, nu, stream=None):
    """u: numpy (D, [nz,] ny, nx) as the mirror holds it == the reference's (nx, ny[, nz], D) column-major."""
    D = u.shape[0]
    n = (C.c_int32 * 3)(*(list(u.shape[1:][::-1]) + [1] * (3 - D)))
    ud = torch.from_numpy(np.ascontiguousarray(u)).cuda()
    out = torch.full((1,), -1.0, dtype=torch.float32, device="cuda")
    rc = wl.load_library().wl_cfl_noalloc(ud.data_ptr(), D, C.byref(n), float(nu), out.data_ptr(), stream)
    assert rc == 0, wl.load_library().wl_last_error(None).decode()
    torch.cuda.synchronize()
    return float(out.item())


@pytest.mark.parametrize("dims,nu", [((48, 40, 24), 0.01), ((96, 64), 0.16)])
def test_cfl_noalloc_matches_wl_cfl_twice(wl, dims, nu):
    import torch
    D = len(dims)
    sim = wl.Simulation(dims, (1.0,) + (0.0,) * (D - 1), 8.0, nu=nu)       # fresh sim: σ's ghost cells are all zero
    rng = np.random.default_rng(11)
    for call, scale in enumerate((1.0, 0.35)):                              # second call: SMALLER velocities, so a stale
        u = (scale * rng.uniform(-1, 1, sim.field_shape("u"))).astype(np.float32)   # row maximum would change the answer
        sim.upload("u", u)
        want = sim.cfl()                                                    # existing entry point: wl_cfl
        got = dt_noalloc(wl, torch, u, nu)
        assert abs(got - want) <= 2e-6 * abs(want), (call, got, want)
        assert got == want or abs(got - want) <= 2e-6 * abs(want)
        # the Δt formula itself, from the definition (src/Flow.jl:234-244)
        inner = tuple(slice(1, -1) for _ in dims)
        s = np.zeros(u.shape[1:], np.float32)[inner]
        for d in range(D):
            ax = D - 1 - d
            up = np.roll(u[d], -1, axis=ax)[inner]
            s = s + (np.maximum(np.float32(0), up) + np.maximum(np.float32(0), -u[d][inner]))
        ref = min(10.0, 1.0 / (float(s.max()) + 5 * nu))
        assert abs(got - ref) <= 1e-5 * ref, (got, ref)
    sim.close()


def test_cfl_noalloc_on_a_stream_and_limits(wl):
    import torch
    lib = wl.load_library()
    u = np.random.default_rng(5).uniform(-1, 1, (3, 10, 12, 14)).astype(np.float32)
    a = dt_noalloc(wl, torch, u, 0.02)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        b = dt_noalloc(wl, torch, u, 0.02, stream=st.cuda_stream)
    assert a == b and 0 < a <= 10
    # beyond the static scratch: 600 x 600 interior rows > 262144 -> a clear error, nothing launched
    n = (C.c_int32 * 3)(8, 602, 602)
    x = torch.zeros(4, device="cuda")
    rc = lib.wl_cfl_noalloc(x.data_ptr(), 3, C.byref(n), 0.1, x.data_ptr(), None)
    assert rc == 1 and "262144" in lib.wl_last_error(None).decode()
    assert lib.wl_cfl_noalloc(None, 3, C.byref(n), 0.1, x.data_ptr(), None) == 1
    torch.cuda.synchronize()
