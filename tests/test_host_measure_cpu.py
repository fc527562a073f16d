"""CPU checks of the host-side `measure!` / `nds` a closure body takes (waterlily.jl_b200.host_measure / host_nds,
src/Body.jl:28-48, src/Metrics.jl:110-113): against the oracle's measure! and pressure_force of the same sphere."""
import numpy as np
import pytest

from oracle import wl_oracle as wo


@pytest.fixture(scope="module")
def wl():
    import __graft_entry__ as ge
    ge.build()
    import waterlily_jl_b200 as wl
    return wl


@pytest.mark.parametrize("dims,c,R", [((32, 24), (12.3, 11.1), 4.2), ((24, 16, 16), (10.4, 8.2, 7.7), 3.6)])
def test_host_measure_matches_oracle_sphere(wl, dims, c, R):
    D = len(dims)
    cc = np.float32(c)
    rr = lambda x: np.sqrt(sum((xx - q) ** 2 for xx, q in zip(x, cc)))
    body = wl.ClosureBody.from_sdf(lambda x, t: (rr(x) - np.float32(R)).astype(np.float32),
                                   lambda x, t: [((xx - q) / rr(x)).astype(np.float32) for xx, q in zip(x, cc)])
    mu0, mu1, V, sig = wl.host_measure(body, dims, 0.0, 1.0)
    ref = wo.OracleSim(dims, (1.0,) + (0.0,) * (D - 1), 8.0, body=wo.make_body("sphere", c, R), build_poisson=False)
    inner = tuple(slice(1, -1) for _ in dims)
    # the oracle's fields already carry the two BC! of Body.jl:49-50: apply them to the host-measured ones too
    import ctypes as C
    ng = (C.c_int * 3)(*([d + 2 for d in dims] + [1] * (3 - D)))
    per = (C.c_int * 3)(0, 0, 0)
    zero = (C.c_float * 3)(0, 0, 0)
    ref.lib.wlo_BC(mu0.ctypes.data, D, ng, zero, 0, per)
    ref.lib.wlo_BC(V.ctypes.data, D, ng, zero, 0, per)
    for a, nm in ((mu0, "mu0"), (V, "V"), (mu1, "mu1")):
        assert float(np.abs(a - ref.field(nm)).max()) <= 2e-6, nm
    assert float(np.abs(sig[inner] - ref.field("sigma")[inner]).max()) <= 2e-6
    assert np.count_nonzero(mu1) > 50
    # nds weights reproduce the oracle's pressure_force for any p
    rng = np.random.default_rng(5)
    p = ref.field("p")
    p[...] = rng.standard_normal(p.shape).astype(np.float32)
    idx, w = wl.host_nds(body, dims, 0.0)
    f = (p.ravel()[idx][:, None] * w).astype(np.float32).astype(np.float64).sum(axis=0)
    f_ref = ref.pressure_force(0.0)
    assert np.allclose(f, f_ref[:D], rtol=1e-5, atol=1e-5 * np.abs(f_ref).max()), (f, f_ref)
