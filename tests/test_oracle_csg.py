"""SetBody programs over RigidMap leaves: the C++ oracle against the reference's known answers
(test/test_bodies.jl:22-25,28,62-106) and against the independent Python restatement."""
import numpy as np
import pytest

from oracle import restate2 as r2
from oracle import wl_oracle as wo

L = wo.CsgBody.leaf
PI = np.pi


def approx(got, want, dtype):
    """`all(measure(...) .≈ (d, n, V))`: scalar isapprox for d, norm-based isapprox for the vectors."""
    rt = float(np.sqrt(np.finfo(np.float32).eps))   # the node descriptor carries Float32 map fields (θ = Float32(π)) in both builds
    d, n, V = got
    wd, wn, wV = want
    assert abs(d - wd) <= rt * max(abs(d), abs(wd)), (d, wd)
    for a, b in ((n, wn), (V, wV)):
        a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
        assert np.linalg.norm(a - b) <= rt * max(np.linalg.norm(a), np.linalg.norm(b)), (a, b)


@pytest.mark.parametrize("T", [np.float32, np.float64])
def test_rigidmap_known_answers(T):
    """test/test_bodies.jl:62-94, line by line."""
    m = lambda b, x: wo.measure_point_csg(b, x, dtype=T)
    body = L("circle", (0, 0), 0.0, radius=1)
    approx(m(body, (1.5, 0)), (0.5, (1, 0), (0, 0)), T)
    body = body.setmap(theta=PI / 4, V=(1.0, 0))
    approx(m(body, (1.5, 0)), (0.5, (1, 0), (1, 0)), T)
    body = body.setmap(omega=0.1)
    approx(m(body, (1.5, 0)), (0.5, (1, 0), (1, 1.5 * 0.1)), T)
    b3 = L("sphere", (0, 0, 0), (0, 0, 0), xp=(-.5, 0, 0), radius=1)
    approx(m(b3, (1.5, 0, 0)), (0.5, (1, 0, 0), (0, 0, 0)), T)
    approx(m(b3.setmap(theta=(PI, 0, 0)), (1.5, 0, 0)), (0.5, (1, 0, 0), (0, 0, 0)), T)
    approx(m(b3.setmap(theta=(0, PI, 0), V=(1.0, 0, 0)), (1.5, 0, 0)), (1.5, (1, 0, 0), (1, 0, 0)), T)
    approx(m(b3.setmap(theta=(0, 0, PI), V=(1.0, 0, 0)), (1.5, 0, 0)), (1.5, (1, 0, 0), (1, 0, 0)), T)
    b3 = b3.setmap(theta=(0, 0, 0), V=(1.0, 0, 0), omega=(0, 0, 0.1))
    approx(m(b3, (1.5, 0, 0)), (0.5, (1, 0, 0), (1, 0.2, 0)), T)
    approx(m(b3, (0, 1.5, 0)), (0.5, (0, 1, 0), (0.85, 0.05, 0)), T)
    s3 = np.sqrt(1 / 3)
    approx(m(b3, (1.5, 1.5, 1.5)), (np.sqrt(3 * 1.5 ** 2) - 1, (s3, s3, s3), (.85, 0.2, 0)), T)
    b3 = b3.setmap(V=(1.0, 0, 0), omega=(0, -0.1, 0.1))
    approx(m(b3, (1.5, 0, 0)), (0.5, (1, 0, 0), (1, 0.2, 0.2)), T)
    s2 = np.sqrt(1 / 2)
    approx(m(b3, (0, 1.5, 1.5)), (np.sqrt(2 * 1.5 ** 2) - 1, (0, s2, s2), (0.7, 0.05, 0.05)), T)


def test_setbody_known_answers():
    """test/test_bodies.jl:22-28 (booleans; the moving bodies there are frozen at t=1) and :104-106 (annulus)."""
    T = np.float64
    # body1 = circle of radius 2+t at the origin, body2 = circle of radius 2 centred at -t², velocity -2t; t = 1
    b1 = L("circle", (0, 0), 0.0, radius=3)
    b2 = L("circle", (-1, -1), 0.0, radius=2, V=(-2, -2))
    x = (-np.sqrt(2.), -np.sqrt(2.))
    h = np.sqrt(.5)
    approx(wo.measure_point_csg(b1 + b2, x, dtype=T), (-np.sqrt(2.), (-h, -h), (-2, -2)), T)
    approx(wo.measure_point_csg(b1 | b2, x, dtype=T), (-np.sqrt(2.), (-h, -h), (-2, -2)), T)
    approx(wo.measure_point_csg(b1 - b2, x, dtype=T), (np.sqrt(2.), (h, h), (-2, -2)), T)
    # exactly equal distances (test_bodies.jl:28)
    two = L("circle", (0, 0), 0.0, radius=2) + L("circle", (6, 0), 0.0, radius=2)
    assert wo.measure_point_csg(two, (3., 0.), fastd2=0.0, dtype=T)[0] == 1
    ring = L("circle", (0, 0), PI / 4, radius=1) - L("circle", (0, 0), PI / 4, radius=0.5)
    approx(wo.measure_point_csg(ring.setmap(omega=1.), (0.25, 0.), dtype=T), (0.25, (-1, 0), (0, 0.25)), T)
    # set identity (test_bodies.jl:45-50): (b1 ∩ b2) ∪ b1 == b1
    for x in [(0.5, 1.5), (2.5, 0.5), (-3.0, 2.0)]:
        a = wo.measure_point_csg((b1 & b2) | b1, x, fastd2=0.0, dtype=T)[0]
        assert np.isclose(a, wo.measure_point_csg(b1, x, fastd2=0.0, dtype=T)[0])


def test_simulation_rigid_sphere():
    """test/test_bodies.jl:95-103: rotating sphere in a 32³ box: extrema(V) ≈ ∓0.9, and μ₀ = 0 inside after a shift."""
    body = L("sphere", (16, 16, 16), (0, 0, 0), radius=4, omega=(0, -0.1, 0.1))
    sim = wo.OracleSim((32, 32, 32), (1, 0, 0), 8.0, body=body, build_poisson=False)
    V = sim.field("V")
    assert np.isclose(V.min(), -0.9, rtol=1e-3) and np.isclose(V.max(), 0.9, rtol=1e-3)
    sim.body = sim.body.setmap(x0=(16, 16, 12))                # no remeasure in the reference test either
    assert np.all(sim.field("mu0")[:, 16, 16, 16] == 0)        # Julia [17,17,17,:]
    sim._measure(0.0)                                          # after measure! the shifted sphere's band reaches that cell
    assert np.all(sim.field("mu0")[:, 16, 16, 16] > 0) and np.all(sim.field("mu0")[:, 12, 16, 16] == 0)


def _tree(body):
    """postfix CsgBody -> the nested tuples restate2.measure_set takes."""
    names = {v: k for k, v in wo.LEAF_KINDS.items() if k != "circle"}
    st = []
    for p in body.prog:
        if p["op"] == 0:
            st.append(("leaf", dict(p, kind=names[p["kind"]])))
        elif p["op"] == 3:
            st.append(("neg", st.pop()))
        else:
            b, a = st.pop(), st.pop()
            st.append(("min" if p["op"] == 1 else "max", a, b))
    return st[0]


CASES = [
    ((16, 12), lambda: L("circle", (6.3, 5.7), 0.0, radius=3.2, V=(0.25, 0.0), omega=0.1), False, ()),
    ((16, 12), lambda: L("capsule", (7.5, 6.2), 0.6, radius=1.1, minor=3.0, axis=0, xp=(0.5, 0.0), omega=-0.2)
                       + L("circle", (10.0, 4.0), 0.0, radius=2.0), True, (2,)),
    ((16, 12), lambda: L("circle", (8.0, 6.0), 0.3, radius=4.0) - L("circle", (8.0, 6.0), 0.3, radius=2.0), False, ()),
    ((10, 8, 8), lambda: L("torus", (5.0, 4.0, 4.0), (0.3, -0.2, 0.5), radius=2.5, minor=1.0, axis=2, V=(0.1, 0, 0),
                           omega=(0.0, 0.05, -0.1)) & -L("sphere", (7.0, 4.0, 4.0), (0, 0, 0), radius=1.5), False, (3,)),
    ((10, 8, 8), lambda: L("cylinder", (5.0, 4.0, 4.0), (0.0, 0.4, 0.0), radius=2.0, axis=2)
                         & L("capsule", (5.0, 4.0, 4.0), (0, 0, 0), radius=3.0, minor=1.0, axis=1), True, ()),
]


@pytest.mark.parametrize("dims,mk,exitBC,perdir", CASES)
def test_measure_csg_two_restatements(dims, mk, exitBC, perdir):
    """measure! over a SetBody of RigidMap leaves by the C++ oracle and by restate2 (band logic, sign fix,
    measure_sdf! through measure(...)[1] for sets, lexicographic min/max)."""
    D = len(dims)
    body = mk()
    sim = wo.OracleSim(dims, (1.0,) + (0.0,) * (D - 1), perdir=perdir, exitBC=exitBC, body=body, build_poisson=False)
    tree = _tree(body)
    fl = {nm: np.zeros_like(sim.field(nm)).T for nm in ("V", "mu0", "mu1", "sigma")}
    r2.measure_flow_fn(fl, lambda x, fd2: r2.measure_set(tree, x, fd2), lambda x, fd2: r2.sdf_set(tree, x, fd2),
                       1.0, exitBC, perdir)
    inner = tuple(slice(1, -1) for _ in dims)
    assert np.allclose(sim.field("sigma")[inner], fl["sigma"].T[inner], rtol=0, atol=5e-6)
    assert (np.abs(sim.field("mu1")) > 0).sum() > 20
    for nm in ("V", "mu0", "mu1"):
        a, c = sim.field(nm), fl[nm].T
        assert np.allclose(a, c, rtol=0, atol=5e-6), (nm, np.abs(a - c).max())


@pytest.mark.parametrize("exitBC", [True, False])
def test_rotating_plate_simulation(exitBC):
    """test/test_simulation.jl:32-36 ("non-uniform V doesn't break"): a plate (capsule of half-length radius-2 and
    thickness radius 2) under `rotate(x,t) = R(t/radius+1)·(x − 2radius)`, i.e. a RigidMap with θ(t) = t/radius+1 and
    ω = 1/radius, impulsively started in still fluid: two Poisson solves of < 5 iterations and 0.5 < Δt < 1."""
    radius = 8
    nu = radius / 250
    theta = lambda t: np.float32(t / radius + 1)
    plate = L("capsule", (2 * radius, 2 * radius), theta(0.0), radius=2, minor=radius - 2, axis=0, omega=1 / radius)
    sim = wo.OracleSim((4 * radius, 4 * radius), (0.0, 0.0), radius, U=1, body=plate, nu=nu, exitBC=exitBC)
    sim.body = sim.body.setmap(theta=theta(float(np.sum(sim.dts))))      # measure!(sim) at t = sum(Δt)
    sim.sim_step()
    assert len(sim.n) == 2 and all(n < 5 for n in sim.n), sim.n
    assert 1 > sim.dts[-1] > 0.5, sim.dts
    V = sim.field("V")
    assert V.max() > 0.5 and abs(V.max() + V.min()) < 0.2                 # |V| = ω·r up to the plate tip, both senses


def _random_body(rng, D, depth=0):
    """A random SetBody tree of at most 8 postfix nodes over random RigidMap leaves."""
    kinds = ["sphere", "cylinder", "torus", "capsule"] if D == 3 else ["circle", "capsule"]

    def leaf():
        k = kinds[rng.integers(len(kinds))]
        v = lambda s: tuple(float(x) for x in rng.uniform(-s, s, D))
        th = tuple(float(x) for x in rng.uniform(-1.5, 1.5, 3)) if D == 3 else float(rng.uniform(-3, 3))
        om = tuple(float(x) for x in rng.uniform(-0.2, 0.2, 3)) if D == 3 else float(rng.uniform(-0.2, 0.2))
        return L(k, v(2.0), th, xp=v(1.0), V=v(0.5), omega=om, radius=float(rng.uniform(1.0, 3.0)),
                 minor=float(rng.uniform(0.3, 1.5)), axis=int(rng.integers(D)))
    b = leaf()
    while len(b.prog) < 6 and rng.random() < 0.7:
        op = rng.integers(4)
        b = -b if op == 3 else (b + leaf() if op == 0 else b & leaf() if op == 1 else b - leaf())
    return b if len(b.prog) <= 8 else leaf()


@pytest.mark.parametrize("D", [2, 3])
def test_random_programs_two_restatements(D):
    """Random trees × random points × both fastd² regimes: C++ stack machine vs the recursive Python restatement
    (lexicographic tuple min/max, complement, the fastd² early return per leaf)."""
    rng = np.random.default_rng(100 + D)
    n_checked = 0
    for _ in range(40):
        body = _random_body(rng, D)
        tree = _tree(body)
        for _ in range(25):
            x = [np.float32(v) for v in rng.uniform(-5, 5, D)]
            for fd2 in (np.float32(9.0), np.float32(0.0), np.float32(np.inf)):
                d, n, V = wo.measure_point_csg(body, x, fastd2=fd2)
                d2, n2, V2 = r2.measure_set(tree, x, fd2)
                assert np.isclose(d, d2, rtol=2e-6, atol=2e-6), (body.prog, x, d, d2)
                assert np.allclose(n, n2, rtol=0, atol=5e-6) and np.allclose(V, V2, rtol=0, atol=5e-6), (body.prog, x)
                n_checked += 1
    assert n_checked == 40 * 25 * 3
