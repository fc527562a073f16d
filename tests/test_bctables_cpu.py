"""The table form of Function boundary data (waterlily.jl_b200/bctables.py) carries everything BC!/accelerate! need:
applying the tables pass by pass (component outer, direction inner, like core.jl:203-218) reproduces the oracle's
function-valued `BC!` and `accelerate!` bit for bit."""
import numpy as np
import pytest

import waterlily_jl_b200  # noqa: F401
from waterlily_jl_b200 import bctables as bt
from test_oracle_fnbc import accel_fn, bc_fn


def apply_tables(a, tables, saveexit, perdir):
    """BC! driven by tables; `a` shaped (D, [nz,] ny, nx)."""
    D = a.shape[0]
    for i in range(D):
        for j in range(D):
            ax = D - 1 - j
            v = np.moveaxis(a[i], ax, 0)                     # v[k] = plane k of direction j (a view)
            if (j + 1) in perdir:
                v[0] = v[-2]
                v[-1] = v[1]
            elif i == j:
                tb = tables[i, j]
                v[0], v[1] = tb["lo0"], tb["lo1"]
                if not saveexit or i > 0:
                    v[-1] = tb["hi"]
            else:
                tb = tables[i, j]
                v[0] = v[1] + tb["dlo"]
                v[-1] = v[-2] + tb["dhi"]


CASES = [((7, 6), (), False), ((7, 6), (2,), True), ((5, 6, 4), (3,), False), ((5, 6, 4), (), True), ((5, 6, 4), (1, 2), False)]


@pytest.mark.parametrize("dims,perdir,saveexit", CASES)
def test_tables_reproduce_function_bc(dims, perdir, saveexit):
    D = len(dims)
    rng = np.random.default_rng(11)
    ub = lambda i, x, t: np.float32(0.3 * i) + np.float32(0.1) * x[i % D] * x[(i + 1) % D] + t * x[0]
    a = rng.standard_normal((D,) + tuple(n + 2 for n in dims[::-1])).astype(np.float32)
    b = a.copy()
    bc_fn(a, ub, saveexit=saveexit, perdir=perdir, t=0.75)
    apply_tables(b, bt.bc_tables(ub, dims, 0.75, perdir), saveexit, perdir)
    assert np.array_equal(a, b)
    # a boundary value that does not vary along the tangential direction keeps the exact copy
    const = lambda i, x, t: np.float32(1.0 if i == 1 else 0.5) + 0 * x[0]
    tb = bt.bc_tables(const, dims, 0.0, perdir)
    assert all(np.all(v["dlo"] == 0) and np.all(v["dhi"] == 0) for (i, j), v in tb.items() if i != j)


def test_scalar_only_closures_and_acceleration():
    dims = (6, 5)
    ub = lambda i, x, t: t if i == 1 else (np.float32(0.25) * x[0] if x[1] > 2 else np.float32(0))   # Python `if`: not vectorisable
    rng = np.random.default_rng(12)
    a = rng.standard_normal((2, 7, 8)).astype(np.float32)
    b = a.copy()
    bc_fn(a, ub, t=0.5)
    apply_tables(b, bt.bc_tables(ub, dims, 0.5), False, ())
    assert np.array_equal(a, b)
    g = lambda i, x, t: np.float32(i) * x[0] - t
    dU = lambda i, x, t: x[1] * np.float32(0.5)
    for gg, dd in ((g, None), (None, dU), (g, dU)):
        r = rng.standard_normal((2, 7, 8)).astype(np.float32)
        q = r.copy()
        accel_fn(r, 0.5, gg, dd)
        q += bt.accel_field(gg, dd, dims, 0.5)
        assert np.array_equal(r, q)
    assert bt.accel_field(None, None, dims, 0.0) is None


def resolve(a, tables, ci, c, saveexit, perdir):
    """The per-cell form a fused device BC! needs (the constant-uBC kernel k_bc_vec already works this way): instead
    of D² sequential passes, walk the passes BACKWARDS from the cell, collecting the tangential differences met on
    the way, until a Dirichlet table value or an untouched cell is reached; then add the differences in pass order.
    `c` is (x, y[, z]); arrays are indexed [z][y][x]."""
    D = a.shape[0]
    N = a.shape[1:][::-1]
    c = list(c)
    ds = []
    at = lambda tb, dj: tb[tuple(c[d] for d in range(D - 1, -1, -1) if d != dj)]      # in-plane index, (z, y, x) order
    base = None
    for dj in range(D - 1, -1, -1):
        x, n = c[dj], N[dj]
        if (dj + 1) in perdir:
            if x == 0:
                c[dj] = n - 2
            elif x == n - 1:
                c[dj] = 1
        elif ci == dj:
            tb = tables[ci, dj]
            if x <= 1:
                base = at(tb["lo0"] if x == 0 else tb["lo1"], dj)
                break
            if x == n - 1 and not (saveexit and ci == 0):
                base = at(tb["hi"], dj)
                break
        else:
            tb = tables[ci, dj]
            if x == 0:
                ds.append(at(tb["dlo"], dj))
                c[dj] = 1
            elif x == n - 1:
                ds.append(at(tb["dhi"], dj))
                c[dj] = n - 2
    if base is None:
        base = a[(ci,) + tuple(c[::-1])]
    for d in reversed(ds):                                   # lowest direction first, like the sequential passes
        base = np.float32(base + d)
    return base


@pytest.mark.parametrize("dims,perdir,saveexit", CASES)
def test_backward_walk_equals_sequential_passes(dims, perdir, saveexit):
    D = len(dims)
    rng = np.random.default_rng(13)
    ub = lambda i, x, t: np.float32(0.3 * i) + np.float32(0.1) * x[i % D] * x[(i + 1) % D] + t * x[0]
    a = rng.standard_normal((D,) + tuple(n + 2 for n in dims[::-1])).astype(np.float32)
    tables = bt.bc_tables(ub, dims, 0.25, perdir)
    seq = a.copy()
    apply_tables(seq, tables, saveexit, perdir)
    out = a.copy()
    for ci in range(D):
        for idx in np.ndindex(*a.shape[1:]):
            out[(ci,) + idx] = resolve(a, tables, ci, idx[::-1], saveexit, perdir)
    assert np.array_equal(out, seq)


def resolve_flat(a, flat, ci, c, saveexit, per):
    """`bc_resolve_tab` of wl_kernels_mom.cuh transliterated (same offsets, plane sizes and in-plane index) on the
    packed array `wl_set_bc_tables` receives: checks the packing against the kernel's addressing."""
    D = a.shape[0]
    N = list(a.shape[1:][::-1]) + [1] * (3 - D)
    psz = [N[0] * N[1] * N[2] // N[dj] if dj < D else 0 for dj in range(3)]
    off, o = {}, 0
    for i in range(D):
        for j in range(D):
            off[i * D + j] = o
            o += 3 * psz[j]
    c = list(c) + [0] * (3 - D)
    ds, base, have = [], None, False
    for dj in range(D - 1, -1, -1):
        if have:
            continue
        x, n = c[dj], N[dj]
        if per[dj]:
            c[dj] = n - 2 if x == 0 else 1 if x == n - 1 else x
            continue
        ia, ib = (1 if dj == 0 else 0), (1 if dj == 2 else 2)
        idx = c[ia] + (N[ia] * c[ib] if D == 3 else 0)
        T, ps = off[ci * D + dj], psz[dj]
        if ci == dj:
            if x <= 1:
                base, have = flat[T + (0 if x == 0 else ps) + idx], True
            elif x == n - 1 and not (saveexit and ci == 0):
                base, have = flat[T + 2 * ps + idx], True
        else:
            if x == 0:
                ds.append(flat[T + idx]); c[dj] = 1
            elif x == n - 1:
                ds.append(flat[T + 2 * ps + idx]); c[dj] = n - 2
    if not have:
        base = a[(ci,) + tuple(c[:D][::-1])]
    if len(ds) == 2:
        base = np.float32(base + ds[1])
    if len(ds) >= 1:
        base = np.float32(base + ds[0])
    return base


@pytest.mark.parametrize("dims,perdir,saveexit", CASES)
def test_packed_tables_match_kernel_addressing(dims, perdir, saveexit):
    D = len(dims)
    rng = np.random.default_rng(14)
    ub = lambda i, x, t: np.float32(0.3 * i) + np.float32(0.1) * x[i % D] * x[(i + 1) % D] + t * x[0]
    a = rng.standard_normal((D,) + tuple(n + 2 for n in dims[::-1])).astype(np.float32)
    tables = bt.bc_tables(ub, dims, 0.25, perdir)
    flat = bt.pack(tables, dims)
    seq = a.copy()
    apply_tables(seq, tables, saveexit, perdir)
    per = [(d + 1) in perdir for d in range(3)]
    out = a.copy()
    for ci in range(D):
        for idx in np.ndindex(*a.shape[1:]):
            out[(ci,) + idx] = resolve_flat(a, flat, ci, idx[::-1], saveexit, per)
    assert np.array_equal(out, seq)


@pytest.mark.parametrize("nranks", [2, 4])
def test_slab_tables_are_slices_of_the_global_tables(nranks):
    """A rank of a slab-decomposed run samples ITS slab (local plane indices, global coordinates, `k_offset`/`n_last`):
    the x- and y-normal tables and the acceleration field must be the corresponding z-slices of the global ones, and the
    z-normal tables of the first / last rank must equal the global lower / upper planes (interface ranks never read theirs)."""
    T = np.float32
    dims = (10, 6, 8)
    ub = lambda i, x, t: T(0.3 * i) + T(0.1) * x[i % 3] * x[(i + 1) % 3] + T(0.01) * t * x[2]
    g = lambda i, x, t: T(0.01) * x[i % 3] - T(0.02) * t * x[(i + 1) % 3]
    t = 0.7
    glob = bt.bc_tables(ub, dims, t)
    gacc = bt.accel_field(g, None, dims, t)
    nl = dims[2] // nranks
    for r in range(nranks):
        k0 = r * nl
        loc = bt.bc_tables(ub, dims, t, k_offset=k0, n_last=nl)
        acc = bt.accel_field(g, None, dims, t, k_offset=k0, n_last=nl)
        assert np.array_equal(acc, gacc[:, k0:k0 + nl + 2])
        for (i, j), tb in loc.items():
            for name, plane in tb.items():
                want = glob[i, j][name]
                if j < 2:                                  # planes over (z, ·): the slab's z range, ghosts included
                    assert np.array_equal(plane, want[k0:k0 + nl + 2]), (r, i, j, name)
                elif (r == 0 and name in ("lo0", "lo1", "dlo")) or (r == nranks - 1 and name in ("hi", "dhi")):
                    assert np.array_equal(plane, want), (r, i, j, name)
        flat = bt.pack(loc, dims, nl)
        assert flat.size == sum(3 * 3 * (np.prod([10 + 2, 6 + 2, nl + 2]) // n) for n in (10 + 2, 6 + 2, nl + 2))
