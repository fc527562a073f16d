"""Host-side evaluation of Function boundary data into tables (groundwork).

Closures cannot cross the C ABI, so a `uBC(i,x,t)` / `g(i,x,t)` given as a function is sampled by the caller on exactly
the points the reference's `BC!` (src/core.jl:201-219) and `accelerate!` (src/Flow.jl:69-73) evaluate it at, in Float32:

* normal component (i == j): the Dirichlet values on planes 1, 2 and N of direction j      -> `lo0`, `lo1`, `hi`
* tangential components (i != j): `uBC(I) - uBC(I±δ)` on the two ghost planes               -> `dlo`, `dhi`
  (the ghost becomes `a[I±δ] + d`; d is exactly 0 where uBC does not vary along j, which keeps the plain copy the
  reference's own test demands, test/test_core.jl:24-27)
* acceleration: `g(i,loc(i,I),t) + ∂ₜuBC(i,loc(i,I),t)` on every index of every component  -> one field

Each table is a numpy array shaped like the field with axis j removed, in the mirror's (z, y, x) order.  The tables are
what a device-side BC!/accelerate! consumes (one upload per step when the data depend on t, once otherwise);
`tests/test_bctables_cpu.py` shows that applying them pass by pass reproduces the CPU checker's `BC!`/`accelerate!` with
functions bit for bit.  Consumers: `wl_set_bc_tables` (k_bc_vec_tab) and `wl_set_accel_field` (K-MOM's field slot).
"""
import numpy as np


def _loc(i, dims, fixed=None, dtype=np.float32, k_offset=0):
    """`loc(i,I)` (src/core.jl:177) for every index: D broadcastable arrays (x, y[, z]) in (z, y, x) array order;
    `fixed = (j, index)` pins direction j to one plane (the axis stays, with length 1).  `k_offset`: the slab's first
    plane in the global grid (indices of the last dimension are local, coordinates global)."""
    D = len(dims)
    out = []
    for d in range(D):
        n = dims[d] + 2
        idx = np.arange(n) if not (fixed and fixed[0] == d) else np.array([fixed[1]])
        if d == D - 1:
            idx = idx + k_offset
        x = idx.astype(dtype) - dtype(0.5) - (dtype(0.5) if i == d else dtype(0))
        shp = [1] * D
        shp[D - 1 - d] = len(idx)
        out.append(x.reshape(shp))
    return out


def _eval(fn, i, x, t, shape, dtype):
    """fn(i+1, x, t) on broadcastable coordinate arrays; falls back to a point-by-point loop for scalar-only closures."""
    try:
        v = np.asarray(fn(i + 1, x, dtype(t)), dtype)
        return np.ascontiguousarray(np.broadcast_to(v, shape))
    except (ValueError, TypeError):
        xs = [np.broadcast_to(a, shape) for a in x]
        out = np.empty(shape, dtype)
        for I in np.ndindex(*shape):
            out[I] = fn(i + 1, [a[I] for a in xs], dtype(t))
        return out


def _local(dims, n_last):
    return tuple(dims) if n_last is None else tuple(dims[:-1]) + (int(n_last),)


def bc_tables(uBC, dims, t, perdir=(), dtype=np.float32, k_offset=0, n_last=None):
    """{(i, j): {...}} for every component i and non-periodic direction j (both 0-based).  A slab of a decomposed run
    passes its plane offset and plane count: the tables then cover the slab (one ghost plane per side) at its global
    coordinates; the planes of the last direction are only meaningful on the ranks that own a physical end."""
    dims = _local(dims, n_last)
    D = len(dims)
    full = tuple(n + 2 for n in dims[::-1])
    tables = {}
    for i in range(D):
        for j in range(D):
            if (j + 1) in perdir:
                continue
            ax = D - 1 - j
            shape = tuple(1 if a == ax else full[a] for a in range(D))
            at = lambda k: _eval(uBC, i, _loc(i, dims, (j, k), dtype, k_offset), t, shape, dtype).squeeze(ax)
            N = dims[j] + 2
            if i == j:
                tables[i, j] = dict(lo0=at(0), lo1=at(1), hi=at(N - 1))
            else:
                tables[i, j] = dict(dlo=at(0) - at(1), dhi=at(N - 1) - at(N - 2))
    return tables


def accel_field(g, dUdt, dims, t, dtype=np.float32, k_offset=0, n_last=None):
    """`g + ∂ₜuBC` on every index of every component, shaped (D, [nz+2,] ny+2, nx+2); None when both are absent.
    `k_offset`, `n_last`: the slab of a decomposed run (see bc_tables)."""
    if g is None and dUdt is None:
        return None
    dims = _local(dims, n_last)
    D = len(dims)
    full = tuple(n + 2 for n in dims[::-1])
    out = np.empty((D,) + full, dtype)
    for i in range(D):
        x = _loc(i, dims, None, dtype, k_offset)
        if g is not None and dUdt is not None:
            out[i] = _eval(g, i, x, t, full, dtype) + _eval(dUdt, i, x, t, full, dtype)
        else:
            out[i] = _eval(g if g is not None else dUdt, i, x, t, full, dtype)
    return out


def pack(tables, dims, n_last=None):
    """The flat array `wl_set_bc_tables` takes (include/wl_b200.h): component outer, direction inner, three planes each
    (`lo0, lo1, hi` or `dlo, 0, dhi`; zeros for periodic directions), every plane in (z, y, x)-minus-j C order."""
    dims = _local(dims, n_last)
    D = len(dims)
    full = tuple(n + 2 for n in dims[::-1])
    parts = []
    for i in range(D):
        for j in range(D):
            shape = tuple(full[a] for a in range(D) if a != D - 1 - j)
            tb = tables.get((i, j))
            if tb is None:
                planes = [np.zeros(shape, np.float32)] * 3
            elif i == j:
                planes = [tb["lo0"], tb["lo1"], tb["hi"]]
            else:
                planes = [tb["dlo"], np.zeros(shape, np.float32), tb["dhi"]]
            for p in planes:
                assert p.shape == shape, (p.shape, shape)
                parts.append(np.ascontiguousarray(p, np.float32).ravel())
    return np.concatenate(parts)
