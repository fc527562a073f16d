"""Second, independent restatement of the momentum-side loops in plain Python (TEST INFRASTRUCTURE).

SURVEY §8c names the residual risk of the oracle: the C++ restatement (oracle/wl_oracle.cpp) could
deviate from the Julia semantics.  This file restates the same `@loop`s a second time, written
directly from the Julia source with Julia's own 1-based CartesianIndex arithmetic (no shared code, no
shared index conventions with the C++ file), one Python loop per cell, Float32 scalars.  It is only
usable on tiny grids; tests/test_oracle_restatement2.py compares it with the C++ oracle bit for bit.

Arrays are numpy views in Julia order: a[i-1, j-1(, k-1)(, c-1)] == Julia a[i,j(,k)(,c)]
(`OracleSim.field(name).T`).  Only tests/ may import this module.
"""
import itertools

import numpy as np

f32 = np.float32


class J:
    """1-based view of a numpy array: J(a)[I] with I a tuple of Julia indices."""

    def __init__(self, a):
        self.a = a

    def __getitem__(self, I):
        return self.a[tuple(x - 1 for x in I)]

    def __setitem__(self, I, v):
        self.a[tuple(x - 1 for x in I)] = v


def CR(*ranges):      # CartesianIndices((lo:hi, ...)), first index fastest like Julia
    for I in itertools.product(*[range(lo, hi + 1) for lo, hi in reversed(ranges)]):
        yield tuple(reversed(I))


def d_(j, n):          # δ(j, Val{n}) core.jl:39
    return tuple(1 if k == j else 0 for k in range(1, n + 1))


def plus(I, K, s=1):
    return tuple(a + s * b for a, b in zip(I, K))


def CIj(j, I, k):      # core.jl:31
    return tuple(k if i == j else I[i - 1] for i in range(1, len(I) + 1))


def inside_u(N, j):    # core.jl:55-57
    return CR(*[(3, N[i - 1] - 1) if i == j else (2, N[i - 1]) for i in range(1, len(N) + 1)])


def slice_(N, i, j, low=1):   # core.jl:188-190
    return CR(*[(i, i) if k == j else (low, N[k - 1]) for k in range(1, len(N) + 1)])


def median(a, b, c):   # Flow.jl:27-36
    if a > b:
        if b >= c:
            return b
        if a > c:
            return c
    else:
        if b <= c:
            return b
        if a < c:
            return c
    return a


def quick(u, c, d):    # Flow.jl:4
    return median((f32(5) * c + f32(2) * d - u) / f32(6), c, median(f32(10) * c - f32(9) * u, c, d))


def vanLeer(u, c, d):  # Flow.jl:5
    return c if (c <= min(u, d) or c >= max(u, d)) else c + (d - c) * (c - u) / (d - u)


def cds(u, c, d):      # Flow.jl:6
    return (c + d) / f32(2)


SCHEME = dict(quick=quick, vanLeer=vanLeer, cds=cds)


def conv_diff(r, u, Phi, lam, nu, perdir):
    """`conv_diff!(r,u,Φ,λ;ν,perdir)` Flow.jl:38-62.  r,u: (N...,n); Φ: (N...)."""
    r[...] = 0
    N, n = u.shape[:-1], u.shape[-1]
    R, U, F = J(r), J(u), J(Phi)
    nu = f32(nu)

    def phi(a, I, f):                                   # ϕ(a,I,f) Flow.jl:3 ; I carries the component
        return (f[I] + f[plus(I, d_(a, n + 1), -1)]) / f32(2)

    def dd(a, I, f):                                    # ∂(a,I::CI{d},f::Array{T,d}) Flow.jl:1
        return f[I] - f[plus(I, d_(a, n + 1), -1)]

    def phiu(a, I, f, ua):                              # ϕu Flow.jl:8
        e = d_(a, n + 1)
        if ua > 0:
            return ua * lam(f[plus(I, e, -2)], f[plus(I, e, -1)], f[I])
        return ua * lam(f[plus(I, e)], f[I], f[plus(I, e, -1)])

    def phiuP(a, Ip, I, f, ua):                         # ϕuP Flow.jl:9
        e = d_(a, n + 1)
        if ua > 0:
            return ua * lam(f[Ip], f[plus(I, e, -1)], f[I])
        return ua * lam(f[plus(I, e)], f[I], f[plus(I, e, -1)])

    def phiuL(a, I, f, ua):                             # ϕuL Flow.jl:10
        e = d_(a, n + 1)
        if ua > 0:
            return ua * phi(a, I, f)
        return ua * lam(f[plus(I, e)], f[I], f[plus(I, e, -1)])

    def phiuR(a, I, f, ua):                             # ϕuR Flow.jl:11
        e = d_(a, n + 1)
        if ua < 0:
            return ua * phi(a, I, f)
        return ua * lam(f[plus(I, e, -2)], f[plus(I, e, -1)], f[I])

    for i in range(1, n + 1):
        for j in range(1, n + 1):
            ej = d_(j, n)
            if j in perdir:                              # lowerBoundary! Val{true} Flow.jl:60-61
                for I in slice_(N, 2, j, 2):
                    Ii = I + (i,)
                    F[I] = phiuP(j, CIj(j, Ii, N[j - 1] - 2), Ii, U, phi(i, I + (j,), U)) - nu * dd(j, Ii, U)
                    R[Ii] = R[Ii] + F[I]
            else:                                        # lowerBoundary! Val{false} Flow.jl:56
                for I in slice_(N, 2, j, 2):
                    Ii = I + (i,)
                    R[Ii] = R[Ii] + (phiuL(j, Ii, U, phi(i, I + (j,), U)) - nu * dd(j, Ii, U))
            for I in inside_u(N, j):                     # Flow.jl:46-48
                Ii = I + (i,)
                F[I] = phiu(j, Ii, U, phi(i, I + (j,), U)) - nu * dd(j, Ii, U)
                R[Ii] = R[Ii] + F[I]
            for I in inside_u(N, j):                     # Flow.jl:49
                Im = plus(I, ej, -1) + (i,)
                R[Im] = R[Im] - F[I]
            if j in perdir:                              # upperBoundary! Val{true} Flow.jl:62
                for I in slice_(N, N[j - 1], j, 2):
                    Im = plus(I, ej, -1) + (i,)
                    R[Im] = R[Im] - F[CIj(j, I, 2)]
            else:                                        # upperBoundary! Val{false} Flow.jl:57
                for I in slice_(N, N[j - 1], j, 2):
                    Ii = I + (i,)
                    Im = plus(I, ej, -1) + (i,)
                    R[Im] = R[Im] + (-phiuR(j, Ii, U, phi(i, I + (j,), U)) + nu * dd(j, Ii, U))


def BDIM(u, u0, f, V, mu0, mu1, dt):
    """`BDIM!(a)` Flow.jl:176-180 with `μddn` :20-26.  mu1: (N...,n,n)."""
    N, n = u.shape[:-1], u.shape[-1]
    dt = f32(dt)
    f[...] = u0 + dt * f - V                              # all indices
    Uu, Ff, Vv, M0, M1 = J(u), J(f), J(V), J(mu0), J(mu1)
    for Ii in CR(*[(2, k - 1) for k in N], (1, n)):        # inside_u(size(p)) core.jl:58
        I, i = Ii[:-1], Ii[-1]
        s = f32(0)
        for j in range(1, n + 1):
            e = d_(j, n)
            s = s + M1[I + (i, j)] * (Ff[plus(I, e) + (i,)] - Ff[plus(I, e, -1) + (i,)])
        Uu[Ii] = Uu[Ii] + (s / f32(2) + Vv[Ii] + M0[Ii] * Ff[Ii])


def BC(a, A, saveexit, perdir):
    """`BC!(a,A,saveexit,perdir)` with a constant tuple A, core.jl:200-219."""
    N, n = a.shape[:-1], a.shape[-1]
    Aa = J(a)
    for i in range(1, n + 1):
        for j in range(1, n + 1):
            e = d_(j, n)
            if j in perdir:
                for I in slice_(N, 1, j):
                    Aa[I + (i,)] = Aa[CIj(j, I, N[j - 1] - 1) + (i,)]
                for I in slice_(N, N[j - 1], j):
                    Aa[I + (i,)] = Aa[CIj(j, I, 2) + (i,)]
            elif i == j:
                for s in (1, 2):
                    for I in slice_(N, s, j):
                        Aa[I + (i,)] = f32(A[i - 1])
                if (not saveexit) or i > 1:
                    for I in slice_(N, N[j - 1], j):
                        Aa[I + (i,)] = f32(A[i - 1])
            else:
                # `uBC + a[I±δ] - uBC` with a constant tuple: the reference's own test compares the result
                # with == against the neighbour (test/test_core.jl:24-27), i.e. the copy is exact
                for I in slice_(N, 1, j):
                    Aa[I + (i,)] = Aa[plus(I, e) + (i,)]
                for I in slice_(N, N[j - 1], j):
                    Aa[I + (i,)] = Aa[plus(I, e, -1) + (i,)]


def loc_(i, I, T=f32):
    """`loc(i,I,T)` core.jl:177 for a 1-based CartesianIndex tuple."""
    return [T(I[k] - 1.5) - (T(0.5) if k + 1 == i else T(0)) for k in range(len(I))]


def BC_fn(a, uBC, saveexit, perdir, t, T=f32):
    """`BC!(a,uBC::Function,saveexit,perdir,t)` core.jl:201-219.  The tangential sum is written as
    a[I±δ] + (uBC(I) − uBC(I±δ)): an exact copy when uBC does not vary along j (see `BC`)."""
    N, n = a.shape[:-1], a.shape[-1]
    Aa = J(a)
    ub = lambda i, I: T(uBC(i, loc_(i, I, T), T(t)))
    for i in range(1, n + 1):
        for j in range(1, n + 1):
            e = d_(j, n)
            if j in perdir:
                for I in slice_(N, 1, j):
                    Aa[I + (i,)] = Aa[CIj(j, I, N[j - 1] - 1) + (i,)]
                for I in slice_(N, N[j - 1], j):
                    Aa[I + (i,)] = Aa[CIj(j, I, 2) + (i,)]
            elif i == j:
                for s in (1, 2):
                    for I in slice_(N, s, j):
                        Aa[I + (i,)] = ub(i, I)
                if (not saveexit) or i > 1:
                    for I in slice_(N, N[j - 1], j):
                        Aa[I + (i,)] = ub(i, I)
            else:
                for I in slice_(N, 1, j):
                    Aa[I + (i,)] = Aa[plus(I, e) + (i,)] + (ub(i, I) - ub(i, plus(I, e)))
                for I in slice_(N, N[j - 1], j):
                    Aa[I + (i,)] = Aa[plus(I, e, -1) + (i,)] + (ub(i, I) - ub(i, plus(I, e, -1)))


def accelerate_fn(r, t, g, dUdt, T=f32):
    """`accelerate!(r,t,g,U)` Flow.jl:69-73 over CartesianIndices(r); `dUdt` is ∂ₜU in closed form."""
    N, n = r.shape[:-1], r.shape[-1]
    R = J(r)
    for i in range(1, n + 1):
        for I in CR(*[(1, m) for m in N]):
            x = loc_(i, I, T)
            if g is not None and dUdt is not None:
                R[I + (i,)] += T(g(i, x, T(t))) + T(dUdt(i, x, T(t)))
            elif g is not None:
                R[I + (i,)] += T(g(i, x, T(t)))
            elif dUdt is not None:
                R[I + (i,)] += T(dUdt(i, x, T(t)))


def CFL(u, sigma, nu):
    """`CFL(a)` Flow.jl:234-244: σ = flux_out inside, then the maximum over the WHOLE σ."""
    N, n = u.shape[:-1], u.shape[-1]
    Uu, S = J(u), J(sigma)
    for I in CR(*[(2, k - 1) for k in N]):
        s = f32(0)
        for i in range(1, n + 1):
            s = s + (max(f32(0), Uu[plus(I, d_(i, n)) + (i,)]) + max(f32(0), -Uu[I + (i,)]))
        S[I] = s
    return min(f32(10), f32(1) / (sigma.max() + f32(5) * f32(nu)))


# ---------------------------------------------------------------------------------------------------
# Pressure side: Poisson.jl:43-148, MultiLevelPoisson.jl:1-50,88-101 (one level = dict of Julia-order arrays)

def inside(a):         # core.jl:47
    return CR(*[(2, k - 1) for k in a.shape])


def perBC(a, perdir):  # core.jl:239-243
    N = a.shape
    A = J(a)
    for j in perdir:
        for I in slice_(N, 1, j):
            A[I] = A[CIj(j, I, N[j - 1] - 1)]
        for I in slice_(N, N[j - 1], j):
            A[I] = A[CIj(j, I, 2)]


def diag(I, L):        # Poisson.jl:49-55
    d = len(I)
    s = f32(0)
    for i in range(1, d + 1):
        s = s - (L[I + (i,)] + L[plus(I, d_(i, d)) + (i,)])
    return s


def set_diag(p):       # Poisson.jl:43-46
    D, iD, L = J(p["D"]), J(p["iD"]), J(p["L"])
    for I in inside(p["D"]):
        D[I] = diag(I, L)
    for I in inside(p["D"]):
        iD[I] = D[I] if D[I] == 0 else f32(1) / D[I]


def mult(I, L, D, x):  # Poisson.jl:70-76
    d = len(I)
    s = x[I] * D[I]
    for i in range(1, d + 1):
        e = d_(i, d)
        s = s + (x[plus(I, e, -1)] * L[I + (i,)] + x[plus(I, e)] * L[plus(I, e) + (i,)])
    return s


def residual(p):       # Poisson.jl:92-98
    perBC(p["x"], p["perdir"])
    r, iD, z, L, D, x = (J(p[k]) for k in ("r", "iD", "z", "L", "D", "x"))
    for I in inside(p["r"]):
        r[I] = f32(0) if iD[I] == 0 else z[I] - mult(I, L, D, x)
    n_in = int(np.prod([k - 2 for k in p["r"].shape]))
    s = f32(p["r"].sum(dtype=np.float64)) / f32(n_in)
    if abs(s) <= 2 * np.finfo(np.float32).eps:
        return
    for I in inside(p["r"]):
        r[I] = r[I] - s


def increment(p, w=1):  # Poisson.jl:100-104
    w = f32(w)
    perBC(p["eps"], p["perdir"])
    r, x, e, L, D = (J(p[k]) for k in ("r", "x", "eps", "L", "D"))
    for I in inside(p["x"]):
        r[I] = r[I] - w * mult(I, L, D, e)
        x[I] = x[I] + w * e[I]


def jacobi(p):         # Poisson.jl:111-114 (it=1, ω=1)
    e, r, iD = J(p["eps"]), J(p["r"]), J(p["iD"])
    for I in inside(p["eps"]):
        e[I] = r[I] * iD[I]
    increment(p, 1)


def gauss_seidel_rb(p, it=4, w=1):   # Poisson.jl:116-148
    e, r, iD, L = J(p["eps"]), J(p["r"]), J(p["iD"]), J(p["L"])
    for I in inside(p["eps"]):
        e[I] = r[I] * iD[I]
    perBC(p["eps"], p["perdir"])
    sh = p["eps"].shape
    d = len(sh)
    half = [(2, sh[i] // 2) if i == d - 1 else (2, sh[i] - 1) for i in range(d)]   # half_rangek :130-132
    for k0 in range(1, it + 1):
        for Iv in CR(*half):
            k = 2 * Iv[-1] - 1 - (sum(Iv[:-1]) + k0) % 2
            I = Iv[:-1] + (k,)
            s = r[I]                                                               # gauss :116-122
            for i in range(1, d + 1):
                ei = d_(i, d)
                s = s - (e[plus(I, ei, -1)] * L[I + (i,)] + e[plus(I, ei)] * L[plus(I, ei) + (i,)])
            e[I] = s * iD[I]
    increment(p, w)


def coarsen_mask(fine, coarse):     # MultiLevelPoisson.jl:29
    return tuple(coarse.shape[j] < fine.shape[j] for j in range(fine.ndim))


def up(I, c):          # MultiLevelPoisson.jl:6
    return CR(*[(2 * I[j] - 2, 2 * I[j] - 1) if c[j] else (I[j], I[j]) for j in range(len(I))])


def down(I, c):        # MultiLevelPoisson.jl:7
    return tuple((I[j] + 2) // 2 if c[j] else I[j] for j in range(len(I)))


def restrict(a, b, c):  # restrict!(a,b,c) MultiLevelPoisson.jl:13-19,49
    A, B = J(a), J(b)
    for I in inside(a):
        s = f32(0)
        for K in up(I, c):
            s = s + B[K]
        A[I] = s


def prolongate(a, b, c):  # MultiLevelPoisson.jl:50
    A, B = J(a), J(b)
    for I in inside(a):
        A[I] = B[down(I, c)]


def restrictL(a, b, c, perdir):   # MultiLevelPoisson.jl:9-11,20-26,42-48
    Na, n = a.shape[:-1], a.shape[-1]
    A, B = J(a), J(b)
    for i in range(1, n + 1):
        for I in CR(*[(2, k - 1) for k in Na]):
            rng = [((2 * I[j] - 2, 2 * I[j] - 2) if c[i - 1] else (I[j], I[j])) if j == i - 1 else
                   ((2 * I[j] - 2, 2 * I[j] - 1) if c[j] else (I[j], I[j])) for j in range(n)]
            s = f32(0)
            for K in CR(*rng):
                s = s + B[K + (i,)]
            A[I + (i,)] = s / f32(2) if c[i - 1] else s
    BC(a, (0.0,) * n, False, perdir)


def vcycle(levels, l=1, w=1):     # Vcycle!(ml;l,ω) MultiLevelPoisson.jl:88-101
    fine, coarse = levels[l - 1], levels[l]
    c = coarsen_mask(fine["x"], coarse["x"])
    jacobi(fine)
    restrict(coarse["r"], fine["r"], c)
    coarse["x"][...] = 0
    if l + 1 < len(levels):
        vcycle(levels, l + 1, w)
    gauss_seidel_rb(coarse, 4, w)
    prolongate(fine["eps"], coarse["x"], c)
    increment(fine, w)


# ---------------------------------------------------------------------------------------------------
# The whole step: exitBC! (core.jl:226-233), solver! (MultiLevelPoisson.jl:108-128), mom_project!
# (Flow.jl:223-232), mom_step! (Flow.jl:156-167).  `fl` = dict of Julia-order arrays u,u0,f,p,sigma,V,mu0,mu1
# plus scalars; levels[0] aliases x≡p, L≡mu0, z≡sigma like the reference (WaterLily.jl:97).

def exitBC(u, u0, dt):
    N, n = u.shape[:-1], u.shape[-1]
    dt = f32(dt)
    Nm = tuple(k - 1 for k in N)
    exitR = list(slice_(Nm, N[0], 1, 2))
    inR = list(slice_(Nm, 2, 1, 2))
    Uu, U0 = J(u), J(u0)
    U = f32(sum(np.float64(Uu[I + (1,)]) for I in inR)) / f32(len(exitR))
    for I in exitR:
        Uu[I + (1,)] = U0[I + (1,)] - U * dt * (U0[I + (1,)] - U0[plus(I, d_(1, n), -1) + (1,)])
    flux = f32(sum(np.float64(Uu[I + (1,)]) for I in exitR)) / f32(len(exitR)) - U
    for I in exitR:
        Uu[I + (1,)] = Uu[I + (1,)] - flux


def solver(levels, n_log, tol=2e-3, itmx=32):
    p = levels[0]
    n_in = int(np.prod([k - 2 for k in p["r"].shape]))
    r1tol, rinftol = (float(tol) / 10) * n_in, tol
    residual(p)
    L1 = lambda: f32(np.abs(p["r"]).sum(dtype=np.float64))
    Linf = lambda: f32(np.abs(p["r"]).max())
    r1, rinf, w = L1(), Linf(), f32(1)
    npn = 0
    while npn < itmx:
        vcycle(levels, 1, w)
        gauss_seidel_rb(p, 4, w)
        rnew, rinf = L1(), Linf()
        npn += 1
        if rnew >= r1:
            w = f32(max(0.2, 0.9 * float(w)))
        elif rnew < r1:
            w = f32(min(1.0, 1.02 * float(w)))
        r1 = rnew
        if float(r1) < r1tol and float(rinf) < rinftol:
            break
    perBC(p["x"], p["perdir"])
    n_log.append(npn)


def mom_project(fl, levels, w, n_log):
    u, x, z, L = fl["u"], levels[0]["x"], levels[0]["z"], levels[0]["L"]
    n = u.shape[-1]
    dt = f32(w) * f32(fl["dt"][-1])
    Uu, X, Z, Ll = J(u), J(x), J(z), J(L)
    for I in inside(z):                                   # div(I,u) Flow.jl:13-19
        s = f32(0)
        for i in range(1, n + 1):
            s = s + (Uu[plus(I, d_(i, n)) + (i,)] - Uu[I + (i,)])
        Z[I] = s
    x *= dt
    solver(levels, n_log)
    for i in range(1, n + 1):
        for I in inside(x):
            Uu[I + (i,)] = Uu[I + (i,)] - Ll[I + (i,)] * (X[I] - X[plus(I, d_(i, n), -1)])
    x /= dt
    BC(u, fl["uBC"], fl["exitBC"], fl["perdir"])


def mom_step(fl, levels, n_log):
    u, u0, f = fl["u"], fl["u0"], fl["f"]
    lam = SCHEME[fl["scheme"]]
    u0[...] = u
    n = u.shape[-1]
    u[tuple(slice(1, -1) for _ in range(n)) + (slice(None),)] = 0        # scale_u!(a,0) over inside_u(size(p))
    # predictor
    conv_diff(f, u0, fl["sigma"], lam, fl["nu"], fl["perdir"])
    BDIM(u, u0, f, fl["V"], fl["mu0"], fl["mu1"], fl["dt"][-1])
    BC(u, fl["uBC"], fl["exitBC"], fl["perdir"])
    if fl["exitBC"]:
        exitBC(u, u0, fl["dt"][-1])
    mom_project(fl, levels, 1, n_log)
    # corrector
    conv_diff(f, u, fl["sigma"], lam, fl["nu"], fl["perdir"])
    BDIM(u, u0, f, fl["V"], fl["mu0"], fl["mu1"], fl["dt"][-1])
    inner = tuple(slice(1, -1) for _ in range(n)) + (slice(None),)
    u[inner] = u[inner] * f32(0.5)
    BC(u, fl["uBC"], fl["exitBC"], fl["perdir"])
    mom_project(fl, levels, 0.5, n_log)
    fl["dt"].append(CFL(u, fl["sigma"], fl["nu"]))


# ---------------------------------------------------------------------------------------------------
# Body side: measure! (Body.jl:28-51), kernels (:54-60), measure(::AutoBody) (AutoBody.jl:29-37) for a
# sphere/circle `sdf(ξ)=|ξ|-R` under the translation map `ξ = x - (c + v t)`, pressure_force (Metrics.jl:116-133).
# The gradient of the sdf is written in closed form (the reference differentiates it with ForwardDiff).

def sinpi(d):
    return f32(np.sin(np.pi * np.float64(d)))


def cospi(d):
    return f32(np.cos(np.pi * np.float64(d)))


PI32 = f32(np.pi)


def kern(d):
    return (f32(1) + cospi(d)) / f32(2)


def kern0(d):
    return (f32(1) + d + sinpi(d) / PI32) / f32(2)


def kern1(d):
    return (f32(1) - d * d) / f32(4) - (d * sinpi(d) + (f32(1) + cospi(d)) / PI32) / (f32(2) * PI32)


def mu0(d, e):
    d, e = f32(d), f32(e)
    return f32(0) if d / e < f32(-1) + np.sqrt(np.spacing(abs(d))) else kern0(min(d / e, f32(1)))


def mu1(d, e):
    d, e = f32(d), f32(e)
    return e * kern1(max(f32(-1), min(d / e, f32(1))))


def measure_sphere(center, R, vel, x, t, fastd2):
    n = len(x)
    t = f32(t)
    xi = [f32(x[k]) - (f32(center[k]) + f32(vel[k]) * t) for k in range(n)]
    s = f32(0)
    for k in range(n):
        s = s + xi[k] * xi[k]
    rt = np.sqrt(s)
    d = rt - f32(R)
    zero = [f32(0)] * n
    if d * d > fastd2:
        return d, zero, zero
    g = [xi[k] / rt for k in range(n)]
    if any(np.isnan(v) for v in g):
        return d, zero, zero
    m = f32(0)
    for k in range(n):
        m = m + g[k] * g[k]
    m = np.sqrt(m)
    return d / m, [v / m for v in g], [f32(v) for v in vel]


def measure_flow(fl, center, R, vel, t, eps, exitBC, perdir):
    """`measure!(a,body;t,ϵ)` for the translating sphere."""
    measure_flow_fn(fl, lambda x, fd2: measure_sphere(center, R, vel, x, t, fd2),
                    lambda x, fd2: measure_sphere(center, R, vel, x, t, f32(0))[0], eps, exitBC, perdir)


def measure_flow_fn(fl, meas, sdf, eps, exitBC, perdir):
    """`measure!(a,body;t,ϵ)`: fills V, μ₀, μ₁, σ (Julia-order arrays in `fl`), then the two BC! calls.
    `meas(x,fastd²)` is `measure(body,x,t;fastd²)`, `sdf(x,fastd²)` is `sdf(body,x,t;fastd²)` of measure_sdf!."""
    V, m0, m1, sg = fl["V"], fl["mu0"], fl["mu1"], fl["sigma"]
    n = V.shape[-1]
    V[...] = 0; m0[...] = 1; m1[...] = 0
    eps = f32(eps)
    d2 = (f32(2) + eps) * (f32(2) + eps)
    Vv, M0, M1, S = J(V), J(m0), J(m1), J(sg)
    loc = lambda i, I: [f32(I[k] - 1.5) - (f32(0.5) if k + 1 == i else f32(0)) for k in range(n)]   # core.jl:177
    for I in inside(sg):
        S[I] = sdf(loc(0, I), d2)                                                # measure_sdf!(a.σ,body,t;fastd²=d²)
    for I in inside(sg):
        if S[I] * S[I] < d2:
            for i in range(1, n + 1):
                di, ni, Vi = meas(loc(i, I), d2)
                di = di if abs(di) <= 0.5 else np.copysign(di, S[I])
                Vv[I + (i,)] = Vi[i - 1]
                M0[I + (i,)] = mu0(di, eps)
                for j in range(1, n + 1):
                    M1[I + (i, j)] = mu1(di, eps) * ni[j - 1]
        elif S[I] < 0:
            for i in range(1, n + 1):
                M0[I + (i,)] = 0
    BC(m0, (0.0,) * n, False, perdir)
    BC(V, (0.0,) * n, exitBC, perdir)


# ---- RigidMap.jl + SetBody (Body.jl:91-107), written from the Julia source with 1-based matrices ----------
def rotation(theta, n):
    """`rotation(θ)` RigidMap.jl:45-49 as a dict {(i,j): R̂[i,j]} (1-based)."""
    c, s_ = (lambda a: f32(np.cos(np.float64(a)))), (lambda a: f32(np.sin(np.float64(a))))
    if n == 2:
        th = f32(theta)
        return {(1, 1): c(th), (1, 2): s_(th), (2, 1): -s_(th), (2, 2): c(th)}
    t1, t2, t3 = (f32(v) for v in theta)
    return {(1, 1): c(t3) * c(t2), (1, 2): c(t3) * s_(t2) * s_(t1) + s_(t3) * c(t1), (1, 3): -c(t3) * s_(t2) * c(t1) + s_(t3) * s_(t1),
            (2, 1): -s_(t3) * c(t2), (2, 2): -s_(t3) * s_(t2) * s_(t1) + c(t3) * c(t1), (2, 3): s_(t3) * s_(t2) * c(t1) + c(t3) * s_(t1),
            (3, 1): s_(t2), (3, 2): -c(t2) * s_(t1), (3, 3): c(t2) * c(t1)}


def cross(w, b):
    if len(b) == 2:                                   # ×(a::Number,b) = a*SA[-b[2],b[1]]
        return [f32(w) * -b[1], f32(w) * b[0]]
    return [f32(w[1]) * b[2] - f32(w[2]) * b[1], f32(w[2]) * b[0] - f32(w[0]) * b[2], f32(w[0]) * b[1] - f32(w[1]) * b[0]]


def lib_sdf(kind, radius, minor, axis, xi):
    """The library sdfs and their body-frame gradients (what ForwardDiff returns away from kinks)."""
    n = len(xi)
    if kind == "capsule":
        y = list(xi)
        y[axis] = xi[axis] - max(-f32(minor), min(xi[axis], f32(minor)))
        s = f32(0)
        for v in y:
            s = s + v * v
        ln = np.sqrt(s)
        return ln - f32(radius), [v / ln for v in y]
    if kind in ("sphere", "circle", "cylinder"):
        s = f32(0)
        for k in range(n):
            if kind != "cylinder" or k != axis:
                s = s + xi[k] * xi[k]
        rt = np.sqrt(s)
        return rt - f32(radius), [f32(0) if (kind == "cylinder" and k == axis) else xi[k] / rt for k in range(n)]
    s = f32(0)                                            # torus
    for k in range(n):
        if k != axis:
            s = s + xi[k] * xi[k]
    rho = np.sqrt(s); q = rho - f32(radius)
    ln = np.sqrt(q * q + xi[axis] * xi[axis])
    return ln - f32(minor), [xi[k] / ln if k == axis else (q / ln) * (xi[k] / rho) for k in range(n)]


def measure_rigid(leaf, x, fastd2):
    """`measure(AutoBody(sdf,RigidMap),x,t;fastd²)` AutoBody.jl:29-37 with RigidMap.jl:38-41."""
    n = len(x)
    with np.errstate(all="ignore"):
        R = rotation(leaf["theta"], n)
        x0, xp = [f32(v) for v in leaf["x0"]], [f32(v) for v in leaf.get("xp") or (0,) * n]
        r = [f32(x[k]) - x0[k] - xp[k] for k in range(n)]
        xi = []
        for i in range(1, n + 1):                         # m.R̂*(x-m.x₀-m.xₚ)+m.xₚ
            a = R[i, 1] * r[0]
            for j in range(2, n + 1):
                a = a + R[i, j] * r[j - 1]
            xi.append(a + xp[i - 1])
        d, g = lib_sdf(leaf["kind"], leaf.get("radius", 0), leaf.get("minor", 0), leaf.get("axis", 2), xi)
        zero = [f32(0)] * n
        if d * d > fastd2 or any(np.isnan(v) for v in g):
            return d, zero, zero
        nn = []
        for j in range(1, n + 1):                         # n = J'n
            a = R[1, j] * g[0]
            for i in range(2, n + 1):
                a = a + R[i, j] * g[i - 1]
            nn.append(a)
        m = f32(0)
        for v in nn:
            m = m + v * v
        m = np.sqrt(m)
        Vc = [f32(v) for v in leaf.get("V") or (0,) * n]
        w = leaf.get("omega", 0.0)
        if n == 3 and np.isscalar(w):
            w = (0.0, 0.0, w)
        cr = cross(w, r)                                   # -J\∂ₜmap = V + ω×(x-x₀-xₚ) in exact arithmetic
        return d / m, [v / m for v in nn], [Vc[k] + cr[k] for k in range(n)]


def _isless(a, b):
    return (not np.isnan(a)) and (np.isnan(b) or (np.signbit(a) and not np.signbit(b) and a == b) or a < b)


def _tuple_less(A, B):
    """`isless` on (d, n::SVector, V::SVector) tuples: lexicographic."""
    fa, fb = [A[0]] + list(A[1]) + list(A[2]), [B[0]] + list(B[1]) + list(B[2])
    for a, b in zip(fa, fb):
        if _isless(a, b):
            return True
        if _isless(b, a):
            return False
    return False


def measure_set(body, x, fastd2):
    """`measure(body::SetBody,...)` Body.jl:104-107 on a nested tuple: ("leaf", {...}) | ("min"|"max", a, b) | ("neg", a)."""
    if body[0] == "leaf":
        return measure_rigid(body[1], x, fastd2)
    if body[0] == "neg":
        d, nn, V = measure_set(body[1], x, fastd2)
        return -d, [-v for v in nn], V
    a, b = measure_set(body[1], x, fastd2), measure_set(body[2], x, fastd2)
    if body[0] == "min":
        return b if _tuple_less(b, a) else a
    return b if _tuple_less(a, b) else a


def sdf_set(body, x, fastd2):
    """`sdf(body,x,t;fastd²)`: AutoBody.jl:19 (raw) for a lone AutoBody, Body.jl:66 (`measure(...)[1]`) for a SetBody."""
    if body[0] == "leaf":
        return measure_rigid(body[1], x, f32(0))[0]
    return measure_set(body, x, fastd2)[0]


def pressure_force(p, center, R, vel, t):
    n = p.ndim
    P = J(p)
    out = np.zeros(n, np.float64)
    for I in inside(p):
        x = [f32(I[k] - 1.5) for k in range(n)]
        d, nn, _ = measure_sphere(center, R, vel, x, t, f32(1))
        kq = kern(max(f32(-1), min(d, f32(1))))
        for c in range(n):
            out[c] += np.float64(P[I] * (nn[c] * kq))
    return out


def viscous_force(u, nu, center, R, vel, t):
    """`viscous_force(u,ν,df,body,t)` Metrics.jl:140-152 with `∂(i,j,I,u)` :42-45 and `S(I,u)` :134."""
    N, n = u.shape[:-1], u.shape[-1]
    Uu = J(u)

    def dudx(i, j, I):
        ei, ej = d_(i, n), d_(j, n)
        if i == j:
            return Uu[plus(I, ei) + (i,)] - Uu[I + (i,)]
        return (Uu[plus(I, ej) + (i,)] + Uu[plus(plus(I, ej), ei) + (i,)]
                - Uu[plus(I, ej, -1) + (i,)] - Uu[plus(plus(I, ej, -1), ei) + (i,)]) / f32(4)

    m2nu = f32(-2) * f32(nu)
    out = np.zeros(n, np.float64)
    for I in CR(*[(2, k - 1) for k in N]):
        x = [f32(I[k] - 1.5) for k in range(n)]
        d, nn, _ = measure_sphere(center, R, vel, x, t, f32(1))
        kq = kern(max(f32(-1), min(d, f32(1))))
        nds = [nn[c] * kq for c in range(n)]
        for a in range(1, n + 1):
            acc = f32(0)
            for c in range(1, n + 1):
                S = (dudx(a, c, I) + dudx(c, a, I)) / f32(2)
                acc = acc + (m2nu * S) * nds[c - 1]
            out[a - 1] += np.float64(acc)
    return out


def force_moment(p, u, nu, x0, center, R, vel, t, viscous):
    """`pressure_moment` / `viscous_moment` (Metrics.jl:168-192), x₀ in Float32."""
    N, n = u.shape[:-1], u.shape[-1]
    Uu, P = J(u), J(p)

    def dudx(i, j, I):
        ei, ej = d_(i, n), d_(j, n)
        if i == j:
            return Uu[plus(I, ei) + (i,)] - Uu[I + (i,)]
        return (Uu[plus(I, ej) + (i,)] + Uu[plus(plus(I, ej), ei) + (i,)]
                - Uu[plus(I, ej, -1) + (i,)] - Uu[plus(plus(I, ej, -1), ei) + (i,)]) / f32(4)

    def cross(a, b):
        if n == 3:
            return [a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]]
        c = a[0] * b[1] - a[1] * b[0]
        return [c, c]

    out = np.zeros(n, np.float64)
    for I in CR(*[(2, k - 1) for k in N]):
        x = [f32(I[k] - 1.5) for k in range(n)]
        d, nn, _ = measure_sphere(center, R, vel, x, t, f32(1))
        kq = kern(max(f32(-1), min(d, f32(1))))
        v = [nn[c] * kq for c in range(n)]
        r = [x[c] - f32(x0[c]) for c in range(n)]
        if viscous:
            w = []
            for a in range(1, n + 1):
                acc = f32(0)
                for c in range(1, n + 1):
                    acc = acc + ((dudx(a, c, I) + dudx(c, a, I)) / f32(2)) * v[c - 1]
                w.append(acc)
            df = [f32(-2) * f32(nu) * q for q in cross(r, w)]
        else:
            df = [P[I] * q for q in cross(r, v)]
        for c in range(n):
            out[c] += np.float64(df[c])
    return out


def mean_update(P, U, UU, p, u, eps):
    """`update!(meanflow,flow)` Metrics.jl:239-249 on Julia-order arrays (UU may be None)."""
    eps = f32(eps)
    om = f32(1) - eps
    P[...] = eps * p + om * P
    U[...] = eps * u + om * U
    if UU is not None:
        n = u.shape[-1]
        for i in range(n):
            for j in range(n):
                UU[..., i, j] = eps * (u[..., i] * u[..., j]) + om * UU[..., i, j]
