"""ctypes wrapper around the CPU oracle (oracle/wl_oracle.cpp).

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may
import this module; the product package (waterlily.jl_b200/) never does.

The class mirrors the reference's `Simulation` / `Flow` / `MultiLevelPoisson` surface
(src/WaterLily.jl:86-149) closely enough that the parity tests read like the reference's own.
Arrays are exposed as numpy views in Julia layout: a Julia array of size (n1,n2,n3,D) is the
numpy array of shape (D,n3,n2,n1) (C order), i.e. `a[i-1, k-1, j-1, ii-1] == a_julia[ii,j,k,i]`.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

FIELDS = dict(u=0, u0=1, f=2, p=3, sigma=4, V=5, mu0=6, mu1=7,
              x=10, r=11, eps=12, z=13, L=14, D=15, iD=16)
SCHEMES = dict(quick=0, cds=1, vanLeer=2)


class Body(C.Structure):
    """Parametric AutoBody: sdf of a primitive composed with map(x,t) = x - centre(t),
    centre(t) = center + vel t + acc t^2 + amp sin(omega t).  (src/AutoBody.jl:1-37)"""
    _fields_ = [("kind", C.c_int), ("axis", C.c_int), ("center", C.c_float * 3),
                ("radius", C.c_float), ("minor", C.c_float), ("vel", C.c_float * 3),
                ("acc", C.c_float * 3), ("amp", C.c_float * 3), ("omega", C.c_float)]


KINDS = dict(none=0, sphere=1, circle=1, cylinder=2, torus=3)


def make_body(kind="none", center=(0, 0, 0), radius=0.0, minor=0.0, axis=2, vel=(0, 0, 0),
              acc=(0, 0, 0), amp=(0, 0, 0), omega=0.0):
    def v3(a):
        a = list(a) + [0.0] * (3 - len(a))
        return (C.c_float * 3)(*[float(x) for x in a])
    return Body(KINDS[kind], axis, v3(center), float(radius), float(minor), v3(vel), v3(acc),
                v3(amp), float(omega))


class Node(C.Structure):
    """One entry of a postfix SetBody program over `AutoBody(sdf, RigidMap(x₀,θ;xₚ,V,ω))` leaves
    (src/Body.jl:91-107, src/RigidMap.jl:28-50); layout of `wlo_node`."""
    _fields_ = [("op", C.c_int), ("kind", C.c_int), ("axis", C.c_int), ("radius", C.c_float), ("minor", C.c_float),
                ("x0", C.c_float * 3), ("theta", C.c_float * 3), ("xp", C.c_float * 3), ("V", C.c_float * 3),
                ("omega", C.c_float * 3)]


LEAF_KINDS = dict(KINDS, capsule=4)


class CsgBody:
    """`AutoBody(sdf, RigidMap(...))` leaves combined with `+ ∪ ∩ −` (lazy `SetBody`, Body.jl:91-101) as a postfix list."""

    def __init__(self, prog):
        self.prog = [dict(p) for p in prog]

    @classmethod
    def leaf(cls, kind, x0, theta=0.0, *, xp=None, V=None, omega=0.0, radius=0.0, minor=0.0, axis=2):
        return cls([dict(op=0, kind=LEAF_KINDS[kind], axis=axis, radius=radius, minor=minor, x0=x0, theta=theta,
                         xp=xp, V=V, omega=omega)])

    kind = property(lambda self: 1)                      # "not NoBody" for sim_step's remeasure test
    __add__ = __or__ = lambda a, b: CsgBody(a.prog + b.prog + [dict(op=1)])
    __and__ = lambda a, b: CsgBody(a.prog + b.prog + [dict(op=2)])
    __neg__ = lambda a: CsgBody(a.prog + [dict(op=3)])
    __sub__ = lambda a, b: a & (-b)

    def setmap(self, **kw):
        """`setmap(body; x₀, θ, xₚ, V, ω)` (RigidMap.jl:54-56): every leaf gets the new map fields."""
        names = {"x0": "x0", "x₀": "x0", "theta": "theta", "θ": "theta", "xp": "xp", "xₚ": "xp", "V": "V", "omega": "omega", "ω": "omega"}
        out = CsgBody(self.prog)
        for p in out.prog:
            if p["op"] == 0:
                for k, v in kw.items():
                    p[names[k]] = v
        return out

    def nodes(self):
        def v3(a, rot=False):
            if a is None:
                a = ()
            if np.isscalar(a):
                a = (0.0, 0.0, a) if rot else (a,)       # 2D: θ and ω are rotations about z
            a = [float(x) for x in a]
            return (C.c_float * 3)(*(a + [0.0] * (3 - len(a))))
        arr = (Node * len(self.prog))()
        for q, p in zip(arr, self.prog):
            q.op = p["op"]
            if p["op"] == 0:
                q.kind, q.axis, q.radius, q.minor = p["kind"], p["axis"], p["radius"], p["minor"]
                q.x0, q.theta, q.xp, q.V, q.omega = v3(p["x0"]), v3(p["theta"], True), v3(p["xp"]), v3(p["V"]), v3(p["omega"], True)
        return arr, len(self.prog)


def measure_point_csg(body, x, fastd2=np.inf, dtype=np.float32):
    """`measure(body,x,t;fastd²)` for a CsgBody at one point -> (d, n, V)."""
    lib = load(dtype)
    D = len(x)
    xx = (lib.real * 3)(*(list(map(float, x)) + [0.0] * (3 - D)))
    d, n, V = lib.real(), (lib.real * 3)(), (lib.real * 3)()
    arr, k = body.nodes()
    lib.wlo_measure_point_csg(arr, k, D, xx, float(fastd2), C.byref(d), n, V)
    return d.value, np.array(n[:D]), np.array(V[:D])


class Config(C.Structure):
    _fields_ = [("D", C.c_int), ("N", C.c_int * 3), ("perdir", C.c_int * 3), ("exitBC", C.c_int),
                ("scheme", C.c_int), ("nu", C.c_float), ("dt0", C.c_float), ("uBC", C.c_float * 3),
                ("maxlevels", C.c_int), ("tol", C.c_float), ("itmx", C.c_int)]


def build(force=False):
    """Compile the oracle libraries with the committed Makefile (g++ is in the image)."""
    need = force or not all(os.path.exists(os.path.join(_HERE, f))
                            for f in ("liboracle_f32.so", "liboracle_f64.so"))
    if not need:
        src = os.path.getmtime(os.path.join(_HERE, "wl_oracle.cpp"))
        need = any(os.path.getmtime(os.path.join(_HERE, f)) < src
                   for f in ("liboracle_f32.so", "liboracle_f64.so"))
    if need:
        subprocess.run(["make", "-C", _HERE, "-s", "-B", "all"], check=True)


_LIBS = {}


def load(dtype=np.float32, fast_path=None):
    key = fast_path or np.dtype(dtype).name
    if key in _LIBS:
        return _LIBS[key]
    if fast_path is None:
        build()
        path = os.path.join(_HERE, "liboracle_f32.so" if np.dtype(dtype) == np.float32 else "liboracle_f64.so")
    else:
        path = fast_path
    lib = C.CDLL(path)
    R = C.c_float if lib.wlo_sizeof_real() == 4 else C.c_double
    RP = C.POINTER(R)
    lib.real = R
    lib.wlo_create.restype = C.c_void_p
    lib.wlo_create.argtypes = [C.POINTER(Config)]
    lib.wlo_destroy.argtypes = [C.c_void_p]
    lib.wlo_finish_ic.argtypes = [C.c_void_p, RP]
    lib.wlo_measure.argtypes = [C.c_void_p, C.POINTER(Body), C.c_double, C.c_double]
    lib.wlo_build_poisson.argtypes = [C.c_void_p]
    lib.wlo_update_poisson.argtypes = [C.c_void_p]
    lib.wlo_field.restype = C.c_void_p
    lib.wlo_field.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_long), C.POINTER(C.c_int)]
    lib.wlo_nlevels.argtypes = [C.c_void_p]
    lib.wlo_dt.restype = R
    lib.wlo_dt.argtypes = [C.c_void_p]
    lib.wlo_set_dt.argtypes = [C.c_void_p, R]
    lib.wlo_mom_step.restype = R
    lib.wlo_mom_step.argtypes = [C.c_void_p, RP, RP, RP, C.POINTER(C.c_int)]
    lib.wlo_get_trace.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)] + [C.POINTER(C.c_float)] * 3
    lib.wlo_solver.argtypes = [C.c_void_p]
    lib.wlo_poisson_solver.argtypes = [C.c_void_p, C.c_int]
    lib.wlo_mult.argtypes = [C.c_void_p, C.c_void_p]
    lib.wlo_cfl.restype = R
    lib.wlo_cfl.argtypes = [C.c_void_p]
    lib.wlo_pressure_force.argtypes = [C.c_void_p, C.POINTER(Body), C.c_double, C.POINTER(C.c_double)]
    lib.wlo_viscous_force.argtypes = [C.c_void_p, C.POINTER(Body), C.c_double, C.POINTER(C.c_double)]
    lib.wlo_meanflow_update.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, R]
    lib.wlo_force_moment.argtypes = [C.c_void_p, C.POINTER(Body), C.c_double, C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double)]
    lib.wlo_op_conv_diff.argtypes = [C.c_void_p, C.c_int]
    lib.wlo_op_bdim.argtypes = [C.c_void_p]
    lib.wlo_op_scale_u.argtypes = [C.c_void_p, R]
    lib.wlo_op_project.argtypes = [C.c_void_p, R]
    for nm in ("residual", "jacobi", "restrict", "prolongate"):
        getattr(lib, "wlo_op_" + nm).argtypes = [C.c_void_p, C.c_int]
    lib.wlo_op_smooth.argtypes = [C.c_void_p, C.c_int, R]
    lib.wlo_op_increment.argtypes = [C.c_void_p, C.c_int, R]
    lib.wlo_op_vcycle.argtypes = [C.c_void_p, R]
    lib.wlo_op_pcg.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.wlo_op_L1.restype = R
    lib.wlo_op_L1.argtypes = [C.c_void_p, C.c_int]
    lib.wlo_op_Linf.restype = R
    lib.wlo_op_Linf.argtypes = [C.c_void_p, C.c_int]
    IP = C.POINTER(C.c_int)
    lib.wlo_BC.argtypes = [C.c_void_p, C.c_int, IP, RP, C.c_int, IP]
    lib.wlo_exitBC.argtypes = [C.c_void_p, C.c_void_p, C.c_int, IP, R]
    lib.wlo_perBC.argtypes = [C.c_void_p, C.c_int, IP, IP]
    lib.wlo_scheme.restype = R
    lib.wlo_scheme.argtypes = [C.c_int, R, R, R]
    lib.wlo_mu0.restype = R
    lib.wlo_mu0.argtypes = [R, R]
    lib.wlo_mu1.restype = R
    lib.wlo_mu1.argtypes = [R, R]
    lib.wlo_measure_point.argtypes = [C.POINTER(Body), C.c_int, RP, R, R, RP, RP, RP]
    NP, DP = C.POINTER(Node), C.POINTER(C.c_double)
    lib.wlo_measure_csg.argtypes = [C.c_void_p, NP, C.c_int, C.c_double]
    lib.wlo_pressure_force_csg.argtypes = [C.c_void_p, NP, C.c_int, DP]
    lib.wlo_viscous_force_csg.argtypes = [C.c_void_p, NP, C.c_int, DP]
    lib.wlo_force_moment_csg.argtypes = [C.c_void_p, NP, C.c_int, DP, C.c_int, DP]
    lib.wlo_measure_point_csg.argtypes = [NP, C.c_int, C.c_int, RP, R, RP, RP, RP]
    FN = C.CFUNCTYPE(R, C.c_int, RP, R)
    lib.FN = FN
    lib.wlo_mom_step_fn.restype = R
    lib.wlo_mom_step_fn.argtypes = [C.c_void_p, FN, FN, FN, C.POINTER(C.c_int)]
    lib.wlo_finish_ic_fn.argtypes = [C.c_void_p, FN]
    lib.wlo_BC_fn.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), FN, C.c_int, C.POINTER(C.c_int), R]
    lib.wlo_accelerate_fn.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), R, FN, FN]
    lib.wlo_apply_fn.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), FN, R]
    _LIBS[key] = lib
    return lib


def _i3(a):
    a = list(a) + [0] * (3 - len(a))
    return (C.c_int * 3)(*[int(x) for x in a])


class OracleSim:
    """`Simulation(dims,uBC,L;...)` restated (src/WaterLily.jl:93-106): flow ctor, measure!, pois ctor."""

    def __init__(self, dims, uBC, L=1.0, *, dt=0.25, nu=0.0, U=None, eps=1.0, perdir=(), exitBC=False,
                 scheme="quick", body=None, u0=None, dtype=np.float32, maxlevels=10, tol=2e-3,
                 itmx=32, lib=None, build_poisson=True, g=None, dUdt=None):
        """`uBC` may be a function `(i,x,t)` (then `U` is required, WaterLily.jl:98, and `dUdt(i,x,t)` is its time
        derivative, which the reference takes by ForwardDiff); `g(i,x,t)` is the acceleration field (Flow.jl:69-73)."""
        self.lib = lib or load(dtype)
        self.uBC_fn = uBC if callable(uBC) else None
        self.g_fn, self.dUdt_fn = g, dUdt
        if self.uBC_fn is not None:
            assert U is not None, "`U` must be specified if `uBC` is a Function"
            uBC = (0.0,) * len(dims)
        self.dtype = np.float32 if self.lib.real is C.c_float else np.float64
        self.D = len(dims)
        self.dims = tuple(int(n) for n in dims)
        self.L = L
        self.U = float(np.sqrt(sum(float(x) ** 2 for x in uBC))) if U is None else U
        self.eps = eps
        self.uBC = tuple(float(x) for x in uBC)
        self.body = body if body is not None else make_body("none")
        per = [1 if (d + 1) in perdir else 0 for d in range(3)]   # perdir is 1-based like Julia
        cfg = Config(self.D, _i3(self.dims), _i3(per), int(exitBC), SCHEMES[scheme], float(nu), float(dt),
                     (C.c_float * 3)(*(list(self.uBC) + [0.0] * (3 - self.D))), maxlevels, float(tol), itmx)
        self.cfg = cfg
        self.h = C.c_void_p(self.lib.wlo_create(C.byref(cfg)))
        self.dts = [self.dtype(dt)]
        self.n = []
        if u0 is not None:
            self.apply_ic(u0)
        elif self.uBC_fn is not None:                                  # ic_function(uBC) = (i,x)->uBC(i,x,0)
            u = self.field("u")
            ng = (C.c_int * 3)(*([n + 2 for n in self.dims] + [1] * (3 - self.D)))
            self.lib.wlo_apply_fn(u.ctypes.data, self.D, ng, self._cb(self.uBC_fn), 0.0)
            self.lib.wlo_finish_ic_fn(self.h, self._cb(self.uBC_fn))
        self._measure(0.0)
        self.nlevels = 0
        if build_poisson:
            self.nlevels = self.lib.wlo_build_poisson(self.h)
            if self.nlevels < 0:
                raise AssertionError("MultiLevelPoisson requires size=a2ⁿ, where n>2")

    def __del__(self):
        try:
            self.lib.wlo_destroy(self.h)
        except Exception:
            pass

    # ---- data surface
    def field(self, name, level=0):
        cnt = C.c_long()
        dims = (C.c_int * 3)()
        ptr = self.lib.wlo_field(self.h, FIELDS[name], level, C.byref(cnt), dims)
        if not ptr:
            raise IndexError(name)
        n = [dims[d] for d in range(self.D)]
        buf = (self.lib.real * cnt.value).from_address(ptr)
        a = np.frombuffer(buf, dtype=self.dtype)
        per = cnt.value // int(np.prod(n))
        shape = ([self.D, self.D] if per == self.D * self.D and per > 1 and name == "mu1" else
                 [per] if per > 1 else []) + n[::-1]
        return a.reshape(shape)

    def loc(self, i):
        """Coordinates `loc(i,I)` (core.jl:177) of component i (0 = cell centre) for every cell,
        returned as D arrays broadcastable against a scalar field view."""
        n = [d + 2 for d in self.dims]
        out = []
        for d in range(self.D):
            x = np.arange(n[d], dtype=self.dtype) - self.dtype(0.5) - (self.dtype(0.5) if i == d + 1 else 0)
            shp = [1] * self.D
            shp[self.D - 1 - d] = n[d]
            out.append(x.reshape(shp))
        return out

    def apply_ic(self, fn, U=None):
        """`apply!(u0,u)` over ALL indices, then BC!, exitBC!(u,u,0), u⁰=u (Flow.jl:140-142)."""
        u = self.field("u")
        for i in range(self.D):
            u[i][...] = np.broadcast_to(np.asarray(fn(i + 1, self.loc(i + 1)), dtype=self.dtype), u[i].shape)
        if self.uBC_fn is not None and U is None:
            self.lib.wlo_finish_ic_fn(self.h, self._cb(self.uBC_fn))
            return
        U = self.uBC if U is None else U
        arr = (self.lib.real * 3)(*(list(U) + [0.0] * (3 - self.D)))
        self.lib.wlo_finish_ic(self.h, arr)

    def _cb(self, fn):
        """ctypes thunk for `f(i,x,t)`; None -> null pointer.  Kept alive on the instance."""
        if fn is None:
            return self.lib.FN()
        cache = self.__dict__.setdefault("_cbs", {})
        if fn not in cache:
            D, T = self.D, self.dtype
            cache[fn] = self.lib.FN(lambda i, x, t: float(fn(i, [T(x[k]) for k in range(D)], T(t))))
        return cache[fn]

    # ---- time stepping
    def time(self):
        return float(np.sum(np.asarray(self.dts[:-1], dtype=self.dtype), dtype=self.dtype)) if len(self.dts) > 1 else 0.0

    def sim_time(self):
        return self.time() * self.U / self.L

    def measure(self, t=None):
        """`measure!(sim,t=sum(Δt))` (WaterLily.jl:146-149)."""
        if t is None:
            t = float(np.sum(np.asarray(self.dts, dtype=self.dtype), dtype=self.dtype))
        self._measure(t)
        self.lib.wlo_update_poisson(self.h)

    def _measure(self, t):
        if isinstance(self.body, CsgBody):
            arr, n = self.body.nodes()
            self.lib.wlo_measure_csg(self.h, arr, n, float(self.eps))
        else:
            self.lib.wlo_measure(self.h, C.byref(self.body), float(t), float(self.eps))

    def mom_step(self, U1=None, acc0=None, acc1=None):
        def arr(v):
            if v is None:
                return None
            return (self.lib.real * 3)(*(list(v) + [0.0] * (3 - len(v))))
        n = (C.c_int * 2)()
        if self.uBC_fn is not None or self.g_fn is not None:
            assert U1 is None and acc0 is None and acc1 is None
            dt = self.lib.wlo_mom_step_fn(self.h, self._cb(self.uBC_fn), self._cb(self.g_fn), self._cb(self.dUdt_fn), n)
        else:
            dt = self.lib.wlo_mom_step(self.h, arr(U1), arr(acc0), arr(acc1), n)
        self.dts.append(self.dtype(dt))
        self.n += [n[0], n[1]]
        return dt

    def sim_step(self, t_end=None, remeasure=True, max_steps=None, uBC_t=None, dUdt=None):
        """`sim_step!(sim[,t_end];remeasure)` (WaterLily.jl:128-139).  `uBC_t(t)`/`dUdt(t)` give a
        space-uniform time-dependent boundary velocity and its derivative (Function uBC)."""
        steps = 0
        while True:
            if t_end is not None and not (self.sim_time() < t_end):
                break
            if max_steps is not None and steps >= max_steps:
                break
            if remeasure and self.body.kind != 0:
                self.measure()
            t1 = float(np.sum(np.asarray(self.dts, dtype=self.dtype), dtype=self.dtype))
            t0 = float(self.dtype(t1) - self.dts[-1])
            self.mom_step(None if uBC_t is None else uBC_t(t1),
                          None if dUdt is None else dUdt(t0), None if dUdt is None else dUdt(t1))
            steps += 1
            if t_end is None:
                break

    def trace(self, which):
        n = C.c_int()
        a, b, c = ((C.c_float * 64)() for _ in range(3))
        self.lib.wlo_get_trace(self.h, which, C.byref(n), a, b, c)
        k = n.value + 1
        return dict(n=n.value, rinf=list(a[:k]), r1=list(b[:k]), omega=list(c[:k]))

    def pressure_force(self, t=None):
        out = (C.c_double * 3)()
        t = self.time() if t is None else t
        if isinstance(self.body, CsgBody):
            arr, n = self.body.nodes()
            self.lib.wlo_pressure_force_csg(self.h, arr, n, out)
            return np.array(out[: self.D])
        self.lib.wlo_pressure_force(self.h, C.byref(self.body), float(t), out)
        return np.array(out[: self.D])

    def viscous_force(self, t=None):
        out = (C.c_double * 3)()
        t = self.time() if t is None else t
        if isinstance(self.body, CsgBody):
            arr, n = self.body.nodes()
            self.lib.wlo_viscous_force_csg(self.h, arr, n, out)
            return np.array(out[: self.D])
        self.lib.wlo_viscous_force(self.h, C.byref(self.body), float(t), out)
        return np.array(out[: self.D])

    def moment(self, x0, viscous=False, t=None):
        """`pressure_moment(x₀,sim)` / `viscous_moment(x₀,sim)` (Metrics.jl:168-192)."""
        out = (C.c_double * 3)()
        t = self.time() if t is None else t
        xx = (C.c_double * 3)(*(list(x0) + [0.0] * (3 - len(x0))))
        if isinstance(self.body, CsgBody):
            arr, n = self.body.nodes()
            self.lib.wlo_force_moment_csg(self.h, arr, n, xx, int(viscous), out)
            return np.array(out[: self.D])
        self.lib.wlo_force_moment(self.h, C.byref(self.body), float(t), xx, int(viscous), out)
        return np.array(out[: self.D])

    def cfl(self):
        return float(self.lib.wlo_cfl(self.h))


class MeanFlow:
    """`MeanFlow(flow; t_init, uu_stats)` + `update!`/`reset!`/`uu`/`copy!` (src/Metrics.jl:205-263) on an OracleSim."""

    def __init__(self, sim, t_init=None, uu_stats=False):
        self.sim = sim
        T = sim.dtype
        self.P = np.zeros_like(sim.field("p"))
        self.U = np.zeros_like(sim.field("u"))
        self.UU = np.zeros((sim.D, sim.D) + self.P.shape, T) if uu_stats else None
        self.t = [T(sim.time() if t_init is None else t_init)]

    def time(self):
        return self.t[-1] - self.t[0]

    def update(self):
        T = self.sim.dtype
        dt = T(self.sim.time()) - self.t[-1]
        eps = dt / (dt + self.time() + np.finfo(T).eps)
        if len(self.t) == 1:
            eps = T(1)
        uu = self.UU.ctypes.data if self.UU is not None else None
        self.sim.lib.wlo_meanflow_update(self.sim.h, self.P.ctypes.data, self.U.ctypes.data, uu, float(eps))
        self.t.append(self.t[-1] + dt)

    def uu(self):
        return self.UU - self.U[None, :] * self.U[:, None]
