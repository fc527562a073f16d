"""waterlily.jl_b200 -- host-side mirror of WaterLily.jl's `Simulation` / `sim_step!` API over the
C ABI of libwaterlily_b200.so (include/wl_b200.h).

The reference is Julia; Julia is not present in this image, so this module plays the role of
the thin `ccall` shim (julia/WaterLilyB200.jl is that shim, unexecuted here).  Names and
argument meaning follow the reference (src/WaterLily.jl:93-149, src/Flow.jl:133-147):

    sim = Simulation((nx,ny,nz), (U,0,0), L; ν, body=AutoBody(...), perdir, exitBC, ...)
    sim_step!(sim, t_end; remeasure)   ->  sim.sim_step(t_end, remeasure=...)
    sim.flow.u, sim.flow.p, sim.flow.Δt, sim.pois.n, pressure_force(sim)

Everything numerical happens in hand-written sm_100a CUDA kernels inside the shared library;
this file only marshals arguments.  There is NO CPU fallback: if the library is missing or no
Blackwell GPU is visible, construction raises.

Array convention: a Julia array of size (n1,n2,n3,D) is returned as a numpy array of shape
(D,n3,n2,n1) (same memory order), i.e. a[c-1,k-1,j-1,i-1] == a_julia[i,j,k,c].
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBNAME = os.path.join(_HERE, "libwaterlily_b200.so")

WL_OK, WL_ERR_INVALID, WL_ERR_UNSUPPORTED, WL_ERR_CUDA, WL_ERR_NO_DEVICE = range(5)
FIELDS = dict(u=0, u0=1, f=2, p=3, sigma=4, V=5, mu0=6, mu1=7,
              x=10, r=11, eps=12, z=13, L=14, D=15, iD=16, mean_P=20, mean_U=21, mean_UU=22, accel_t0=23, accel_t1=24)
SCHEMES = dict(quick=0, cds=1, vanLeer=2)
KINDS = dict(none=0, sphere=1, circle=1, cylinder=2, torus=3)

# symbols include/wl_b200.h declares (tests check the library exports every one of them)
ABI_SYMBOLS = [
    "wl_create", "wl_destroy", "wl_last_error", "wl_version", "wl_field_dims", "wl_download",
    "wl_upload", "wl_nlevels", "wl_finish_ic", "wl_bc_table_size", "wl_set_bc_tables", "wl_set_accel_field", "wl_measure_body", "wl_set_body_coeffs",
    "wl_build_poisson", "wl_update_poisson", "wl_solver", "wl_poisson_solver", "wl_mom_step", "wl_get_dt", "wl_set_dt",
    "wl_cfl", "wl_cfl_noalloc", "wl_tile_classes", "wl_pressure_force", "wl_set_force_weights", "wl_pressure_force_weights", "wl_viscous_force", "wl_force_moment", "wl_measure_csg", "wl_pressure_force_csg",
    "wl_viscous_force_csg", "wl_force_moment_csg", "wl_meanflow_init", "wl_meanflow_update", "wl_meanflow_copy_to_flow", "wl_sync", "wl_launch_count", "wl_enable_timing",
    "wl_last_phase_times", "wl_debug_op", "wl_prof_begin", "wl_prof_end", "wl_event_record",
    "wl_event_elapsed_ms", "wl_create_slab", "wl_nccl_unique_id", "wl_slab_info", "wl_halo_stats", "wl_halo_transport", "wl_halo_timing",
]


class WLConfig(C.Structure):
    _fields_ = [("D", C.c_int32), ("N", C.c_int32 * 3), ("perdir", C.c_int32 * 3), ("exitBC", C.c_int32),
                ("scheme", C.c_int32), ("nu", C.c_float), ("dt0", C.c_float), ("uBC", C.c_float * 3),
                ("maxlevels", C.c_int32), ("tol", C.c_float), ("itmx", C.c_int32), ("device", C.c_int32),
                ("flags", C.c_int32)]


class WLDist(C.Structure):
    _fields_ = [("rank", C.c_int32), ("nranks", C.c_int32), ("nccl_id", C.c_uint8 * 128)]


class WLBody(C.Structure):
    _fields_ = [("kind", C.c_int32), ("axis", C.c_int32), ("center", C.c_float * 3), ("radius", C.c_float),
                ("minor", C.c_float), ("vel", C.c_float * 3), ("acc", C.c_float * 3), ("amp", C.c_float * 3),
                ("omega", C.c_float)]


class WLNode(C.Structure):
    """`wl_body_node` (include/wl_b200.h)."""
    _fields_ = [("op", C.c_int32), ("kind", C.c_int32), ("axis", C.c_int32), ("radius", C.c_float), ("minor", C.c_float),
                ("x0", C.c_float * 3), ("theta", C.c_float * 3), ("xp", C.c_float * 3), ("V", C.c_float * 3),
                ("omega", C.c_float * 3)]


class WLStepInfo(C.Structure):
    _fields_ = [("dt_next", C.c_float), ("n_pred", C.c_int32), ("n_corr", C.c_int32), ("flags", C.c_int32),
                ("trace_len", C.c_int32 * 2), ("rinf", (C.c_float * 34) * 2), ("r1", (C.c_float * 34) * 2),
                ("omega", (C.c_float * 34) * 2)]


class WaterLilyB200Error(RuntimeError):
    pass


_lib = None


def load_library(path=None):
    """dlopen libwaterlily_b200.so and declare the prototypes of include/wl_b200.h.
    Loading needs no GPU; any compute entry point does."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or os.environ.get("WL_B200_LIB") or _LIBNAME   # WL_B200_LIB: an alternative build of the same ABI
    if not os.path.exists(path):
        raise WaterLilyB200Error(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU fallback.")
    lib = C.CDLL(path)
    P, F, I = C.c_void_p, C.c_float, C.c_int32
    FP = C.POINTER(F)
    lib.wl_create.argtypes = [C.POINTER(WLConfig), C.POINTER(P)]
    lib.wl_create_slab.argtypes = [C.POINTER(WLConfig), C.POINTER(WLDist), C.POINTER(P)]
    lib.wl_nccl_unique_id.argtypes = [C.c_char_p]
    lib.wl_slab_info.argtypes = [P] + [C.POINTER(I)] * 5
    lib.wl_halo_stats.argtypes = [P, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.wl_halo_transport.argtypes = [P]
    lib.wl_halo_timing.argtypes = [P, C.POINTER(C.c_double * 4)]
    lib.wl_destroy.argtypes = [P]
    lib.wl_destroy.restype = None
    lib.wl_last_error.argtypes = [P]
    lib.wl_last_error.restype = C.c_char_p
    lib.wl_field_dims.argtypes = [P, I, I, C.POINTER(I * 5), C.POINTER(I)]
    lib.wl_download.argtypes = [P, I, I, P, C.c_size_t]
    lib.wl_upload.argtypes = [P, I, I, P, C.c_size_t]
    lib.wl_nlevels.argtypes = [P]
    lib.wl_finish_ic.argtypes = [P, FP]
    lib.wl_measure_body.argtypes = [P, C.POINTER(WLBody), C.c_double, C.c_double]
    lib.wl_bc_table_size.argtypes = [P, C.POINTER(C.c_size_t)]
    lib.wl_set_bc_tables.argtypes = [P, C.c_void_p, C.c_size_t]
    lib.wl_set_accel_field.argtypes = [P, I, C.c_void_p, C.c_size_t]
    lib.wl_set_body_coeffs.argtypes = [P, P, P, P, P]
    lib.wl_build_poisson.argtypes = [P]
    lib.wl_update_poisson.argtypes = [P]
    lib.wl_solver.argtypes = [P, C.POINTER(I), C.POINTER(I), FP, FP, FP]
    lib.wl_poisson_solver.argtypes = [P, I, C.POINTER(I)]
    lib.wl_mom_step.argtypes = [P, FP, FP, FP, C.POINTER(WLStepInfo)]
    lib.wl_get_dt.argtypes = [P, FP]
    lib.wl_set_dt.argtypes = [P, F]
    lib.wl_cfl.argtypes = [P, FP]
    lib.wl_tile_classes.argtypes = [P, I, C.POINTER(C.c_int64 * 3)]
    lib.wl_cfl_noalloc.argtypes = [P, I, C.POINTER(I * 3), F, P, P]     # device pointers and a cudaStream_t: plain addresses
    lib.wl_pressure_force.argtypes = [P, C.POINTER(WLBody), C.c_double, C.POINTER(C.c_double * 3)]
    lib.wl_viscous_force.argtypes = [P, C.POINTER(WLBody), C.c_double, C.POINTER(C.c_double * 3)]
    lib.wl_set_force_weights.argtypes = [P, C.c_int64, C.c_void_p, C.c_void_p]
    lib.wl_pressure_force_weights.argtypes = [P, C.POINTER(C.c_double * 3)]
    lib.wl_force_moment.argtypes = [P, C.POINTER(WLBody), C.c_double, C.POINTER(C.c_double * 3), I, C.POINTER(C.c_double * 3)]
    D3 = C.POINTER(C.c_double * 3)
    lib.wl_measure_csg.argtypes = [P, C.POINTER(WLNode), I, C.c_double]
    lib.wl_pressure_force_csg.argtypes = [P, C.POINTER(WLNode), I, D3]
    lib.wl_viscous_force_csg.argtypes = [P, C.POINTER(WLNode), I, D3]
    lib.wl_force_moment_csg.argtypes = [P, C.POINTER(WLNode), I, D3, I, D3]
    lib.wl_meanflow_init.argtypes = [P, I]
    lib.wl_meanflow_update.argtypes = [P, F]
    lib.wl_meanflow_copy_to_flow.argtypes = [P]
    lib.wl_sync.argtypes = [P]
    lib.wl_launch_count.argtypes = [P]
    lib.wl_launch_count.restype = C.c_int64
    lib.wl_enable_timing.argtypes = [P, I]
    lib.wl_last_phase_times.argtypes = [P, C.POINTER(C.c_double * 8)]
    lib.wl_debug_op.argtypes = [P, I, I, F]
    lib.wl_prof_begin.argtypes = [P, C.c_char_p]
    lib.wl_prof_end.argtypes = [P, C.c_char_p, C.c_size_t]
    lib.wl_event_record.argtypes = [P, I]
    lib.wl_event_elapsed_ms.argtypes = [P, I, I, FP]
    if path == _LIBNAME:
        _lib = lib
    return lib


def _v3f(a):
    a = list(a) + [0.0] * (3 - len(a))
    return (C.c_float * 3)(*[float(x) for x in a])


def slab_range(rank, nranks, n_last):
    """Interior planes [k0, k0+n) of the last dimension owned by `rank` (slabs along the
    slowest index).  The split must be even: every rank gets n_last/nranks planes, an even number."""
    if n_last % nranks or (n_last // nranks) % 2 or n_last // nranks < 4:
        raise WaterLilyB200Error("WL_ERR_INVALID: the last dimension must split into an even number (>= 4) "
                                 "of planes per rank")
    n = n_last // nranks
    return rank * n, n


def slab_of(a, rank, nranks):
    """The local slab (one ghost plane per side) of a GLOBAL field array shaped ([D[,D]], n3, n2, n1)."""
    nz = a.shape[-3] - 2
    k0, n = slab_range(rank, nranks, nz)
    return np.ascontiguousarray(a[..., k0:k0 + n + 2, :, :])


def join_slabs(parts):
    """Inverse of slab_of: rank-ordered local slabs -> the global array (interface ghost planes dropped)."""
    n = len(parts)
    out = []
    for r, p in enumerate(parts):
        lo = 0 if r == 0 else 1
        hi = p.shape[-3] if r == n - 1 else p.shape[-3] - 1
        out.append(p[..., lo:hi, :, :])
    return np.concatenate(out, axis=-3)


def make_slab_comm(group=None):
    """(rank, nranks, nccl_id) for `Simulation(..., comm=...)` from an initialised torch.distributed
    process group: rank 0 creates the NCCL id inside the library, torch broadcasts the bytes."""
    import torch.distributed as dist
    rank, n = dist.get_rank(group), dist.get_world_size(group)
    box = [None]
    if rank == 0:
        buf = C.create_string_buffer(128)
        if load_library().wl_nccl_unique_id(buf) != WL_OK:
            raise WaterLilyB200Error("wl_nccl_unique_id failed")
        box[0] = buf.raw
    dist.broadcast_object_list(box, src=0, group=group)
    return rank, n, box[0]


class AutoBody:
    """Parametric stand-in for `AutoBody(sdf,map)` (src/AutoBody.jl:1-37): a primitive signed
    distance composed with the rigid translation map(x,t) = x - (center + vel t + acc t² + amp sin(ω t)).
    Arbitrary closures stay on the Julia host and enter through `set_body_coeffs`."""

    def __init__(self, kind="none", center=(0, 0, 0), radius=0.0, minor=0.0, axis=2, vel=(0, 0, 0),
                 acc=(0, 0, 0), amp=(0, 0, 0), omega=0.0):
        self.c = WLBody(KINDS[kind], int(axis), _v3f(center), float(radius), float(minor), _v3f(vel),
                        _v3f(acc), _v3f(amp), float(omega))

    @property
    def kind(self):
        return self.c.kind


LEAF_KINDS = dict(none=0, sphere=1, circle=1, cylinder=2, torus=3, capsule=4)


class RigidBody:
    """`AutoBody(sdf, RigidMap(x₀, θ; xₚ, V, ω))` (src/RigidMap.jl:28-50) for a library sdf, and the lazy `SetBody`
    algebra `+ | & -` over such leaves (src/Body.jl:91-101), held as the postfix program `wl_measure_csg` takes.
    `setmap(...)` (RigidMap.jl:54-56) returns a copy with every leaf's map fields replaced."""

    def __init__(self, kind=None, x0=None, theta=0.0, *, xp=None, V=None, omega=0.0, radius=0.0, minor=0.0, axis=2, prog=None):
        if prog is not None:
            self.prog = [dict(p) for p in prog]
        else:
            self.prog = [dict(op=0, kind=LEAF_KINDS[kind], axis=int(axis), radius=float(radius), minor=float(minor),
                              x0=x0, theta=theta, xp=xp, V=V, omega=omega)]

    kind = property(lambda self: 0 if (len(self.prog) == 1 and self.prog[0].get("kind", 1) == 0) else 1)

    def __add__(self, other):
        return RigidBody(prog=self.prog + other.prog + [dict(op=1)])
    __or__ = __add__

    def __and__(self, other):
        return RigidBody(prog=self.prog + other.prog + [dict(op=2)])

    def __neg__(self):
        return RigidBody(prog=self.prog + [dict(op=3)])

    def __sub__(self, other):
        return self & (-other)

    def setmap(self, **kw):
        names = {"x0": "x0", "x₀": "x0", "theta": "theta", "θ": "theta", "xp": "xp", "xₚ": "xp", "V": "V", "omega": "omega", "ω": "omega"}
        out = RigidBody(prog=self.prog)
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
            return _v3f([float(x) for x in a])
        arr = (WLNode * len(self.prog))()
        for q, p in zip(arr, self.prog):
            q.op = p["op"]
            if p["op"] == 0:
                q.kind, q.axis, q.radius, q.minor = p["kind"], p["axis"], p["radius"], p["minor"]
                q.x0, q.theta, q.xp, q.V, q.omega = v3(p["x0"]), v3(p["theta"], True), v3(p["xp"]), v3(p["V"]), v3(p["omega"], True)
        return arr, len(self.prog)


def NoBody():
    return AutoBody("none")


class ClosureBody:
    """Any `AbstractBody` given by host code: `measure(x, t, fastd2) -> (d, n, V)` (src/Body.jl:1-12), vectorised over
    numpy coordinate arrays (x = D broadcastable arrays; d an array, n and V sequences of D arrays, Float32; where
    d² > fastd2 a fast method may return n = V = 0, as `measure(::AutoBody)` does).  Optional `sdf(x, t)` = the distance
    alone (`sdf(body,x,t)`, src/Body.jl:66-68; default `measure(...)[0]`).  This is the route an arbitrary
    `AutoBody(sdf,map)` closure takes through the C ABI: `measure!` (src/Body.jl:28-51) and `nds`
    (src/Metrics.jl:110-113) run on the HOST (`host_measure`, `host_nds`), their results travel through
    `wl_set_body_coeffs` / `wl_set_force_weights`; everything after that is the device path.  `from_sdf` builds the
    contract from an sdf and its gradient exactly as `measure(::AutoBody)` does (src/AutoBody.jl:29-37, identity map)."""
    kind = 1

    def __init__(self, measure, sdf=None):
        self.measure = measure
        self.sdf = sdf if sdf is not None else (lambda x, t: measure(x, t, np.float32(0))[0])

    @classmethod
    def from_sdf(cls, sdf, grad, vel=None):
        def measure(x, t, fastd2=np.inf):
            d = np.asarray(sdf(x, t), np.float32) + np.float32(0) * x[0]
            far = d * d > np.float32(fastd2)                         # AutoBody.jl:31: return (d, 0, 0) un-normalised
            n = [np.asarray(g, np.float32) for g in grad(x, t)]
            bad = np.zeros(d.shape, bool)
            for g in n:
                bad = bad | np.isnan(g)                              # AutoBody.jl:33: non-differentiable points
            m = np.sqrt(sum(g * g for g in n), dtype=np.float32)
            skip = far | bad
            with np.errstate(invalid="ignore", divide="ignore"):
                d = np.where(skip, d, d / m)
                n = [np.where(skip, np.float32(0), g / m) for g in n]
            V = [np.zeros_like(d) for _ in n] if vel is None else [np.where(skip, np.float32(0), np.asarray(v, np.float32) + 0 * d) for v in vel(x, t)]
            return d, n, V
        return cls(measure, lambda x, t: np.asarray(sdf(x, t), np.float32) + np.float32(0) * x[0])


def _kern0(d):
    d64 = d.astype(np.float64)
    return ((np.float32(1) + d + (np.sin(np.pi * d64) / np.pi).astype(np.float32)) / np.float32(2)).astype(np.float32)


def _kern1(d):
    d64 = d.astype(np.float64)
    sp, cp = np.sin(np.pi * d64).astype(np.float32), np.cos(np.pi * d64).astype(np.float32)
    pi = np.float32(np.pi)
    return ((np.float32(1) - d * d) / np.float32(4) - (d * sp + (np.float32(1) + cp) / pi) / (np.float32(2) * pi)).astype(np.float32)


def host_measure(body, dims, t=0.0, eps=1.0, k_offset=0, n_last=None):
    """`measure!(flow, body; t, ϵ)` (src/Body.jl:28-48) on the host for a ClosureBody, BEFORE its two `BC!` calls (the
    library applies them): returns (μ₀, μ₁, V, σ) in the mirror's array order.  `k_offset`/`n_last`: the slab of a
    slab-decomposed sim."""
    D = len(dims)
    T = np.float32
    n = [int(v) + 2 for v in dims]
    if n_last is not None:
        n[-1] = n_last + 2
    full = tuple(n[::-1])

    def loc(i):                                        # loc(i,I) src/core.jl:177 on 0-based indices; i = 0: cell centre
        out = []
        for d in range(D):
            x = (np.arange(n[d]) + (k_offset if d == D - 1 else 0)).astype(T) - T(0.5) - (T(0.5) if i == d + 1 else T(0))
            shp = [1] * D
            shp[D - 1 - d] = n[d]
            out.append(np.broadcast_to(x.reshape(shp), full))
        return out
    eps = T(eps)
    d2 = (T(2) + eps) ** 2
    inner = tuple(slice(1, -1) for _ in range(D))
    mu0 = np.ones((D,) + full, T)
    mu1 = np.zeros((D, D) + full, T)
    V = np.zeros((D,) + full, T)
    sig = np.zeros(full, T)
    # measure_sdf! (src/Body.jl:73-75): σ = sdf at the cell centres of inside(σ), ghosts stay 0
    dc = np.asarray(body.sdf([a[inner] for a in loc(0)], T(t)), T)
    sig[inner] = dc
    band = np.zeros(full, bool)
    band[inner] = dc * dc < d2
    solid = np.zeros(full, bool)
    solid[inner] = (~(dc * dc < d2)) & (dc < 0)
    for i in range(D):
        x = [a[band] for a in loc(i + 1)]
        if x[0].size:
            di, ni, Vi = body.measure(x, T(t), d2)
            di = np.asarray(di, T)
            bad = np.zeros(di.shape, bool)
            for g in ni:
                bad |= np.isnan(np.asarray(g, T))
            ni = [np.where(bad, T(0), np.asarray(g, T)) for g in ni]
            Vi = [np.where(bad, T(0), np.asarray(v, T) + 0 * di) for v in Vi]
            di = np.where(np.abs(di) <= T(0.5), di, np.copysign(di, sig[band]))      # sign consistency, Body.jl:36
            V[i][band] = Vi[i]
            q = di / eps
            m0 = np.where(q < T(-1) + np.sqrt(np.spacing(np.abs(di)).astype(T)), T(0), _kern0(np.minimum(q, T(1))))
            mu0[i][band] = m0
            k1 = eps * _kern1(np.clip(q, T(-1), T(1)))
            for j in range(D):
                mu1[j, i][band] = k1 * ni[j]           # μ₁[I,i,j]: component i (slow-1), direction j (slowest)
        mu0[i][solid] = 0
    return mu0, mu1, V, sig


def host_nds(body, dims, t=0.0, k_offset=0, n_last=None):
    """`nds(body, loc(0,I), t)` (src/Metrics.jl:110-113) on the interior cells: (idx int64 [nnz], w float32 [nnz, D]) in
    the layout `wl_set_force_weights` takes (entries with |d| >= 1 or a zero weight are dropped)."""
    D = len(dims)
    T = np.float32
    n = [int(v) + 2 for v in dims]
    if n_last is not None:
        n[-1] = n_last + 2
    full = tuple(n[::-1])
    xs = []
    for d in range(D):
        x = (np.arange(n[d]) + (k_offset if d == D - 1 else 0)).astype(T) - T(0.5)
        shp = [1] * D
        shp[D - 1 - d] = n[d]
        xs.append(np.broadcast_to(x.reshape(shp), full))
    inner = np.zeros(full, bool)
    inner[tuple(slice(1, -1) for _ in range(D))] = True
    d, nn, _ = body.measure([a[inner] for a in xs], T(t), T(1))
    d = np.asarray(d, T)
    near = np.ones(d.shape, bool)                       # (fastd² = 1: beyond it the body's measure returns n = 0 itself)
    kern = ((T(1) + np.cos(np.pi * np.clip(d, T(-1), T(1)).astype(np.float64)).astype(T)) / T(2)).astype(T)
    w = np.stack([np.where(near & ~np.isnan(np.asarray(g, T)), np.asarray(g, T) * kern, T(0)) for g in nn], axis=-1).astype(T)
    keep = np.any(w != 0, axis=-1)
    idx = np.flatnonzero(inner.ravel())[keep].astype(np.int64)      # C order of (z,y,x) == first dimension fastest
    return idx, np.ascontiguousarray(w[keep])


class _Handle:
    def __init__(self, lib, h):
        self.lib, self.h = lib, h

    def close(self):
        if self.h:
            self.lib.wl_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class B200Flow:
    """`Flow` (src/Flow.jl:114-148) whose arrays live on the device; attribute access downloads."""

    def __init__(self, sim):
        self._sim = sim
        self.Δt = sim._dts
        self.ν = sim.nu
        self.uBC = sim.uBC
        self.exitBC = sim.exitBC
        self.perdir = sim.perdir

    def __getattr__(self, name):
        alias = {"μ₀": "mu0", "μ₁": "mu1", "σ": "sigma", "u⁰": "u0"}
        name = alias.get(name, name)
        if name in ("u", "u0", "f", "p", "sigma", "V", "mu0", "mu1"):
            return self._sim.download(name)
        raise AttributeError(name)


class _LevelView:
    def __init__(self, sim, l):
        self._sim, self._l = sim, l

    def __getattr__(self, name):
        alias = {"ϵ": "eps"}
        name = alias.get(name, name)
        if name in ("x", "r", "eps", "z", "L", "D", "iD"):
            return self._sim.download(name, self._l)
        raise AttributeError(name)


class B200Poisson:
    """`MultiLevelPoisson` (src/MultiLevelPoisson.jl:61-77): `n` iteration log and `levels[l].x/r/L/D`."""

    def __init__(self, sim):
        self._sim = sim
        self.n = sim._n

    @property
    def levels(self):
        return [_LevelView(self._sim, l) for l in range(self._sim.nlevels)]


class Simulation:
    """`Simulation(dims,uBC,L;Δt,ν,U,ϵ,perdir,u0,exitBC,λ,body,T,mem)` (src/WaterLily.jl:93-106) on a B200."""

    def __init__(self, dims, uBC, L=1.0, *, dt=0.25, nu=0.0, U=None, eps=1.0, perdir=(), exitBC=False,
                 scheme="quick", body=None, u0=None, g=None, maxlevels=10, tol=2e-3, itmx=32, device=0,
                 T=np.float32, lib=None, flags=0, comm=None, dUBCdt=None):
        if np.dtype(T) != np.float32:
            raise WaterLilyB200Error("WL_ERR_UNSUPPORTED: the B200 path is Float32 only")
        # Function uBC(i,x,t) / g(i,x,t) (src/Flow.jl:128,131): closures stay on the host and are SAMPLED into the
        # boundary tables / acceleration fields the device path consumes (set_bc_function, set_accel_function), at
        # t = 0 for the constructor's BC! and at t₁ / (t₀,t₁) before every mom_step (sim_step).  ∂ₜuBC of accelerate!
        # (the reference differentiates uBC by AD, Flow.jl:72-73) is supplied by the caller as `dUBCdt(i,x,t)`.
        bc_fn = uBC if callable(uBC) else None
        if bc_fn is not None:
            if U is None:
                raise AssertionError("U must be specified if uBC is a Function")      # src/WaterLily.jl:99
            uBC = (0.0,) * len(dims)
        if scheme not in SCHEMES:
            raise WaterLilyB200Error("WL_ERR_UNSUPPORTED: convective scheme must be quick, cds or vanLeer")
        self.lib = lib or load_library()
        self.D = len(dims)
        self.dims = tuple(int(n) for n in dims)
        self.L = L
        self.U = float(np.sqrt(sum(float(x) ** 2 for x in uBC))) if U is None else U
        self.ϵ = self.eps = eps
        self.uBC = tuple(float(x) for x in uBC)
        self.nu = float(nu)
        self.exitBC = bool(exitBC)
        self.perdir = tuple(perdir)
        self.body = body if body is not None else NoBody()
        per = [1 if (d + 1) in perdir else 0 for d in range(3)]
        cfg = WLConfig(self.D, (C.c_int32 * 3)(*(list(self.dims) + [0] * (3 - self.D))), (C.c_int32 * 3)(*per),
                       int(exitBC), SCHEMES[scheme], float(nu), float(dt), _v3f(self.uBC), maxlevels, float(tol),
                       itmx, device, int(flags))
        h = C.c_void_p()
        self.rank, self.nranks, self.k_offset = 0, 1, 0
        if comm is not None and comm[1] > 1:
            # one process per GPU, slab along the last dimension: `dims` is the GLOBAL grid
            self.rank, self.nranks = int(comm[0]), int(comm[1])
            self.k_offset, _ = slab_range(self.rank, self.nranks, self.dims[-1]) if self.D == 3 else (0, 0)
            dd = WLDist(self.rank, self.nranks, (C.c_uint8 * 128)(*bytes(comm[2])[:128]))
            rc = self.lib.wl_create_slab(C.byref(cfg), C.byref(dd), C.byref(h))
        else:
            rc = self.lib.wl_create(C.byref(cfg), C.byref(h))
        if rc != WL_OK:
            raise WaterLilyB200Error(f"wl_create failed ({rc}): {self.lib.wl_last_error(None).decode()}")
        self._handle = _Handle(self.lib, h)
        self.h = h
        self._dts = [np.float32(dt)]
        self._n = []
        self.last_info = None
        self._bc_fn = self._accel = None
        if bc_fn is not None:
            self.set_bc_function(bc_fn, 0.0)
            if u0 is None:
                u0 = lambda i, x: bc_fn(i, x, np.float32(0))                              # ic_function(uBC)
        if u0 is not None:
            self.apply_ic(u0)
        if callable(g) or dUBCdt is not None:
            self.set_accel_function(g if callable(g) else None, dUBCdt)
        self._measure(0.0)                                                                      # measure!(flow,body;ϵ)
        self._check(self.lib.wl_build_poisson(self.h))                                        # pois_ctor(flow)
        self.nlevels = self.lib.wl_nlevels(self.h)
        self.flow = B200Flow(self)
        self.pois = B200Poisson(self)

    # ---- plumbing
    def close(self):
        """`wl_destroy` now.  On a slab-decomposed run destruction is COLLECTIVE (the ranks meet once before the peer
        mappings go away): call it on every rank at the same point of the program instead of leaving it to the garbage
        collector (`sim.flow` / `sim.pois` refer back to the simulation, so `del sim` alone does not free it)."""
        self._handle.close()
        self.h = None

    def _check(self, rc):
        if rc != WL_OK:
            msg = self.lib.wl_last_error(self.h).decode()
            if rc == WL_ERR_INVALID and "size=a2" in msg:
                raise AssertionError("MultiLevelPoisson requires size=a2ⁿ, where n>2")
            raise WaterLilyB200Error(f"libwaterlily_b200 error {rc}: {msg}")

    def field_shape(self, name, level=0):
        dims = (C.c_int32 * 5)()
        nd = C.c_int32()
        self._check(self.lib.wl_field_dims(self.h, FIELDS[name], level, C.byref(dims), C.byref(nd)))
        return tuple(dims[i] for i in range(nd.value))[::-1]

    def download(self, name, level=0, out=None):
        """`Array(flow.u)`: device -> host in the reference's (unpadded) layout."""
        shape = self.field_shape(name, level)
        a = out if out is not None else np.empty(shape, np.float32)
        self._check(self.lib.wl_download(self.h, FIELDS[name], level, a.ctypes.data, a.size))
        return a

    def upload(self, name, a, level=0):
        """`copyto!(flow.u, ::Array)`."""
        a = np.ascontiguousarray(a, np.float32)
        assert a.shape == self.field_shape(name, level), (a.shape, self.field_shape(name, level))
        self._check(self.lib.wl_upload(self.h, FIELDS[name], level, a.ctypes.data, a.size))

    def loc(self, i):
        """`loc(i,I)` (src/core.jl:177) for every cell; D arrays broadcastable against a scalar field."""
        n = [d + 2 for d in self.dims]
        off = [0] * self.D
        if self.nranks > 1:
            n[-1] = self.dims[-1] // self.nranks + 2
            off[-1] = self.k_offset
        out = []
        for d in range(self.D):
            x = (np.arange(n[d]) + off[d]).astype(np.float32) - np.float32(0.5) - (np.float32(0.5) if i == d + 1 else 0)
            shp = [1] * self.D
            shp[self.D - 1 - d] = n[d]
            out.append(x.reshape(shp))
        return out

    def apply_ic(self, fn, U=None):
        """`apply!(u0,u)` over all indices + BC! + exitBC!(u,u,0) + u⁰=u (src/Flow.jl:140-142)."""
        shape = self.field_shape("u")
        u = np.empty(shape, np.float32)
        for i in range(self.D):
            u[i][...] = np.broadcast_to(np.asarray(fn(i + 1, self.loc(i + 1)), dtype=np.float32), shape[1:])
        self.upload("u", u)
        self._check(self.lib.wl_finish_ic(self.h, _v3f(self.uBC if U is None else U)))

    def set_bc_function(self, uBC, t=0.0):
        """Boundary data from a Function `uBC(i,x,t)` (src/core.jl:201-219), sampled on the host into the tables
        `wl_set_bc_tables` takes; `None` returns to the constant tuple.  The function is kept: `sim_step` re-samples
        it at t₁ = sum(Δt) before every `mom_step`, the time all four `BC!` of a step use (src/Flow.jl:194,209,231).
        Its `∂ₜuBC` term in `accelerate!` goes through `set_accel_function(dUdt=...)`."""
        self._bc_fn = uBC
        if uBC is None:
            self._check(self.lib.wl_set_bc_tables(self.h, None, 0))
            return
        from . import bctables
        kw = self._slab_kw()
        flat = bctables.pack(bctables.bc_tables(uBC, self.dims, t, self.perdir, **kw), self.dims, kw.get("n_last"))
        n = C.c_size_t()
        self._check(self.lib.wl_bc_table_size(self.h, C.byref(n)))
        assert flat.size == n.value, (flat.size, n.value)
        self._check(self.lib.wl_set_bc_tables(self.h, flat.ctypes.data, flat.size))

    def _slab_kw(self):
        """This rank's slab for the host-side samplers: {} on one GPU, else its plane offset and plane count."""
        if self.nranks > 1 and self.D == 3:
            return dict(k_offset=self.k_offset, n_last=self.dims[-1] // self.nranks)
        return {}

    def _upload_bc(self, t):
        from . import bctables
        kw = self._slab_kw()
        flat = bctables.pack(bctables.bc_tables(self._bc_fn, self.dims, t, self.perdir, **kw), self.dims, kw.get("n_last"))
        self._check(self.lib.wl_set_bc_tables(self.h, flat.ctypes.data, flat.size))

    def set_accel_function(self, g=None, dUdt=None):
        """`g(i,x,t)` and `∂ₜuBC(i,x,t)` of `accelerate!` (src/Flow.jl:69-73) as Functions: sampled on the host on every
        index at t₀ and t₁ before each `mom_step` of `sim_step` and uploaded (`wl_set_accel_field`); both None clears."""
        self._accel = None if (g is None and dUdt is None) else (g, dUdt)
        if self._accel is None:
            self._check(self.lib.wl_set_accel_field(self.h, 0, None, 0))

    def _upload_accel(self, t0, t1):
        from . import bctables
        for which, t in ((0, t0), (1, t1)):
            a = bctables.accel_field(self._accel[0], self._accel[1], self.dims, t, **self._slab_kw())
            self._check(self.lib.wl_set_accel_field(self.h, which, a.ctypes.data, a.size))

    # ---- reference API
    def time(self):
        """`time(flow)` = sum(Δt[1:end-1]) (src/Flow.jl:174)."""
        return float(np.sum(np.asarray(self._dts[:-1], np.float32), dtype=np.float32)) if len(self._dts) > 1 else 0.0

    def sim_time(self):
        return self.time() * self.U / self.L

    def set_dts(self, dts):
        """`empty!(flow.Δt); push!(flow.Δt, dts...)` (the `load!` methods): the host vector and the device-side Δt[end]."""
        self._dts[:] = [np.float32(x) for x in dts]            # in place: `sim.flow.Δt` aliases this list
        self._check(self.lib.wl_set_dt(self.h, float(self._dts[-1])))

    def measure(self, t=None):
        """`measure!(sim,t=sum(Δt))` = measure!(flow,body) + update!(pois) (src/WaterLily.jl:146-149)."""
        if t is None:
            t = float(np.sum(np.asarray(self._dts, np.float32), dtype=np.float32))
        self._measure(t)
        self._check(self.lib.wl_update_poisson(self.h))

    def _measure(self, t):
        if isinstance(self.body, ClosureBody):          # host measure! + upload (src/Body.jl:28-51)
            nl = self.dims[-1] // self.nranks if (self.nranks > 1 and self.D == 3) else None
            mu0, mu1, V, sdf = host_measure(self.body, self.dims, t, self.eps, self.k_offset, nl)
            self._check(self.lib.wl_set_body_coeffs(self.h, mu0.ctypes.data, mu1.ctypes.data, V.ctypes.data, sdf.ctypes.data))
        elif isinstance(self.body, RigidBody):
            arr, n = self.body.nodes()
            self._check(self.lib.wl_measure_csg(self.h, arr, n, float(self.eps)))
        else:
            self._check(self.lib.wl_measure_body(self.h, C.byref(self.body.c), float(t), float(self.eps)))

    def set_body_coeffs(self, mu0=None, mu1=None, V=None, sdf=None):
        """Upload host-measured BDIM coefficients (arbitrary AutoBody closures), then update!(pois)."""
        def p(a):
            return None if a is None else np.ascontiguousarray(a, np.float32).ctypes.data
        keep = [np.ascontiguousarray(a, np.float32) for a in (mu0, mu1, V, sdf) if a is not None]
        self._check(self.lib.wl_set_body_coeffs(self.h, p(mu0), p(mu1), p(V), p(sdf)))
        del keep
        self._check(self.lib.wl_update_poisson(self.h))

    def mom_step(self, U1=None, acc0=None, acc1=None):
        """`mom_step!(flow,pois)` (src/Flow.jl:156-167)."""
        info = WLStepInfo()
        f = lambda v: None if v is None else _v3f(v)
        self._check(self.lib.wl_mom_step(self.h, f(U1), f(acc0), f(acc1), C.byref(info)))
        self._dts.append(np.float32(info.dt_next))
        self._n += [info.n_pred, info.n_corr]
        self.last_info = info
        return info

    def sim_step(self, t_end=None, remeasure=True, max_steps=None, uBC_t=None, dUdt=None, udf=None):
        """`sim_step!(sim[,t_end];remeasure,max_steps)` (src/WaterLily.jl:128-139)."""
        if udf is not None:
            raise WaterLilyB200Error("WL_ERR_UNSUPPORTED: udf hooks cannot cross the C ABI")
        steps = 0
        while True:
            if t_end is not None and not (self.sim_time() < t_end):
                break
            if max_steps is not None and steps >= max_steps:
                break
            if remeasure and self.body.kind != 0:
                self.measure()
            t1 = float(np.sum(np.asarray(self._dts, np.float32), dtype=np.float32))
            t0 = float(np.float32(t1) - self._dts[-1])
            if self._bc_fn is not None:
                if uBC_t is not None:
                    raise WaterLilyB200Error("WL_ERR_INVALID: a Function uBC (set_bc_function) and uBC_t are mutually exclusive")
                self._upload_bc(t1)
            if self._accel is not None:
                self._upload_accel(t0, t1)
            self.mom_step(None if uBC_t is None else uBC_t(t1),
                          None if dUdt is None else dUdt(t0), None if dUdt is None else dUdt(t1))
            steps += 1
            if t_end is None:
                break

    def solver(self):
        """Stand-alone `solver!(pois)` on the current x (=p) and z (=σ)."""
        n, tl = C.c_int32(), C.c_int32()
        a, b, c = ((C.c_float * 34)() for _ in range(3))
        self._check(self.lib.wl_solver(self.h, C.byref(n), C.byref(tl), a, b, c))
        k = tl.value
        return dict(n=n.value, rinf=list(a[:k]), r1=list(b[:k]), omega=list(c[:k]))

    def poisson_solver(self, itmx=1000):
        """Stand-alone single-level `solver!(p::Poisson;itmx)` with `pcg!` (src/Poisson.jl:166-186,212-223)."""
        n = C.c_int32()
        self._check(self.lib.wl_poisson_solver(self.h, int(itmx), C.byref(n)))
        return n.value

    def trace(self, which):
        i = self.last_info
        k = i.trace_len[which]
        return dict(n=k - 1, rinf=list(i.rinf[which][:k]), r1=list(i.r1[which][:k]), omega=list(i.omega[which][:k]))

    def cfl(self):
        dt = C.c_float()
        self._check(self.lib.wl_cfl(self.h, C.byref(dt)))
        return float(dt.value)

    def pressure_force(self, t=None):
        """`pressure_force(sim)` (src/Metrics.jl:116-133)."""
        out = (C.c_double * 3)()
        t = self.time() if t is None else t
        if isinstance(self.body, ClosureBody):          # host nds (src/Metrics.jl:110-113) -> sparse weights -> device sum
            nl = self.dims[-1] // self.nranks if (self.nranks > 1 and self.D == 3) else None
            idx, w = host_nds(self.body, self.dims, t, self.k_offset, nl)
            self._check(self.lib.wl_set_force_weights(self.h, idx.size, idx.ctypes.data, w.ctypes.data))
            self._check(self.lib.wl_pressure_force_weights(self.h, C.byref(out)))
        elif isinstance(self.body, RigidBody):
            arr, n = self.body.nodes()
            self._check(self.lib.wl_pressure_force_csg(self.h, arr, n, C.byref(out)))
        else:
            self._check(self.lib.wl_pressure_force(self.h, C.byref(self.body.c), float(t), C.byref(out)))
        return np.array(out[: self.D])

    def viscous_force(self, t=None):
        """`viscous_force(sim)` (src/Metrics.jl:140-152)."""
        if isinstance(self.body, ClosureBody):
            raise WaterLilyB200Error("WL_ERR_UNSUPPORTED: viscous_force for a host-measured body")
        out = (C.c_double * 3)()
        t = self.time() if t is None else t
        if isinstance(self.body, RigidBody):
            arr, n = self.body.nodes()
            self._check(self.lib.wl_viscous_force_csg(self.h, arr, n, C.byref(out)))
        else:
            self._check(self.lib.wl_viscous_force(self.h, C.byref(self.body.c), float(t), C.byref(out)))
        return np.array(out[: self.D])

    def moment(self, x0, viscous=False, t=None):
        """`pressure_moment(x₀,sim)` / `viscous_moment(x₀,sim)` (src/Metrics.jl:168-192)."""
        if isinstance(self.body, ClosureBody):
            raise WaterLilyB200Error("WL_ERR_UNSUPPORTED: moments for a host-measured body")
        out = (C.c_double * 3)()
        t = self.time() if t is None else t
        xx = (C.c_double * 3)(*(list(x0) + [0.0] * (3 - len(x0))))
        if isinstance(self.body, RigidBody):
            arr, n = self.body.nodes()
            self._check(self.lib.wl_force_moment_csg(self.h, arr, n, C.byref(xx), int(viscous), C.byref(out)))
        else:
            self._check(self.lib.wl_force_moment(self.h, C.byref(self.body.c), float(t), C.byref(xx), int(viscous), C.byref(out)))
        return np.array(out[: self.D])

    def total_moment(self, x0, t=None):
        """`total_moment(x₀,sim)` (src/Metrics.jl:199)."""
        return self.moment(x0, False, t) + self.moment(x0, True, t)

    def total_force(self, t=None):
        """`total_force(sim)` (src/Metrics.jl:159)."""
        return self.pressure_force(t) + self.viscous_force(t)

    def debug_op(self, op, level=0, arg=0.0):
        self._check(self.lib.wl_debug_op(self.h, int(op), int(level), float(arg)))

    def prof_begin(self, filt=""):
        self._check(self.lib.wl_prof_begin(self.h, filt.encode()))

    def prof_end(self):
        """-> {kernel name: (launches, total_ms)} measured with CUDA events on the launching stream
        (`self.prof_blocks[name]` = thread blocks launched, the work size behind those times)."""
        buf = C.create_string_buffer(1 << 16)
        self.lib.wl_prof_end(self.h, buf, len(buf))
        out, self.prof_blocks = {}, {}
        for line in buf.value.decode().splitlines():
            name, n, ms, blocks = line.split("\t")
            out[name] = (int(n), float(ms))
            self.prof_blocks[name] = int(blocks)
        return out

    def event_record(self, slot):
        self._check(self.lib.wl_event_record(self.h, slot))

    def event_elapsed_ms(self, a, b):
        ms = C.c_float()
        self._check(self.lib.wl_event_elapsed_ms(self.h, a, b, C.byref(ms)))
        return float(ms.value)

    def sync(self):
        self._check(self.lib.wl_sync(self.h))

    def halo_transport(self):
        return {0: "none", 1: "nccl", 2: "p2p"}.get(self.lib.wl_halo_transport(self.h), "?")

    def halo_timing(self):
        """peer-store halos: mean µs per exchange spent [waiting for ready, copying, waiting for done], count."""
        out = (C.c_double * 4)()
        self._check(self.lib.wl_halo_timing(self.h, C.byref(out)))
        n = max(out[3], 1.0)
        return [out[0] / n / 1e3, out[1] / n / 1e3, out[2] / n / 1e3, int(out[3])]

    def halo_stats(self):
        """(number of halo exchanges, bytes sent) since creation."""
        n, b = C.c_int64(), C.c_int64()
        self._check(self.lib.wl_halo_stats(self.h, C.byref(n), C.byref(b)))
        return int(n.value), int(b.value)

    def gather(self, name, level=0, group=None):
        """The GLOBAL array of a field on every rank (test/inspection helper; goes through the host)."""
        a = self.download(name, level)
        if self.nranks == 1:
            return a
        import torch.distributed as dist
        parts = [None] * self.nranks
        dist.all_gather_object(parts, a, group=group)
        return join_slabs(parts)

    def tile_classes(self, level=0):
        """(general, constants, constants + boundary) tile counts of a multigrid level (`wl_tile_classes`)."""
        out = (C.c_int64 * 3)()
        self._check(self.lib.wl_tile_classes(self.h, int(level), C.byref(out)))
        return tuple(int(x) for x in out)

    def launch_count(self):
        return int(self.lib.wl_launch_count(self.h))

    def enable_timing(self, on=True):
        self._check(self.lib.wl_enable_timing(self.h, int(on)))

    def phase_times(self):
        out = (C.c_double * 8)()
        self._check(self.lib.wl_last_phase_times(self.h, C.byref(out)))
        return list(out)


def sim_step(sim, *a, **k):
    return sim.sim_step(*a, **k)


def sim_time(sim):
    return sim.sim_time()


def measure(sim, t=None):
    return sim.measure(t)


class MeanFlow:
    """`MeanFlow(flow; t_init, uu_stats)` (src/Metrics.jl:205-227) with `update!` (:239-249), `reset!` (:231-236),
    `uu` (:251-259) and `copy!(flow, meanflow)` (:260-263).  P/U/UU live on the device; `t` is a host list."""

    def __init__(self, sim, t_init=None, uu_stats=False):
        self.sim, self.uu_stats = sim, bool(uu_stats)
        sim._check(sim.lib.wl_meanflow_init(sim.h, int(self.uu_stats)))
        self.t = [np.float32(sim.time() if t_init is None else t_init)]

    P = property(lambda self: self.sim.download("mean_P"))
    U = property(lambda self: self.sim.download("mean_U"))
    UU = property(lambda self: self.sim.download("mean_UU") if self.uu_stats else None)

    def time(self):
        return self.t[-1] - self.t[0]

    def reset(self, t_init=0.0):
        self.sim._check(self.sim.lib.wl_meanflow_init(self.sim.h, int(self.uu_stats)))
        self.t = [np.float32(t_init)]

    def update(self):
        dt = np.float32(self.sim.time()) - self.t[-1]
        eps = dt / (dt + self.time() + np.finfo(np.float32).eps)
        if len(self.t) == 1:
            eps = np.float32(1)          # the first update takes the instantaneous field
        self.sim._check(self.sim.lib.wl_meanflow_update(self.sim.h, float(eps)))
        self.t.append(self.t[-1] + dt)

    def uu(self):
        U = self.U
        return self.UU - U[None, :] * U[:, None]

    def copy_to_flow(self):
        self.sim._check(self.sim.lib.wl_meanflow_copy_to_flow(self.sim.h))


def pressure_force(sim):
    return sim.pressure_force()


def viscous_force(sim):
    return sim.viscous_force()


def total_force(sim):
    return sim.total_force()
