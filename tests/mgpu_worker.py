"""Multi-GPU parity worker (run under torchrun, one rank per GPU; see tests/test_multigpu.py):
the slab-decomposed run must reproduce the single-GPU run of the same library on the same inputs
-- identical Poisson iteration counts and Δt history, fields within 1e-6·scale (SURVEY §8e parity
requirement; reductions fold per-rank partials, so the last bit of a norm may differ).

  torchrun --nproc-per-node 2 tests/mgpu_worker.py [case ...]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def tgv(N, nz):
    k = np.float32(2 * np.pi / N)

    def u0(i, x):
        if i == 1:
            return -np.sin(k * x[0]) * np.cos(k * x[1]) * np.cos(k * x[2])
        if i == 2:
            return np.cos(k * x[0]) * np.sin(k * x[1]) * np.cos(k * x[2])
        return 0 * x[0] * x[1] * x[2]
    return dict(dims=(N, N, nz), uBC=(0.0, 0.0, 0.0), L=float(N),
                kw=dict(perdir=(1, 2, 3), nu=float(1 / (k * 1600)), U=1.0, u0=u0), body=None, remeasure=False)


CASES = {
    # periodic in every direction: ring exchange, stale-ghost sweeps at the wrap, mean removal
    "tgv32": lambda: dict(tgv(32, 32), steps=6),
    "tgv_32x32x64": lambda: dict(tgv(32, 64), steps=4),
    # body band straddling the slab interface, Dirichlet/Neumann physical boundaries in z
    "sphere": lambda: dict(dims=(64, 32, 32), uBC=(1.0, 0.0, 0.0), L=8.0, kw=dict(nu=8.0 / 100),
                           body=dict(kind="sphere", center=(15, 15, 15), radius=4.0), remeasure=False, steps=6),
    "sphere_exit": lambda: dict(dims=(64, 32, 32), uBC=(1.0, 0.0, 0.0), L=8.0, kw=dict(nu=8.0 / 100, exitBC=True),
                                body=dict(kind="sphere", center=(15, 15, 17), radius=4.0), remeasure=False, steps=6),
    # moving body, remeasure every step (measure! + update! across slabs), periodic last dimension
    "cylinder": lambda: dict(dims=(64, 64, 32), uBC=(0.0, 0.0, 0.0), L=8.0,
                             kw=dict(nu=8.0 / 100, U=1.0, perdir=(3,)),
                             body=dict(kind="cylinder", axis=2, center=(32, 32, 0), radius=4.0, amp=(4.0, 0, 0),
                                       omega=2 * np.pi / 32), remeasure=True, steps=5),
    "sphere64": lambda: dict(dims=(64, 32, 64), uBC=(1.0, 0.0, 0.0), L=8.0, kw=dict(nu=8.0 / 100, exitBC=True),
                             body=dict(kind="sphere", center=(15, 15, 31), radius=5.0), remeasure=False, steps=5),
    # torus ring crossing the interface, vanLeer scheme, semi-coarsening hierarchy (96 = 3·32)
    "torus": lambda: dict(dims=(48, 32, 32), uBC=(1.0, 0.0, 0.0), L=8.0, kw=dict(nu=0.05, scheme="vanLeer"),
                          body=dict(kind="torus", axis=0, center=(20, 15, 15), radius=6.0, minor=2.0),
                          remeasure=False, steps=4),
    # the stand-alone Poisson/pcg! solver on slabs (perdot = gathered partial sums), after two ordinary steps
    "pcg": lambda: dict(dims=(32, 32, 32), uBC=(1.0, 0.0, 0.0), L=8.0, kw=dict(nu=8.0 / 100),
                        body=dict(kind="sphere", center=(12, 15, 15), radius=4.0), remeasure=False, steps=2, pcg=True),
    # slab-parallel output (io.VTKWriter): one .vti piece per rank + rank 0's .pvti index, read back = the single-GPU fields
    "vtk": lambda: dict(dims=(32, 32, 32), uBC=(1.0, 0.0, 0.0), L=8.0, kw=dict(nu=8.0 / 100),
                        body=dict(kind="sphere", center=(12, 15, 15), radius=4.0), remeasure=False, steps=2, vtk=True),
    # Function uBC(i,x,t) varying along the decomposed direction + its ∂ₜ in accelerate!: every rank samples its own slab
    # of the tables at global coordinates (physical z ends on the first / last rank, interfaces in between)
    "fn_ubc": lambda: dict(dims=(32, 24, 16), uBC=_ubc, L=8.0, kw=dict(nu=0.02, U=1.0, dt=0.1, dUBCdt=_dub),
                           body=dict(kind="sphere", center=(10, 11, 7.5), radius=3.0), remeasure=False, steps=6),
    # acceleration field g(i,x,t) through K-MOM's field slot, periodic last dimension (wrap link between the end ranks)
    "fn_g": lambda: dict(dims=(32, 24, 16), uBC=(1.0, 0.0, 0.0), L=8.0, kw=dict(nu=0.02, perdir=(3,), g=_g),
                         body=dict(kind="sphere", center=(10, 11, 8.0), radius=3.0), remeasure=False, steps=5),
}
_T = np.float32
_ubc = lambda i, x, t: (_T(1) + _T(0.2) * t + _T(0.01) * x[1] + _T(0.005) * x[2]) if i == 1 else _T(0) * x[0] * t
_dub = lambda i, x, t: (_T(0.2) + _T(0) * x[1] * t) if i == 1 else _T(0) * x[0] * t
_g = lambda i, x, t: _T(0.01) * x[i % 3] - _T(0.02) * t * x[(i + 1) % 3]


def run_case(wl, name, comm, dev):
    c = CASES[name]()
    mk = lambda **extra: wl.Simulation(c["dims"], c["uBC"], c["L"], body=wl.AutoBody(**c["body"]) if c["body"] else None,
                                       device=dev, **c["kw"], **extra)
    one = mk()                 # the whole grid on this rank's GPU
    slab = mk(comm=comm)       # this rank's slab
    rank, n = comm[0], comm[1]
    for _ in range(c["steps"]):
        one.sim_step(remeasure=c["remeasure"])
        slab.sim_step(remeasure=c["remeasure"])
    out = dict(case=name, rank=rank, n_one=list(one.pois.n), n_slab=list(slab.pois.n))
    out["dt_err"] = float(np.abs(np.array(one.flow.Δt) - np.array(slab.flow.Δt)).max())
    ok = out["n_one"] == out["n_slab"] and out["dt_err"] <= 1e-6
    for f in ("u", "p", "mu0", "V"):
        ref = wl.slab_of(getattr(one.flow, f), rank, n)
        got = getattr(slab.flow, f)
        scale = max(1.0, float(np.abs(ref).max()))
        # ghost planes at the physical ends of periodic directions may legitimately hold other images
        err = float(np.abs(got - ref)[..., 1:-1, :, :].max()) / scale
        out["err_" + f] = err
        ok = ok and err <= 1e-6
    if c["body"]:
        f1, f2 = one.pressure_force(), slab.pressure_force()
        out["force_err"] = float(np.abs(f1 - f2).max() / max(1.0, np.abs(f1).max()))
        ok = ok and out["force_err"] <= 1e-6
    if c.get("pcg"):
        # stand-alone solver!(p::Poisson) with pcg! (src/Poisson.jl:166-186,212-223) on the stepped state: a smooth,
        # zero-mean source in σ, x = 0.  The dot products are `perdot` sums over the slabs (Poisson.jl:150-157); CG
        # amplifies their last-bit differences, so x agrees to the solver's tolerance, not to rounding
        shp = one.field_shape("sigma")
        zz, yy, xx = np.meshgrid(*(np.arange(m, dtype=np.float32) for m in shp), indexing="ij")
        src = (np.sin(0.3 * xx) * np.cos(0.2 * yy) * np.sin(0.25 * zz + 0.4)).astype(np.float32)
        inner = (slice(1, -1),) * 3
        src[inner] -= src[inner].mean(dtype=np.float64).astype(np.float32)
        zero = np.zeros(shp, np.float32)
        one.upload("sigma", src); one.upload("p", zero)
        slab.upload("sigma", wl.slab_of(src, rank, n)); slab.upload("p", wl.slab_of(zero, rank, n))
        n1, n2 = one.poisson_solver(200), slab.poisson_solver(200)
        ref = wl.slab_of(one.flow.p, rank, n)
        got = slab.flow.p
        out["pcg_n"] = [n1, n2]
        out["pcg_err"] = float(np.abs(got - ref)[1:-1].max() / max(1e-30, float(np.abs(ref).max())))
        ok = ok and abs(n1 - n2) <= 1 + n1 // 50 and n1 > 0 and out["pcg_err"] <= 1e-3
    if c.get("vtk"):
        import tempfile
        import torch.distributed as dist
        from waterlily_jl_b200 import io as wio
        box = [tempfile.mkdtemp(prefix="wl_vtk_") if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        cwd = os.getcwd()
        os.chdir(box[0])                                   # `fname` is a name, not a path (as in the reference's vtkWriter)
        w = wio.VTKWriter("slab", dir="OUT")
        path = w.save(slab)
        w.close(rank=rank)
        dist.barrier()                                     # all pieces on disk before anyone reads the index
        grid, data = wio.read_pvti(path)
        out["vtk_files"] = sorted(os.listdir("OUT"))
        out["vtk_err"] = max(float(np.abs(data["Pressure"] - one.flow.p).max()), float(np.abs(data["Velocity"] - one.flow.u).max()))
        ok = ok and tuple(grid) == tuple(one.field_shape("p")) and out["vtk_err"] <= 1e-6 and len(out["vtk_files"]) == n + 1
        ok = ok and [f for _, f in wio._read_pvd("slab.pvd")] == ["OUT/slab_000000.pvti"] if rank == 0 else ok
        # restart from the collection (load_pvd: last file, Δt[end] from the stored time, push CFL): the slab run takes its
        # slab of the reassembled arrays, the single-GPU run the whole of them; one more step must agree again
        slab.upload("p", np.zeros(slab.field_shape("p"), np.float32))
        slab.upload("u", np.zeros(slab.field_shape("u"), np.float32))
        wio.load_pvd(slab, "slab.pvd")
        wio.load_pvd(one, "slab.pvd")
        r_err = 0.0
        for step in range(2):
            for f in ("u", "p"):
                d = np.abs(getattr(slab.flow, f) - wl.slab_of(getattr(one.flow, f), rank, n))[..., 1:-1, :, :]
                r_err = max(r_err, float(d.max()))
            if step == 0:
                one.sim_step(remeasure=False)
                slab.sim_step(remeasure=False)
        out["vtk_restart_err"] = r_err
        out["vtk_restart_dt"] = [float(x) for x in slab.flow.Δt[-2:]]
        ok = ok and r_err <= 1e-6 and [float(x) for x in one.flow.Δt] == [float(x) for x in slab.flow.Δt]
        dist.barrier()
        os.chdir(cwd)
    slab.sync()
    out["halo"] = slab.halo_stats()
    out["transport"] = slab.halo_transport()
    out["ok"] = bool(ok)
    slab.close()               # collective: every rank, at the same point (never left to the garbage collector)
    one.close()
    return out


def main():
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        ge.build()
    dist.barrier()
    import waterlily_jl_b200 as wl
    names = sys.argv[1:] or list(CASES)
    bad = 0
    for nm in names:
        comm = wl.make_slab_comm()
        res = run_case(wl, nm, comm, local)
        allres = [None] * dist.get_world_size()
        dist.all_gather_object(allres, res)
        if rank == 0:
            for r in allres:
                print("MGPU", json.dumps(r), flush=True)
                bad += 0 if r["ok"] else 1
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MGPU_RESULT", "FAIL" if bad else "PASS", flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
