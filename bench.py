#!/usr/bin/env python
"""bench.py -- ns per cell per time step of WaterLily's hot path (`sim_step!` -> `mom_step!` +
MultiLevelPoisson V-cycle) on B200, with the HBM roofline of the dominant kernel and the CPU
baseline timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload tgv|sphere|circle] [--impl reference]

One "step" = one `sim_step!` (= one mom_step!; plus measure!+update! for remeasured bodies).
Default workload: BASELINE.json configs[1], 3D Taylor-Green vortex 128^3, Float32, periodic.
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# Algorithmic HBM bytes per interior cell per launch of each kernel (Float32; every field the
# kernel must consume read once, every field it must produce written once).
B_ALG = {
    # fused predictor/corrector: reads u(adv) D, u⁰ D(corr only), μ₀ D, V D, μ₁ D²; writes u D  
    "k_mom": {3: (84 + 96) / 2, 2: (48 + 56) / 2},
    # un-fused RHS kernel: reads u_adv D, u⁰ D (corrector: both), V D; writes f D
    "k_mom_rhs": {3: (36 + 48) / 2, 2: (24 + 32) / 2},
    "k_bdim": {3: (3 + 3 + 3 + 9 + 3) * 4 + 6, 2: (2 + 2 + 2 + 4 + 2) * 4 + 4},
    "k_gs_half": {3: 4 * (1 + 0.5 + 3 + 0.5), 2: 4 * (1 + 0.5 + 2 + 0.5)},
    "k_smooth": {3: 28, 2: 24},
    "k_increment": {3: 4 * (1 + 1 + 1 + 3 + 1 + 1), 2: 4 * (1 + 1 + 1 + 2 + 1 + 1)},
}


def tgv3d(i, x, k):
    if i == 1:
        return -np.sin(k * x[0]) * np.cos(k * x[1]) * np.cos(k * x[2])
    if i == 2:
        return np.cos(k * x[0]) * np.sin(k * x[1]) * np.cos(k * x[2])
    return 0 * x[0] * x[1] * x[2]


def workload(name, size=None, zmul=1, nz=None):
    """-> (description, dims, uBC, L, kwargs, body dict or None, remeasure)   (SURVEY §8d).
    zmul stretches the last dimension (weak scaling: every GPU keeps the single-GPU slab)."""
    if name == "tgv":
        N = size or 128
        k = np.float32(2 * np.pi / N)
        kw = dict(perdir=(1, 2, 3), nu=float(1 / (k * 1600)), U=1.0, u0=lambda i, x: tgv3d(i, x, k))
        nz = (nz or N) * zmul
        return (f"3D Taylor-Green vortex {N}^3 Float32 periodic Re=1600" if nz == N else
                f"3D Taylor-Green vortex {N}x{N}x{nz} Float32 periodic Re=1600"), (N, N, nz), (0.0, 0.0, 0.0), float(N), kw, None, False
    if name == "sphere":
        m = size or 256
        R = m / 8
        body = dict(kind="sphere", center=(m / 2 - 1, m / 2 - 1, m * zmul / 2 - 1), radius=R)
        kw = dict(nu=2 * R / 3700)
        return f"3D static sphere Re=3700 grid {2*m}x{m}x{m*zmul} BDIM", (2 * m, m, m * zmul), (1.0, 0.0, 0.0), 2 * R, kw, body, False
    if name == "cylinder":
        m = size or 256
        R = m / 8
        body = dict(kind="cylinder", axis=2, center=(m, m, 0), radius=R, amp=(R, 0, 0), omega=2 * np.pi / (8 * R))
        kw = dict(nu=2 * R / 1000, U=1.0, perdir=(3,))
        return (f"3D moving cylinder remeasure grid {2*m}x{2*m}x{m*zmul}", (2 * m, 2 * m, m * zmul), (0.0, 0.0, 0.0), 2 * R,
                kw, body, True)
    if name == "circle":
        n, m = 96, 64
        R = m / 8
        body = dict(kind="circle", center=(m / 2 - 1, m / 2 - 1), radius=R)
        return "2D circle Re=100 grid 96x64", (n, m), (1.0, 0.0), 2 * R, dict(nu=2 * R / 100), body, False
    raise SystemExit(f"unknown workload {name}")


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md): NVML polled every
    millisecond from a thread (the default timed region lasts ~15 ms, too short for `nvidia-smi -lms`), with
    the nvidia-smi loop as the fallback when NVML cannot be loaded."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.samples, self.mask, self.max_mhz, self.power = [], 0, None, 0.0
        self.stop_flag = False
        self.thread = self.p = self.f = None
        try:
            import pynvml
            import threading
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(int(device))
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self._poll()
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.thread = None
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(device), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def _poll(self):
        nv = self.nv
        self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        try:
            get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            self.mask |= int(get(self.h))
            self.power = max(self.power, nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
        except Exception:
            pass

    def _run(self):
        while not self.stop_flag:
            try:
                self._poll()
            except Exception:
                break
            time.sleep(0.001)

    def stop(self):
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[])
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            try:
                self._poll()
            except Exception:
                pass
            sm = sorted(self.samples)
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=self.max_mhz, power_w_max=self.power, samples=len(sm),
                       reasons=[nm for bit, nm in self.REASONS if self.mask & bit], source="nvml, 1 ms polling")
            return out
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.split(",") for r in open(self.f.name).read().strip().splitlines() if r.count(",") >= 8]
        os.unlink(self.f.name)
        if not rows:
            return out
        sm = sorted(float(r[1]) for r in rows)
        out["sm_mhz"] = sm[len(sm) // 2]
        out["sm_max_mhz"] = float(rows[0][2])
        out["power_w_max"] = max(float(r[3]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for i, nm in enumerate(names):
            if any("Active" == r[5 + i].strip() for r in rows):
                out["reasons"].append(nm)
        out["samples"] = len(rows)
        out["source"] = "nvidia-smi -lms 100"
        return out


# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel, from `ncu --set full`
# captures of these exact workloads on one GPU (summaries under profiles/): (bytes per launch, source)
NCU_TRAFFIC = {
    # mean of the two captured launches (predictor, corrector)
    ("tgv", (128, 128, 128)): ((32.50e6 + 0.21e6 + 58.01e6 + 4.11e6) / 2, "profiles/r1_v15_ncu_full_kmom3_tgv128.tsv (most of the 128^3 working set stays in the 126 MB L2)"),
    ("sphere", (512, 256, 256)): ((1273.3e6 + 387.2e6 + 1678.3e6 + 389.9e6) / 2, "profiles/r1_v15_ncu_full_kmom3_sphere512.tsv"),
}


def measured_peak():
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(pk["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def build_fast_oracle():
    """-O3 -march=native build of the oracle on THIS machine, used only as the CPU baseline."""
    from oracle import wl_oracle as wo
    out = os.path.join(tempfile.gettempdir(), "wl_oracle_fast_%d" % os.getuid())
    os.makedirs(out, exist_ok=True)
    lib = os.path.join(out, "liboracle_f32_fast.so")
    src = os.path.join(ROOT, "oracle", "wl_oracle.cpp")
    if not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(src):
        try:
            subprocess.run(["g++", "-O3", "-march=native", "-fopenmp", "-shared", "-fPIC", "-std=c++17",
                            "-o", lib, src], check=True, capture_output=True)
        except Exception:
            wo.build()
            return os.path.join(ROOT, "oracle", "liboracle_f32.so"), "-O2 strict (fast build failed)"
    return lib, "-O3 -march=native -fopenmp"


def cpu_run(wname, size, steps, warmup):
    """Time the CPU restatement of the reference (oracle port) on all host cores."""
    from oracle import wl_oracle as wo
    path, flags = build_fast_oracle()
    lib = wo.load(fast_path=path)
    desc, dims, uBC, L, kw, body, remeasure = workload(wname, size)
    b = wo.make_body(**body) if body else None
    sim = wo.OracleSim(dims, uBC, L, body=b, lib=lib, **kw)
    for _ in range(warmup):
        sim.sim_step(remeasure=remeasure)
    t0 = time.perf_counter()
    for _ in range(steps):
        sim.sim_step(remeasure=remeasure)
    dt = time.perf_counter() - t0
    cells = int(np.prod(dims))
    return dict(ns=dt / steps / cells * 1e9, ms_per_step=dt / steps * 1e3, n=list(sim.n[-2 * steps:]), desc=desc,
                cores=os.cpu_count(), flags=flags, cells=cells)


def reference_arm(args, rank):
    """`--impl reference`: the reference's own CPU path for this workload.  The reference is Julia and
    Julia is not installed (here or on the GPU box), so this times the oracle port (oracle/wl_oracle.cpp,
    a loop-for-loop restatement) with all host threads.  Rank 0 only."""
    if rank != 0:
        return
    r = cpu_run(args.workload, args.size, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": "ns_per_cell_per_step", "value": r["ns"], "unit": "ns/cell/step",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": r["desc"], "cells": r["cells"], "poisson_iters_per_step": float(np.mean(r["n"])) * 2},
        "cpu_baseline": {"value": r["ns"], "unit": "ns/cell/step", "cores": r["cores"], "kind": "port",
                         "sample": f"{args.steps} full steps of the same workload after {args.warmup} warm-up, {r['flags']}"},
        "e2e": {"value": r["ns"], "unit": "ns/cell/step", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="tgv")
    ap.add_argument("--size", type=int, default=None)
    ap.add_argument("--nz", type=int, default=None, help="tgv: planes per GPU in the last dimension (default: size)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N>1: weak = every GPU keeps the single-GPU grid (last dimension x N), strong = the same grid split N ways")
    ap.add_argument("--breakdown", default=None, help="write a per-kernel CUDA-event breakdown (3 extra steps) here")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return reference_arm(args, rank)

    import torch
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.barrier()
    import waterlily_jl_b200 as wl

    # N>1: ONE simulation, slab-decomposed along the last dimension, one rank per GPU
    desc, dims, uBC, L, kw, body, remeasure = workload(args.workload, args.size, world if args.scaling == "weak" else 1, args.nz)
    D = len(dims)
    cells = int(np.prod(dims))                       # whole job
    comm = wl.make_slab_comm() if world > 1 else None
    sim = wl.Simulation(dims, uBC, L, body=wl.AutoBody(**body) if body else None, device=local, comm=comm, **kw)

    def barrier():
        sim.sync()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        sim.sim_step(remeasure=remeasure)

    # ---- timed region: K steps, state resident in HBM, CUDA events on the launching stream
    dom = "k_mom3" if D == 3 else "k_mom_rhs"
    balg_key = "k_mom" if D == 3 else "k_mom_rhs"
    clocks = ClockSampler(local)
    barrier()
    l0 = sim.launch_count()
    sim.prof_begin(dom)
    sim.event_record(0)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sim.sim_step(remeasure=remeasure)
    sim.event_record(1)
    sim.sync()
    wall = time.perf_counter() - t0
    dev_ms = sim.event_elapsed_ms(0, 1)
    prof = sim.prof_end()
    launches = sim.launch_count() - l0
    clk = clocks.stop()
    n_iters = list(sim.pois.n[-2 * args.steps:])
    t_max = dev_ms
    if dist is not None:
        t = torch.tensor([dev_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_max = float(t.item())
    ms_per_step = t_max / args.steps
    value = t_max * 1e6 / args.steps / cells               # ns per cell per step, whole job

    # ---- roofline of the dominant kernel (live CUDA events inside the timed region)
    peak, peak_src = measured_peak()
    kn, kms = 0, 0.0
    for name, (n, ms) in prof.items():
        kn += n
        kms += ms
    balg = B_ALG[balg_key][D]
    achieved = (balg * cells / world) / (kms / max(kn, 1) * 1e-3) / 1e9 if kn else None     # per GPU (rank 0's kernel)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None,
                "traffic": NCU_TRAFFIC.get((args.workload, tuple(dims)), (None, None))[0] if world == 1 else None,
                "traffic_source": NCU_TRAFFIC.get((args.workload, tuple(dims)), (None, None))[1] if world == 1 else None,
                "alg_bytes_per_cell": balg, "launches": kn, "avg_ms": kms / max(kn, 1), "peak_source": peak_src,
                "share_of_step": kms / dev_ms}

    # ---- e2e: the same step through the public API with HOST buffers (pinned): upload state, step, download
    shape_u, shape_p = sim.field_shape("u"), sim.field_shape("p")
    hu = torch.empty(shape_u, dtype=torch.float32, pin_memory=True).numpy()
    hp = torch.empty(shape_p, dtype=torch.float32, pin_memory=True).numpy()
    sim.download("u", out=hu)
    sim.download("p", out=hp)
    barrier()
    e2e_steps = max(3, min(args.steps, 10))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        sim.upload("u", hu)
        sim.upload("p", hp)
        sim.sim_step(remeasure=remeasure)
        sim.download("u", out=hu)
        sim.download("p", out=hp)
    sim.sync()
    e2e_wall = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e2e_wall], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_wall = float(t.item())
    e2e = {"value": e2e_wall / e2e_steps / cells * 1e9, "unit": "ns/cell/step",
           "h2d_bytes_per_step": int(hu.nbytes + hp.nbytes) * world, "d2h_bytes_per_step": int(hu.nbytes + hp.nbytes) * world,
           "steps": e2e_steps, "what": "upload u,p from pinned host + sim_step + download u,p"}

    # ---- optional per-kernel breakdown (outside the timed region)
    phases = None
    if args.breakdown:                      # every rank steps (collective); rank 0 writes
        sim.enable_timing(True)
        sim.sim_step(remeasure=remeasure)
        phases = [round(t * 1e3, 4) for t in sim.phase_times()[:5]]
        sim.enable_timing(False)
        sim.prof_begin("")
        sim.event_record(2)
        for _ in range(3):
            sim.sim_step(remeasure=remeasure)
        sim.event_record(3)
        allp = sim.prof_end()
        tot = sim.event_elapsed_ms(2, 3)
        with open(args.breakdown if rank == 0 else os.devnull, "w") as f:
            f.write(f"# phases of one step [ms]: predictor {phases[0]}, project {phases[1]}, corrector {phases[2]}, "
                    f"project {phases[3]}, CFL {phases[4]}\n")
            f.write(f"# {desc}; 3 steps, {tot:.3f} ms total with per-launch events (events add overhead)\n")
            f.write("kernel\tlaunches\ttotal_ms\tavg_us\tshare_of_kernel_time\n")
            ksum = sum(ms for _, ms in allp.values())
            for name, (n, ms) in sorted(allp.items(), key=lambda kv: -kv[1][1]):
                f.write(f"{name}\t{n}\t{ms:.4f}\t{ms / n * 1e3:.2f}\t{ms / ksum:.3f}\n")

    halo = sim.halo_stats()
    halo_us = sim.halo_timing() if world > 1 else None
    nsteps_total = len(sim.pois.n) // 2
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_size = {"tgv": min(args.size or 128, 128), "sphere": 64, "cylinder": 64, "circle": None}[args.workload]
        r = cpu_run(args.workload, cpu_size, 3, 1)
        big = r["cells"] != cells
        cpu = {"value": r["ns"], "unit": "ns/cell/step", "cores": r["cores"], "kind": "port",
               "sample": f"3 steps of {r['desc']} after 1 warm-up, oracle port {r['flags']}" +
                         (" (reduced grid, per-cell figure)" if big else "")}

    if rank == 0:
        line = {
            "metric": "ns_per_cell_per_step", "value": value, "unit": "ns/cell/step", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": False,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "cells": cells, "cells_per_gpu": cells // world, "remeasure": remeasure,
                       "parallelism": "single GPU" if world == 1 else
                       f"{world} slabs along the last dimension, halo exchange by {sim.halo_transport()} ({halo[0] // max(1, nsteps_total)} exchanges, "
                       f"{halo[1] / max(1, nsteps_total) / 1e6:.1f} MB sent per rank per step), coarse levels replicated",
                       "poisson_iters_per_step": float(np.mean(n_iters)) * 2,
                       "l2": "working set (%.0f MB per GPU) exceeds L2, no flush" % (30 * 4 * cells / world / 1e6)
                             if cells / world > 1.5e6 else "fits L2"},
            "halo_us_per_exchange": halo_us, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
            "wall_ms_per_step": wall / args.steps * 1e3,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
