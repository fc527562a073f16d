/* wl_b200.h -- C ABI of libwaterlily_b200.so: the B200 (sm_100a) implementation of
 * WaterLily.jl's time-step hot path (sim_step! -> mom_step! + MultiLevelPoisson V-cycle).
 *
 * This is the drop-in boundary.  The Julia side (waterlily.jl_b200/julia/WaterLilyB200.jl)
 * binds these symbols with `ccall`; waterlily.jl_b200/__init__.py binds the same symbols with
 * ctypes.  No torch / C++ types cross this boundary: plain pointers, sizes and POD structs.
 *
 * Conventions
 *  - Host arrays passed in/out use the reference's layout exactly: Julia column-major,
 *    size (N1+2, N2+2[, N3+2][, D[, D]]), Float32, one ghost layer per side, component slowest
 *    (reference src/Flow.jl:133-147).  The library repacks into its padded device layout.
 *  - Every function returns 0 on success or a WL_ERR_* code; wl_last_error() gives the text.
 *    Nothing aborts or exits.  One host thread per wl_sim (distinct sims are independent).
 *  - All device memory is owned by the wl_sim; host buffers are borrowed for the call only.
 *  - Unsupported through this boundary (WL_ERR_UNSUPPORTED, never a silent CPU fallback):
 *    space-varying Function uBC / g, udf hooks, user-defined convective schemes, T != Float32.
 */
#ifndef WL_B200_H
#define WL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WL_OK 0
#define WL_ERR_INVALID 1       /* bad argument / size rule violated                         */
#define WL_ERR_UNSUPPORTED 2   /* feature of the reference API this path does not implement */
#define WL_ERR_CUDA 3          /* CUDA runtime error (text in wl_last_error)                 */
#define WL_ERR_NO_DEVICE 4     /* no usable sm_100 device: the library never falls back      */

typedef struct wl_sim wl_sim;

/* Replaces the keyword arguments of `Simulation(dims,uBC,L;...)` / `Flow(N,uBC;...)`
 * (reference src/WaterLily.jl:93-98, src/Flow.jl:133-134) and of
 * `MultiLevelPoisson(x,L,z;maxlevels,perdir)` / `solver!(ml;tol,itmx)`
 * (src/MultiLevelPoisson.jl:68,108). */
typedef struct wl_config {
  int32_t D;            /* 2 or 3                                                    */
  int32_t N[3];         /* interior cells per direction (dims)                       */
  int32_t perdir[3];    /* 1 if direction is periodic                                */
  int32_t exitBC;       /* convective exit in direction 1                            */
  int32_t scheme;       /* λ: 0 quick, 1 cds, 2 vanLeer (src/Flow.jl:4-6)            */
  float nu;             /* ν                                                         */
  float dt0;            /* Δt[1]                                                     */
  float uBC[3];         /* constant-tuple boundary velocity                          */
  int32_t maxlevels;    /* 10                                                        */
  float tol;            /* 2e-3                                                      */
  int32_t itmx;         /* 32                                                        */
  int32_t device;       /* CUDA device ordinal                                       */
  int32_t flags;        /* WL_FLAG_*                                                 */
} wl_config;

/* (flag bit 1 is unused: round 1 reserved it for a CUDA-graph replay that was never built) */
#define WL_FLAG_NO_FUSED 2    /* 3D: use the un-fused K-RHS + K-BDIM kernels (materialises flow.f) */
#define WL_FLAG_FUSED_SMOOTH 4 /* 3D: K-GS2, two red-black sweeps per pass (wl_kernels_gs2.cuh); bit-identical, opt-in: slower for now */

/* Parametric AutoBody: primitive sdf composed with map(x,t) = x - centre(t),
 * centre(t) = center + vel*t + acc*t^2 + amp*sin(omega*t)  (src/AutoBody.jl:1-37).
 * Arbitrary Julia closures cannot cross a C ABI; they are measured on the host by the
 * reference's own measure! and uploaded with wl_set_body_coeffs. */
typedef struct wl_body {
  int32_t kind;         /* 0 NoBody, 1 sphere/circle, 2 cylinder along `axis`, 3 torus with ring normal `axis` */
  int32_t axis;
  float center[3];
  float radius;
  float minor;
  float vel[3], acc[3], amp[3];
  float omega;
} wl_body;

/* One entry of a postfix `SetBody` program (src/Body.jl:91-107) whose leaves are library sdfs under a
 * `RigidMap(x0, theta; xp, V, omega)` (src/RigidMap.jl:28-50).  `a + b`, `a ∪ b` -> a b {op 1}; `a ∩ b` -> a b {op 2};
 * `-a` -> a {op 3}; `a - b` -> a b {op 3} {op 2}.  At most WL_CSG_MAX entries.  In 2D theta[2] / omega[2] are the
 * scalar angle / angular velocity.  `setmap(body; ...)` (RigidMap.jl:54-56) = rewrite the fields and measure again. */
#define WL_CSG_MAX 8
typedef struct wl_body_node {
  int32_t op;           /* 0 leaf, 1 union (min), 2 intersection (max), 3 complement                             */
  int32_t kind;         /* leaf: 0 NoBody, 1 sphere/circle, 2 cylinder, 3 torus, 4 capsule (segment +-minor on `axis`) */
  int32_t axis;
  float radius, minor;
  float x0[3], theta[3], xp[3], V[3], omega[3];
} wl_body_node;

/* Fields a caller can read or write (sim.flow.* and sim.pois.levels[l].*). */
enum wl_field {
  WL_U = 0, WL_U0 = 1, WL_F = 2, WL_P = 3, WL_SIGMA = 4, WL_V = 5, WL_MU0 = 6, WL_MU1 = 7,
  WL_X = 10, WL_R = 11, WL_EPS = 12, WL_Z = 13, WL_L = 14, WL_DIAG = 15, WL_IDIAG = 16,
  WL_MEAN_P = 20, WL_MEAN_U = 21, WL_MEAN_UU = 22,  /* MeanFlow.P/.U/.UU, after wl_meanflow_init */
  WL_ACCEL_T0 = 23, WL_ACCEL_T1 = 24                /* accelerate! fields, after wl_set_accel_field */
};

/* Per-step results: replaces `push!(a.Δt,CFL(a))` (src/Flow.jl:166), the two `push!(ml.n,nᵖ)`
 * (src/MultiLevelPoisson.jl:127) and the `@log` solver trace (:112,117). */
typedef struct wl_step_info {
  float dt_next;
  int32_t n_pred, n_corr;
  int32_t flags;                       /* bit0: non-finite residual seen, bit1: itmx reached */
  int32_t trace_len[2];                /* entries valid in each trace (n+1)                 */
  float rinf[2][34], r1[2][34], omega[2][34];
} wl_step_info;

/* ---- lifetime.  wl_create = Flow constructor (src/Flow.jl:133-147) with the default IC.      */
int wl_create(const wl_config* cfg, wl_sim** out);

/* ---- multi-GPU (the reference has no distributed path, its nearest hook is the
 *      `mem`/`flow_ctor` injection of src/WaterLily.jl:93-98).  One process per GPU; the grid is
 *      slab-decomposed along the LAST dimension.  cfg describes the GLOBAL grid; every rank calls
 *      wl_create_slab with the same cfg and the same NCCL id (made once by wl_nccl_unique_id and
 *      broadcast by the host) and owns N[D-1]/nranks planes.  All data-surface calls then refer to
 *      the local slab: size (N1+2, N2+2, N3/nranks+2), whose first/last planes mirror the
 *      neighbouring slabs.  Every entry point is collective: all ranks call it in the same order.  */
typedef struct wl_dist {
  int32_t rank, nranks;
  char nccl_id[128];
} wl_dist;
int wl_nccl_unique_id(char id[128]);
int wl_create_slab(const wl_config* cfg, const wl_dist* dist, wl_sim** out);
int wl_slab_info(const wl_sim* s, int32_t* rank, int32_t* nranks, int32_t* k_offset, int32_t* n_local, int32_t* n_global);
int wl_halo_stats(const wl_sim* s, int64_t* exchanges, int64_t* bytes);
/* 0 single GPU, 1 NCCL send/recv, 2 direct peer stores over NVLink (CUDA IPC; WL_NO_P2P=1 disables) */
int wl_halo_transport(const wl_sim* s);
/* peer-store halos: accumulated ns [waiting for the neighbours' ready, copy + fence, waiting for their done]
 * and the number of exchanges, as seen by this rank's halo kernels */
int wl_halo_timing(wl_sim* s, double out[4]);
void wl_destroy(wl_sim* s);
const char* wl_last_error(const wl_sim* s);   /* s may be NULL: last error of wl_create */
int wl_version(void);

/* ---- data surface: Array(field) / copyto!(field, ::Array)  (ext/WaterLilyJLD2Ext.jl:13-14,43-44) */
int wl_field_dims(const wl_sim* s, int field, int level, int32_t dims[5], int32_t* ndims);
int wl_download(wl_sim* s, int field, int level, float* host, size_t count);
int wl_upload(wl_sim* s, int field, int level, const float* host, size_t count);
int wl_nlevels(const wl_sim* s);

/* ---- initial condition: after uploading WL_U this applies BC!, exitBC!(u,u,0), u⁰=u (src/Flow.jl:141-142) */
int wl_finish_ic(wl_sim* s, const float uBC[3]);

/* ---- body: measure!(flow,body;t,ϵ) (src/Body.jl:28-51)                                      */
/* Function `uBC(i,x,t)` (src/core.jl:201-219; the Flow ctor and `mom_step!` call `BC!(a.u,a.uBC,a.exitBC,a.perdir,t)`,
 * src/Flow.jl:141,195,209,231).  A closure cannot cross this boundary, so the caller samples it at the points BC!
 * evaluates it at and hands over the table; every later BC! on the velocity (wl_finish_ic, wl_mom_step) uses it
 * until it is replaced or cleared (tables == NULL).  Layout: for component ci (outer) and direction dj (inner),
 * three planes over the other dims, ghosts included, lowest remaining dim fastest:
 *   ci == dj: uBC at index 0, 1 and n-1 of dj;   ci != dj: uBC(I)-uBC(I+e_dj) at index 0, (unused), uBC(I)-uBC(I-e_dj)
 *   at index n-1.  Periodic directions keep their (ignored) slots.  `wl_bc_table_size` gives the element count.
 * On a slab-decomposed sim every rank passes the tables of ITS slab (local plane indices, global coordinates); the planes
 * of the last direction are only read on the ranks that own a physical end.  `∂ₜuBC` and `g(i,x,t)` of `accelerate!`
 * go through wl_set_accel_field.  GPU parity: tests/test_fnbc_gpu.py, tests/test_fullsize_gpu.py, tests/mgpu_worker.py. */
int wl_bc_table_size(const wl_sim* s, size_t* count);
int wl_set_bc_tables(wl_sim* s, const float* tables, size_t count);

/* `accelerate!(r,t,g,U)` (src/Flow.jl:69-73) with Function `g(i,x,t)` / `uBC(i,x,t)`: the caller samples
 * `g + ∂ₜuBC` on every index of every component (layout of WL_F) at t₀ = sum(Δt)-Δt[end] (`which` = 0, predictor) and
 * t₁ = sum(Δt) (`which` = 1, corrector) before wl_mom_step; the fields are added to the convective-diffusive right-hand
 * side after the uniform term (3D: a slot of the fused K-MOM and of its band patch; 2D: the per-operator K-RHS).
 * `field == NULL` clears both.  Slab-decomposed sims pass their slab of the field (no exchange is needed: only the
 * cell's own value is read). */
int wl_set_accel_field(wl_sim* s, int which, const float* field, size_t count);

int wl_measure_body(wl_sim* s, const wl_body* body, double t, double eps);
/* host-measured coefficients (values BEFORE the two BC! calls of Body.jl:49-50, which are applied here) */
int wl_set_body_coeffs(wl_sim* s, const float* mu0, const float* mu1, const float* V, const float* sdf);

/* ---- Poisson: pois_ctor(flow) (src/WaterLily.jl:97), update! (src/MultiLevelPoisson.jl:79-86),
 *      solver! (src/MultiLevelPoisson.jl:108-128) on the current x (=p) and z (=σ).            */
int wl_build_poisson(wl_sim* s);
int wl_update_poisson(wl_sim* s);
int wl_solver(wl_sim* s, int32_t* n_iter, int32_t* trace_len, float* rinf, float* r1, float* omega);

/* ---- stand-alone single-level solver: `solver!(p::Poisson;tol,itmx)` with the Jacobi-preconditioned
 *      `pcg!(p;it=6)` smoother (src/Poisson.jl:166-186,212-223) on level 1 (x = p, z = σ, L = μ₀), for
 *      users who pass `pois_ctor = flow->Poisson(flow.p,flow.μ₀,flow.σ)`.  tol comes from the config.
 *      On a slab-decomposed sim the dot products are global (`perdot`, src/Poisson.jl:150-157): per-rank partial sums
 *      folded in rank order, identical on every rank.                                            */
int wl_poisson_solver(wl_sim* s, int32_t itmx, int32_t* n_iter);

/* ---- time step: mom_step! (src/Flow.jl:156-167).  uBC_t1 = boundary velocity at t₁ (NULL: cfg.uBC);
 *      acc0/acc1 = space-uniform g+dU/dt at t₀/t₁ for accelerate! (src/Flow.jl:69-73), may be NULL.
 *      Uses and replaces the stored Δt (wl_get_dt / wl_set_dt).                                */
int wl_mom_step(wl_sim* s, const float* uBC_t1, const float* acc0, const float* acc1, wl_step_info* info);
int wl_get_dt(const wl_sim* s, float* dt);
int wl_set_dt(wl_sim* s, float dt);
int wl_cfl(wl_sim* s, float* dt);                       /* CFL(a) src/Flow.jl:234-237 */

/* ---- allocation-free CFL.  NO COUNTERPART IN THE REFERENCE: added for callers
 *      that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 *      Computes what wl_cfl computes -- Δt = min(10, 1/(max_I flux_out(I,u) + 5ν)), src/Flow.jl:234-244, the maximum
 *      taken over the interior cells and the zero of the lower ghost cells -- but needs no wl_sim and allocates
 *      nothing: its reduction scratch is a file-scope `static __device__` array (1 MiB + 256 B on every GPU that
 *      loads the library, held until exit).  The call makes no cudaMalloc / cudaFree / allocating library call; it
 *      enqueues one 4-byte memset and one kernel on `cuda_stream` (a cudaStream_t; NULL = the legacy default stream)
 *      and returns without synchronising.  The first call on each device resolves the array's device address
 *      (cudaGetSymbolAddress): make one warm-up call per device before capturing a graph.
 *        u_dev   device pointer, Float32, the reference's dense layout (n[0], n[1][, n[2]], D), ghosts included
 *        n       extents INCLUDING the two ghost cells per direction (n[2] ignored for D == 2)
 *        dt_dev  device pointer to one float (output)
 *      LIMIT: (n[1]-2)·(n[2]-2) <= 262144 interior rows (e.g. 512x512x512 cells); larger grids return WL_ERR_INVALID
 *      with the limit in wl_last_error(NULL).  All calls on a GPU SHARE the array, which keeps the previous call's
 *      contents (the call clears what must start at zero): calls that overlap on one GPU race -- serialise them.   */
int wl_cfl_noalloc(const float* u_dev, int32_t D, const int32_t n[3], float nu, float* dt_dev, void* cuda_stream);

/* ---- pressure_force(sim) (src/Metrics.jl:116-133); clobbers flow.f like the reference        */
int wl_pressure_force(wl_sim* s, const wl_body* body, double t, double out[3]);
/* ---- pressure_force for a HOST-measured body (any `AutoBody(sdf,map)` closure; replaces `pressure_force(flow,body)`,
 *      src/Metrics.jl:126-133, for bodies the library cannot evaluate itself): the caller evaluates
 *      `nds(body,loc(0,I),t)` (:110-113) on the interior cells with |d| < 1 and hands over the non-zero entries --
 *      idx = 0-based linear index into the (N1+2,N2+2[,N3+2]) scalar array, first dimension fastest (the local slab
 *      on a slab-decomposed sim), w = D floats per entry.  wl_pressure_force_weights forms p[I]*w in Float32 and
 *      sums in Float64 like the reference; the weights stay valid until replaced (moving bodies: every step). */
int wl_set_force_weights(wl_sim* s, int64_t nnz, const int64_t* idx, const float* w);
int wl_pressure_force_weights(wl_sim* s, double out[3]);
/* ---- viscous_force(sim) (src/Metrics.jl:140-152): -2ν Σ S(I,u)·nds; clobbers flow.f like the reference.
 *      total_force(sim) (:159) is the sum of the two, formed by the caller.                     */
int wl_viscous_force(wl_sim* s, const wl_body* body, double t, double out[3]);
/* ---- pressure_moment(x₀,sim) (viscous = 0) / viscous_moment(x₀,sim) (viscous = 1) (src/Metrics.jl:168-192);
 *      x₀ is converted to Float32 (the reference promotes to the type of x₀).  In 2D both components of
 *      `out` hold the scalar moment, as the reference's broadcast leaves them.                   */
int wl_force_moment(wl_sim* s, const wl_body* body, double t, const double x0[3], int viscous, double out[3]);

/* The same four calls for a `SetBody` / `RigidMap` body given as a postfix program of `n` nodes:
 * `measure!(flow, body; ϵ)` (src/Body.jl:28-51 with :104-107 and src/AutoBody.jl:29-37 per leaf), `pressure_force`,
 * `viscous_force` and the moments (src/Metrics.jl:116-192).  The maps carry no time: the caller rewrites the nodes. */
int wl_measure_csg(wl_sim* s, const wl_body_node* nodes, int32_t n, double eps);
int wl_pressure_force_csg(wl_sim* s, const wl_body_node* nodes, int32_t n, double out[3]);
int wl_viscous_force_csg(wl_sim* s, const wl_body_node* nodes, int32_t n, double out[3]);
int wl_force_moment_csg(wl_sim* s, const wl_body_node* nodes, int32_t n, const double x0[3], int viscous, double out[3]);

/* ---- running statistics: MeanFlow (src/Metrics.jl:205-263).  wl_meanflow_init = the constructor / reset!
 *      (P, U and, with uu_stats, UU allocated and zeroed); wl_meanflow_update = the three loops of update!
 *      with ε = dt/(dt + time(meanflow) + eps(T)) formed by the caller, who owns the time vector `t` like
 *      the reference; wl_meanflow_copy_to_flow = copy!(flow, meanflow).  P/U/UU are read and written
 *      through wl_download / wl_upload (WL_MEAN_P, WL_MEAN_U, WL_MEAN_UU).                          */
int wl_meanflow_init(wl_sim* s, int uu_stats);
int wl_meanflow_update(wl_sim* s, float eps);
int wl_meanflow_copy_to_flow(wl_sim* s);

/* ---- asynchronous stepping for host-overlap: enqueue `nsteps` mom_steps without reading
 *      results back between them is not possible (Δt and the solver loop are host-steered in
 *      the reference); wl_sync waits for all queued device work.                               */
int wl_sync(wl_sim* s);

/* ---- measurement: how many 128x4x2-cell tiles of multigrid level `level` the last update! classified as
 *      0 general (coefficients L, iD loaded), 1 body-free (the pressure kernels use the level's constants instead of
 *      loading L / iD), 2 body-free next to a boundary or slab interface (constants with the boundary faces zeroed).
 *      Same values either way: results do not depend on the classes.                */
int wl_tile_classes(wl_sim* s, int32_t level, int64_t counts[3]);

/* ---- counters for benchmarks: kernels launched by this sim since creation, and per-phase
 *      device time of the last step when enabled by wl_enable_timing.                         */
int64_t wl_launch_count(const wl_sim* s);
int wl_enable_timing(wl_sim* s, int on);
/* phases: 0 mom predictor, 1 project(w=1), 2 mom corrector, 3 project(w=.5), 4 cfl; seconds */
int wl_last_phase_times(const wl_sim* s, double out[8]);

/* ---- measurement hooks (bench.py): per-kernel CUDA-event timing on the launching stream.
 *      wl_prof_begin(filter): time every kernel whose name contains `filter` ("" = all) until
 *      wl_prof_end, which writes "<kernel>\t<launches>\t<total_ms>\t<total_blocks>\n" lines into buf.  A filter that starts
 *      with '@' ("@", "@k_down") reports one line per (kernel, grid size) as "<kernel>@<blocks>", i.e. per multigrid level.
 *      wl_event_record / wl_event_elapsed_ms: events on the sim's own stream, slots 0..7.      */
int wl_prof_begin(wl_sim* s, const char* filter);
int wl_prof_end(wl_sim* s, char* buf, size_t buflen);
int wl_event_record(wl_sim* s, int slot);
int wl_event_elapsed_ms(wl_sim* s, int a, int b, float* ms);

/* ---- test hook: run ONE constituent operation of the step so that kernel-level parity tests can
 *      compare it with the oracle.  op: 0 BC!(u,uBC,exitBC) 1 conv_diff!+BDIM! predictor (u from u⁰)
 *      2 mom_project!(w=arg) 3 residual!(level) 4 Vcycle!(ω=arg) 5 smooth!(level,ω=arg)
 *      6 Jacobi!(level) 7 exitBC!(u,u⁰,Δt) 8 conv_diff!+BDIM!+scale corrector 9 u⁰ .= u
 *      10 pcg!(levels[1];it=arg)                                                                 */
int wl_debug_op(wl_sim* s, int op, int level, float arg);

#ifdef __cplusplus
}
#endif
#endif /* WL_B200_H */
