// wl_b200.cu -- host side of libwaterlily_b200.so: device state, kernel orchestration of
// mom_step! / solver! / update! / measure!, and the C ABI declared in include/wl_b200.h.
// There is no CPU fallback anywhere in this file: without an sm_100 device wl_create fails.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/wl_b200.h"
#include "wl_common.cuh"
#include "wl_kernels_body.cuh"
#include "wl_kernels_mg.cuh"
#include "wl_kernels_mg2.cuh"
#include "wl_kernels_mom.cuh"
#include "wl_kernels_mom3.cuh"

using namespace wl;

namespace {

std::string g_create_error;

struct DevField {
  float* alloc = nullptr;
  float* base = nullptr;   // element (0,0,0) of component 0
  Lay lay{};
  int ncomp = 0;
  size_t bytes = 0;
};

struct Level {
  Lay lay{};
  DevField L, x, z, r, eps, Dg, iD;
  bool owns = false;
  int c[3] = {0, 0, 0};
};

constexpr int NRED = 8;

struct ProfRec { const char* name; cudaEvent_t a, b; };

}  // namespace

struct wl_sim {
  wl_config cfg{};
  int D = 3;
  Lay lay{};
  Per per{};
  DevField u, u0, f, p, sigma, V, mu0, mu1;
  DevField w;                 // third rotating velocity buffer of the fused 3D K-MOM path
  int* band_list = nullptr;   // interior cells with μ₁ ≠ 0 (patched after K-MOM)
  int* band_count = nullptr;
  int band_cap = 0;
  bool band_dirty = false;
  std::vector<Level> lv;
  cudaStream_t st = nullptr;
  RedWork redw{};
  Red* d_red = nullptr;
  Red* h_red = nullptr;
  float dt = 0.25f;
  std::string err;
  int64_t launches = 0;
  int device = 0;
  bool timing = false;
  cudaEvent_t ev[8] = {};
  double phase_s[8] = {};
  int status = WL_OK;
  // profiling hooks
  bool prof_on = false;
  std::string prof_filter;
  std::vector<cudaEvent_t> evpool;
  std::vector<ProfRec> recs;
  cudaEvent_t uev[8] = {};
};

namespace {

#define WL_CUDA(s, call)                                                                   \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      (s)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                       \
      (s)->status = WL_ERR_CUDA;                                                           \
      return WL_ERR_CUDA;                                                                  \
    }                                                                                      \
  } while (0)

int alloc_field(wl_sim* s, DevField& f, const Lay& lay, int ncomp) {
  f.lay = lay; f.ncomp = ncomp;
  f.bytes = (size_t)lay.cstride * ncomp * sizeof(float) + 256;
  WL_CUDA(s, cudaMalloc(&f.alloc, f.bytes));
  WL_CUDA(s, cudaMemsetAsync(f.alloc, 0, f.bytes, s->st));
  f.base = f.alloc + LPAD;
  return WL_OK;
}
void free_field(DevField& f) { if (f.alloc) cudaFree(f.alloc); f.alloc = f.base = nullptr; }

dim3 blk(int D) {
  static int bx = 0, by = 4, bz = 2;
  if (!bx) {   // WL_BLK3="bx,by,bz" overrides the 3D interior tile (experiments); bx*by*bz must be 256, bz even
    bx = 32;
    if (const char* e = getenv("WL_BLK3")) { int a, b, c; if (sscanf(e, "%d,%d,%d", &a, &b, &c) == 3 && a * b * c == 256 && a % 32 == 0 && c % 2 == 0) { bx = a; by = b; bz = c; } }
  }
  return D == 3 ? dim3(bx, by, bz) : dim3(32, 8, 1);
}
dim3 grid_inside(const Lay& l) {
  dim3 b = blk(l.D);
  return dim3((l.n0 - 2 + b.x - 1) / b.x, (l.n1 - 2 + b.y - 1) / b.y, l.D == 3 ? (l.n2 - 2 + b.z - 1) / b.z : 1);
}
dim3 grid_all(const Lay& l) {
  dim3 b = blk(l.D);
  return dim3((l.n0 + LPAD + 31) / 32, (l.n1 + b.y - 1) / b.y, l.D == 3 ? (l.n2 + b.z - 1) / b.z : 1);
}
// 1-D persistent grid for the kernels built on for_tiles_inside / for_tiles_all: at most 8 blocks per SM
dim3 grid_tiles(const Lay& l, bool all) {
  const dim3 g = all ? grid_all(l) : grid_inside(l);
  const long long nt = (long long)g.x * g.y * g.z;
  return dim3((unsigned)(nt < 148 * 8 ? nt : 148 * 8));
}
// plane kernels: threads over (fast in-plane dim, slow in-plane dim); sized for the largest plane
void plane_launch(const Lay& l, dim3& g, dim3& b) {
  int m = l.n0 > l.n1 ? l.n0 : l.n1;
  int m2 = l.D == 3 ? (l.n1 > l.n2 ? l.n1 : l.n2) : 1;
  b = l.D == 3 ? dim3(32, 8, 1) : dim3(128, 1, 1);
  g = dim3((m + b.x - 1) / b.x, (m2 + b.y - 1) / b.y, 1);
}

// Optional per-kernel CUDA-event timing on the launching stream (wl_prof_begin/wl_prof_end).
inline bool prof_pre(wl_sim* s, const char* name, cudaEvent_t* a, cudaEvent_t* b) {
  if (!s->prof_on) return false;
  if (!s->prof_filter.empty() && !strstr(name, s->prof_filter.c_str())) return false;
  if (s->evpool.size() < 2) {
    for (int i = 0; i < 64; ++i) { cudaEvent_t e; cudaEventCreate(&e); s->evpool.push_back(e); }
  }
  *a = s->evpool.back(); s->evpool.pop_back();
  *b = s->evpool.back(); s->evpool.pop_back();
  cudaEventRecord(*a, s->st);
  return true;
}
inline void prof_post(wl_sim* s, const char* name, cudaEvent_t a, cudaEvent_t b) {
  cudaEventRecord(b, s->st);
  s->recs.push_back({name, a, b});
}

#define LAUNCH(s, kern, g, b, ...)                         \
  do {                                                     \
    cudaEvent_t pa_, pb_;                                  \
    const bool prof_ = prof_pre(s, #kern, &pa_, &pb_);     \
    kern<<<g, b, 0, (s)->st>>>(__VA_ARGS__);               \
    if (prof_) prof_post(s, #kern, pa_, pb_);              \
    ++(s)->launches;                                       \
  } while (0)

#define DISPATCH_D(s, kern, g, b, ...)                                    \
  do {                                                                    \
    if ((s)->D == 3) LAUNCH(s, kern<3>, g, b, __VA_ARGS__);               \
    else LAUNCH(s, kern<2>, g, b, __VA_ARGS__);                           \
  } while (0)

#define DISPATCH_DS(s, kern, g, b, ...)                                                   \
  do {                                                                                    \
    const int sc_ = (s)->cfg.scheme;                                                      \
    if ((s)->D == 3) {                                                                    \
      if (sc_ == 0) LAUNCH(s, (kern<3, 0>), g, b, __VA_ARGS__);                           \
      else if (sc_ == 1) LAUNCH(s, (kern<3, 1>), g, b, __VA_ARGS__);                      \
      else LAUNCH(s, (kern<3, 2>), g, b, __VA_ARGS__);                                    \
    } else {                                                                              \
      if (sc_ == 0) LAUNCH(s, (kern<2, 0>), g, b, __VA_ARGS__);                           \
      else if (sc_ == 1) LAUNCH(s, (kern<2, 1>), g, b, __VA_ARGS__);                      \
      else LAUNCH(s, (kern<2, 2>), g, b, __VA_ARGS__);                                    \
    }                                                                                     \
  } while (0)

// ---------------------------------------------------------------- boundary conditions
void bc_vec(wl_sim* s, DevField& a, const float* U, bool saveexit) {
  BcArgs args;
  args.l = a.lay; args.a = a.base; args.per = s->per; args.saveexit = saveexit ? 1 : 0;
  for (int d = 0; d < 3; ++d) args.U.v[d] = U ? U[d] : 0.f;
  dim3 g, b;
  plane_launch(a.lay, g, b);
  g.z = 3 * s->D;
  DISPATCH_D(s, k_bc_vec, g, b, args);
}
void perbc(wl_sim* s, const Lay& l, float* a) {
  for (int d = 0; d < s->D; ++d) {
    if (!s->per.p[d]) continue;
    PerArgs args{l, a, d};
    dim3 g, b;
    plane_launch(l, g, b);
    DISPATCH_D(s, k_perbc, g, b, args);
  }
}
void exitbc(wl_sim* s, DevField& u, DevField& u0, float dt) {
  ExitArgs args{u.lay, u.base, u0.base, dt};
  DISPATCH_D(s, k_exitbc, dim3(1), dim3(1024), args);
}

// ---------------------------------------------------------------- multigrid pieces
LevArgs lev_args(wl_sim* s, Level& p) {
  LevArgs a{};
  a.l = p.lay; a.L = p.L.base; a.x = p.x.base; a.r = p.r.base; a.eps = p.eps.base; a.z = p.z.base;
  a.omega = 1.f; a.k0 = 0; a.w = s->redw; a.out = s->d_red; a.in = s->d_red; a.want_norms = 0;
  return a;
}
void set_diag(wl_sim* s, Level& p) {
  DiagArgs a{p.lay, p.L.base, p.Dg.base, p.iD.base};
  DISPATCH_D(s, k_set_diag, grid_inside(p.lay), blk(s->D), a);
}
void increment(wl_sim* s, Level& p, float w, Red* norms) {
  perbc(s, p.lay, p.eps.base);
  LevArgs a = lev_args(s, p);
  a.omega = w; a.want_norms = norms ? 1 : 0; a.out = norms ? norms : s->d_red;
  DISPATCH_D(s, k_increment, grid_inside(p.lay), blk(s->D), a);
}
// Jacobi!(p) with it=1, ω=1 (Poisson.jl:111-114)
void jacobi(wl_sim* s, Level& p) {
  LevArgs a = lev_args(s, p);
  DISPATCH_D(s, k_eps_init, grid_inside(p.lay), blk(s->D), a);
  increment(s, p, 1.f, nullptr);
}
// GaussSeidelRB!(p;it=4,ω) (Poisson.jl:141-148)
void smooth(wl_sim* s, Level& p, float w, Red* norms) {
  LevArgs a = lev_args(s, p);
  DISPATCH_D(s, k_eps_init, grid_inside(p.lay), blk(s->D), a);
  perbc(s, p.lay, p.eps.base);
  for (int k0 = 1; k0 <= 4; ++k0) {
    a.k0 = k0;
    DISPATCH_D(s, k_gs_half, grid_inside(p.lay), blk(s->D), a);
  }
  increment(s, p, w, norms);
}
void restrict_r(wl_sim* s, Level& fi, Level& co) {
  XferArgs a{fi.lay, co.lay, fi.r.base, co.r.base, {co.c[0], co.c[1], co.c[2]}};
  DISPATCH_D(s, k_restrict, grid_inside(co.lay), blk(s->D), a);
}
void prolongate(wl_sim* s, Level& fi, Level& co) {
  XferArgs a{fi.lay, co.lay, co.x.base, fi.eps.base, {co.c[0], co.c[1], co.c[2]}};
  DISPATCH_D(s, k_prolongate, grid_inside(fi.lay), blk(s->D), a);
}
void restrictL(wl_sim* s, Level& fi, Level& co) {
  XferLArgs a{fi.lay, co.lay, fi.L.base, co.L.base, {co.c[0], co.c[1], co.c[2]}};
  DISPATCH_D(s, k_restrictL, grid_inside(co.lay), blk(s->D), a);
  bc_vec(s, co.L, nullptr, false);
}
// Vcycle!(ml;l,ω) (MultiLevelPoisson.jl:88-101), v1 kernels (kept for WL_FLAG_NO_FUSED)
void vcycle_v1(wl_sim* s, size_t l, float w) {
  Level &fine = s->lv[l], &coarse = s->lv[l + 1];
  jacobi(s, fine);
  restrict_r(s, fine, coarse);
  cudaMemsetAsync(coarse.x.alloc, 0, coarse.x.bytes, s->st);
  if (l + 2 < s->lv.size()) vcycle_v1(s, l + 1, w);
  smooth(s, coarse, w, nullptr);
  prolongate(s, fine, coarse);
  increment(s, fine, w, nullptr);
}

// ---- v2: fused kernels (wl_kernels_mg2.cuh)
LevDev levdev(const Level& p) {
  LevDev d{};
  d.l = p.lay; d.L = p.L.base; d.iD = p.iD.base; d.x = p.x.base; d.r = p.r.base; d.eps = p.eps.base;
  for (int q = 0; q < 3; ++q) d.c[q] = p.c[q];
  return d;
}
// one block of 1024 threads sweeps a level in (cells/1024) dependent round trips to L2 per phase:
// worthwhile up to ~18^3, beyond that the per-level launches win
inline bool level_is_small(const Level& p) { return (long long)p.lay.n0 * p.lay.n1 * p.lay.n2 <= 8000; }
// GaussSeidelRB!(p;it=4,ω): ϵ init + ghost images, 4 half sweeps on stale ghosts, wrap-read increment
void smooth2(wl_sim* s, Level& p, float w, Red* norms) {
  SmArgs a{levdev(p), s->per, w, 0, s->redw, norms ? norms : s->d_red, norms ? 1 : 0};
  DISPATCH_D(s, k_eps_init_g, grid_inside(p.lay), blk(s->D), a);
  LevArgs g = lev_args(s, p);
  for (int k0 = 1; k0 <= 4; ++k0) {
    g.k0 = k0;
    DISPATCH_D(s, k_gs_half, grid_inside(p.lay), blk(s->D), g);
  }
  DISPATCH_D(s, k_increment_w, grid_tiles(p.lay, false), blk(s->D), a);
}
// levels first..end in ONE single-block launch: Vcycle!(l=first) (if any coarser level) + smooth!(first)
void coarse_block(wl_sim* s, size_t first, float w) {
  CoarseArgs a{};
  for (size_t l = 0; l < s->lv.size() && l < (size_t)WL_MAXLEV; ++l) a.lv[l] = levdev(s->lv[l]);
  a.first = (int)first; a.nlev = (int)s->lv.size(); a.per = s->per; a.omega = w;
  DISPATCH_D(s, k_coarse, dim3(1), dim3(1024), a);
}
void vcycle(wl_sim* s, size_t l, float w) {
  if (s->cfg.flags & WL_FLAG_NO_FUSED) return vcycle_v1(s, l, w);
  Level &fine = s->lv[l], &coarse = s->lv[l + 1];
  {
    DownArgs a{levdev(fine), levdev(coarse), s->per};
    DISPATCH_D(s, k_down, grid_inside(fine.lay), blk(s->D), a);
    std::swap(fine.r, fine.eps);                     // K-DOWN staged the new residual in ϵ
  }
  if (level_is_small(coarse)) {
    coarse_block(s, l + 1, w);
  } else {
    if (l + 2 < s->lv.size()) vcycle(s, l + 1, w);
    smooth2(s, coarse, w, nullptr);
  }
  UpArgs u{levdev(fine), levdev(coarse), s->per, w, s->redw, s->d_red, 0};
  DISPATCH_D(s, k_up, grid_inside(fine.lay), blk(s->D), u);
}

int fetch_red(wl_sim* s) {
  WL_CUDA(s, cudaMemcpyAsync(s->h_red, s->d_red, sizeof(Red) * NRED, cudaMemcpyDeviceToHost, s->st));
  WL_CUDA(s, cudaStreamSynchronize(s->st));
  return WL_OK;
}

struct Trace { int n = 0; int flags = 0; float rinf[34], r1[34], omega[34]; };

// solver!(ml;tol,itmx) (MultiLevelPoisson.jl:108-128).  The loop exit and ω are steered by two
// scalars per iteration, exactly as in the reference.
int mlp_solver(wl_sim* s, Trace* tr, bool residual_done = false) {
  Level& p = s->lv[0];
  const double tol = s->cfg.tol;
  const double r1tol = (tol / 10) * (double)p.lay.ninside(), rinftol = tol;
  const int itmx = s->cfg.itmx;
  const bool fused = !(s->cfg.flags & WL_FLAG_NO_FUSED);
  {
    LevArgs a = lev_args(s, p);
    if (!residual_done) {                       // K-DIVRES already produced r and Σr in slot 0
      perbc(s, p.lay, p.x.base);
      a.out = s->d_red + 0;
      DISPATCH_D(s, k_residual, grid_inside(p.lay), blk(s->D), a);
    }
    a.in = s->d_red + 0; a.out = s->d_red + 1;
    DISPATCH_D(s, k_shift_norms, grid_tiles(p.lay, false), blk(s->D), a);
  }
  if (int e = fetch_red(s)) return e;
  float r1 = (float)s->h_red[1].sum, rinf = s->h_red[1].maxv, w = 1.f;
  int np = 0, flags = 0;
  if (tr) { tr->rinf[0] = rinf; tr->r1[0] = r1; tr->omega[0] = w; }
  while (np < itmx) {
    if (fused && level_is_small(p)) {            // the whole hierarchy fits the single-block kernel
      coarse_block(s, 0, w);
      LevArgs a = lev_args(s, p);
      a.in = s->d_red + 7; a.out = s->d_red + 1; // slot 7 stays zero: no shift, norms only
      DISPATCH_D(s, k_shift_norms, grid_tiles(p.lay, false), blk(s->D), a);
    } else {
      vcycle(s, 0, w);
      if (fused) smooth2(s, p, w, s->d_red + 1);
      else smooth(s, p, w, s->d_red + 1);
    }
    if (int e = fetch_red(s)) return e;
    const float rnew = (float)s->h_red[1].sum;
    rinf = s->h_red[1].maxv; ++np;
    if (tr && np < 34) { tr->rinf[np] = rinf; tr->r1[np] = rnew; tr->omega[np] = w; }
    if (rnew >= r1) w = (float)std::fmax(0.2, 0.9 * (double)w);
    else if (rnew < r1) w = (float)std::fmin(1.0, 1.02 * (double)w);
    r1 = rnew;
    if (!std::isfinite(rnew)) flags |= 1;
    if ((double)r1 < r1tol && (double)rinf < rinftol) break;
  }
  if (np >= itmx) flags |= 2;
  if (!residual_done) perbc(s, p.lay, p.x.base);   // fused projection wrap-reads x and stores p's images itself
  if (tr) { tr->n = np; tr->flags = flags; }
  return WL_OK;
}

// ---------------------------------------------------------------- momentum pieces
void mom_rhs(wl_sim* s, DevField& uadv, const float* acc) {
  MomArgs a{};
  a.l = s->lay; a.ua = uadv.base; a.u0 = s->u0.base; a.V = s->V.base; a.f = s->f.base;
  a.per = s->per; a.nu = s->cfg.nu; a.dt = s->dt;
  for (int d = 0; d < 3; ++d) a.acc.v[d] = acc ? acc[d] : 0.f;
  DISPATCH_DS(s, k_mom_rhs, grid_all(s->lay), blk(s->D), a);
}
// list of band cells for k_mom_patch; rebuilt when μ₁ changed (measure!, uploads)
int band_rebuild(wl_sim* s) {
  for (int pass = 0; pass < 2; ++pass) {
    WL_CUDA(s, cudaMemsetAsync(s->band_count, 0, sizeof(int), s->st));
    BandArgs a{s->lay, s->mu1.base, s->band_list, s->band_count, s->band_cap};
    DISPATCH_D(s, k_band_scan, grid_inside(s->lay), blk(s->D), a);
    int n = 0;
    WL_CUDA(s, cudaMemcpyAsync(&n, s->band_count, sizeof(int), cudaMemcpyDeviceToHost, s->st));
    WL_CUDA(s, cudaStreamSynchronize(s->st));
    if (n <= s->band_cap) break;
    if (s->band_list) cudaFree(s->band_list);
    s->band_cap = n + n / 4 + 1024;
    WL_CUDA(s, cudaMalloc(&s->band_list, sizeof(int) * (size_t)s->band_cap));
  }
  s->band_dirty = false;
  return WL_OK;
}
// K-MOM (3D fused): out = new velocity from ua (advecting) and u0
void mom3(wl_sim* s, DevField& ua, DevField& u0, DevField& out, int corrector, const float* acc) {
  Mom3Args a{};
  a.l = s->lay; a.ua = ua.base; a.u0 = u0.base; a.V = s->V.base; a.mu0 = s->mu0.base; a.un = out.base;
  a.per = s->per; a.nu = s->cfg.nu; a.dt = s->dt; a.corrector = corrector;
  for (int d = 0; d < 3; ++d) a.acc.v[d] = acc ? acc[d] : 0.f;
  const Lay& l = s->lay;
  // march length: as long as possible (amortises the extra flux plane) while keeping >= 6 waves of
  // 2 blocks/SM on 148 SMs
  a.KZ = 64;
  while (a.KZ > 8 && (long long)((l.n0 - 2 + 30) / 31) * ((l.n1 - 2 + M3_TY - 2) / (M3_TY - 1)) *
                             ((l.n2 - 2 + a.KZ - 1) / a.KZ) < 148 * 2 * 6)
    a.KZ >>= 1;
  dim3 g((l.n0 - 2 + 30) / 31, (l.n1 - 2 + M3_TY - 2) / (M3_TY - 1), (l.n2 - 2 + a.KZ - 1) / a.KZ);
  dim3 b(32, M3_TY, 1);
  const int sc = s->cfg.scheme;
  if (sc == 0) LAUNCH(s, k_mom3<0>, g, b, a);
  else if (sc == 1) LAUNCH(s, k_mom3<1>, g, b, a);
  else LAUNCH(s, k_mom3<2>, g, b, a);
  PatchArgs pa{};
  pa.m.l = s->lay; pa.m.ua = ua.base; pa.m.u0 = u0.base; pa.m.V = s->V.base; pa.m.f = nullptr;
  pa.m.per = s->per; pa.m.nu = s->cfg.nu; pa.m.dt = s->dt; pa.m.acc = a.acc;
  pa.mu0 = s->mu0.base; pa.mu1 = s->mu1.base; pa.un = out.base;
  pa.list = s->band_list; pa.count = s->band_count; pa.cap = s->band_cap; pa.corrector = corrector;
  if (s->band_cap > 0) {
    if (sc == 0) LAUNCH(s, k_mom_patch<0>, dim3(1184), dim3(128), pa);
    else if (sc == 1) LAUNCH(s, k_mom_patch<1>, dim3(1184), dim3(128), pa);
    else LAUNCH(s, k_mom_patch<2>, dim3(1184), dim3(128), pa);
  }
  if (s->cfg.exitBC) {
    ExitCopyArgs e{s->lay, ua.base, out.base};
    LAUNCH(s, k_copy_exit_plane, dim3((l.n1 + 127) / 128, l.n2), dim3(128), e);
  }
}
void bdim(wl_sim* s, int corrector) {
  BdimArgs a{s->lay, s->u.base, s->f.base, s->V.base, s->mu0.base, s->mu1.base, corrector};
  DISPATCH_D(s, k_bdim, grid_inside(s->lay), blk(s->D), a);
}
// mom_project!(a,b,w,t) (Flow.jl:223-232)
int mom_project(wl_sim* s, float w, const float* U, Trace* tr) {
  if (!(s->cfg.flags & WL_FLAG_NO_FUSED)) {
    Level& p0 = s->lv[0];
    const float dt = w * s->dt;
    DivResArgs d{s->lay, s->u.base, s->p.base, p0.eps.base, s->mu0.base, p0.r.base, s->per, dt, s->redw, s->d_red + 0};
    DISPATCH_D(s, k_divres, grid_tiles(s->lay, false), blk(s->D), d);
    std::swap(s->p, p0.eps); p0.x = s->p;                       // scaled x lives in the former ϵ buffer
    if (int e = mlp_solver(s, tr, true)) return e;
    ProjArgs2 q{s->lay, s->u.base, s->p.base, p0.eps.base, s->mu0.base, s->per, dt};
    DISPATCH_D(s, k_proj, grid_inside(s->lay), blk(s->D), q);
    std::swap(s->p, p0.eps); p0.x = s->p;
    bc_vec(s, s->u, U, s->cfg.exitBC != 0);
    return WL_OK;
  }
  ProjArgs a{s->lay, s->u.base, s->p.base, s->sigma.base, s->mu0.base, w * s->dt};
  DISPATCH_D(s, k_div_scale, grid_all(s->lay), blk(s->D), a);
  if (int e = mlp_solver(s, tr)) return e;
  DISPATCH_D(s, k_project, grid_inside(s->lay), blk(s->D), a);
  DISPATCH_D(s, k_unscale, grid_all(s->lay), blk(s->D), a);
  bc_vec(s, s->u, U, s->cfg.exitBC != 0);
  return WL_OK;
}
int cfl(wl_sim* s, float* out) {
  CflArgs a{s->lay, s->u.base, s->sigma.base, s->redw, s->d_red + 2};
  DISPATCH_D(s, k_cfl, grid_tiles(s->lay, true), blk(s->D), a);
  if (int e = fetch_red(s)) return e;
  const float m = s->h_red[2].maxv;
  *out = std::fmin(10.f, 1.f / (m + 5.f * s->cfg.nu));
  return WL_OK;
}
void ghost_phi(wl_sim* s) {
  GhostPhiArgs a{s->lay, s->u.base, s->sigma.base, s->per, s->cfg.nu};
  dim3 g, b;
  plane_launch(s->lay, g, b);
  g.z = s->D;
  DISPATCH_DS(s, k_ghost_phi, g, b, a);
}

BodyDev body_at(const wl_body* b, int D, float t) {
  BodyDev d{};
  d.kind = b->kind; d.axis = b->axis; d.radius = b->radius; d.minor = b->minor;
  for (int k = 0; k < 3; ++k) { d.c[k] = 0.f; d.Vb[k] = 0.f; }
  for (int k = 0; k < D; ++k) {
    d.c[k] = b->center[k] + b->vel[k] * t + b->acc[k] * t * t + b->amp[k] * std::sin(b->omega * t);
    d.Vb[k] = b->vel[k] + 2 * b->acc[k] * t + b->amp[k] * b->omega * std::cos(b->omega * t);
  }
  return d;
}

DevField* flow_field(wl_sim* s, int field, int level, Lay* lay) {
  DevField* f = nullptr;
  if (field < 10) {
    switch (field) {
      case WL_U: f = &s->u; break;
      case WL_U0: f = &s->u0; break;
      case WL_F: f = &s->f; break;
      case WL_P: f = &s->p; break;
      case WL_SIGMA: f = &s->sigma; break;
      case WL_V: f = &s->V; break;
      case WL_MU0: f = &s->mu0; break;
      case WL_MU1: f = &s->mu1; break;
      default: return nullptr;
    }
  } else {
    if (level < 0 || level >= (int)s->lv.size()) return nullptr;
    Level& l = s->lv[level];
    switch (field) {
      case WL_X: f = &l.x; break;
      case WL_R: f = &l.r; break;
      case WL_EPS: f = &l.eps; break;
      case WL_Z: f = &l.z; break;
      case WL_L: f = &l.L; break;
      case WL_DIAG: f = &l.Dg; break;
      case WL_IDIAG: f = &l.iD; break;
      default: return nullptr;
    }
  }
  if (lay) *lay = f->lay;
  return f;
}

inline bool divisible(int N) { return N % 2 == 0 && N > 4; }

}  // namespace

#include "wl_abi.cuh"
