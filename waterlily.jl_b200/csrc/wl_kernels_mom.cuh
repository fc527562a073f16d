// wl_kernels_mom.cuh -- momentum-side kernels: conv_diff!+accelerate!+BDIM! (RHS and update),
// BC!, exitBC!, divergence, projection, CFL.  Reference: src/Flow.jl, src/core.jl:192-243.
#pragma once
#include "wl_common.cuh"

namespace wl {

// ---------------------------------------------------------------- convective schemes (Flow.jl:4-6,27-36)
__device__ __forceinline__ float median3(float a, float b, float c) {
  if (a > b) { if (b >= c) return b; if (a > c) return c; }
  else       { if (b <= c) return b; if (a < c) return c; }
  return a;
}
template <int S>
__device__ __forceinline__ float lam(float u, float c, float d) {
  if (S == 0) return median3((5.f * c + 2.f * d - u) / 6.f, c, median3(10.f * c - 9.f * u, c, d));
  if (S == 1) return (c + d) / 2.f;
  return (c <= fminf(u, d) || c >= fmaxf(u, d)) ? c : c + (d - c) * (c - u) / (d - u);
}

// Φ on the lower j-face of the cell at offset o whose j-coordinate is cj (1..Nj-1):
// ϕu / ϕuL / ϕuP / ϕuR minus ν∂ (Flow.jl:8-11,47,56-61).  `perj` selects the periodic variant
// for cj==1; cj==Nj-1 is only evaluated for non-periodic j.
template <int S>
__device__ __forceinline__ float face_flux(const float* __restrict__ ui, const float* __restrict__ uj,
                                           int o, int si, int sj, int cj, int Nj, bool perj, float nu) {
  const float ua = (uj[o] + uj[o - si]) / 2.f;
  const float c0 = ui[o], cm = ui[o - sj];
  const float dif = nu * (c0 - cm);
  float f;
  if (cj == 1) {
    if (ua > 0.f) f = perj ? ua * lam<S>(ui[o + sj * (Nj - 4)], cm, c0) : ua * ((c0 + cm) / 2.f);
    else f = ua * lam<S>(ui[o + sj], c0, cm);
  } else if (cj == Nj - 1) {
    if (ua < 0.f) f = ua * ((c0 + cm) / 2.f);
    else f = ua * lam<S>(ui[o - 2 * sj], cm, c0);
  } else {
    if (ua > 0.f) f = ua * lam<S>(ui[o - 2 * sj], cm, c0);
    else f = ua * lam<S>(ui[o + sj], c0, cm);
  }
  return f - dif;
}

// Slab interfaces (multi-GPU): a z-face next to an interface is an ordinary inner face whose
// stencil reaches into the halo planes, so the boundary variants must not be selected there.
__device__ __forceinline__ void zface_eff(const Per& per, int dj, int& cj, int& Nj) {
  if (dj == 2 && ((per.zlo && cj <= 1) || (per.zhi && cj >= Nj - 1))) { cj = 2; Nj = 1 << 28; }
}

struct MomArgs {
  Lay l;
  const float* ua;   // advecting/advected velocity (u⁰ in the predictor, u in the corrector)
  const float* u0;   // u⁰
  const float* V;
  float* f;
  Per per;
  float nu, dt;
  Vec3 acc;          // accelerate!: space-uniform g+dU/dt
};

// K-RHS: conv_diff! (Flow.jl:38-62) + accelerate! (:69-73) + first loop of BDIM! (:178):
//   f[I,i] = u⁰[I,i] + Δt·(r[I,i]+acc_i) − V[I,i]  over ALL indices,
// with r accumulated exactly in the reference's order (for j: +Φ_low, −Φ_up) and r==0 on cells
// the reference never writes (any coordinate on the lower ghost).  One thread per cell.
// f[I,ci] for ANY cell (ghosts included) straight from global memory, in the reference's order.
// ACCF adds the caller's acceleration field (g(i,x,t)+∂ₜuBC(i,x,t) sampled on every index, wl_set_accel_field)
// after the uniform term, as accelerate! does (Flow.jl:69-73); the default instantiation is unchanged.
template <int D, int S, bool ACCF = false>
__device__ __forceinline__ float rhs_f(const MomArgs& a, int ci, int c0, int c1, int c2, const float* accf = nullptr) {
  const Lay& l = a.l;
  const int o = l.off(c0, c1, c2);
  const int c[3] = {c0, c1, c2};
  const int N[3] = {l.n0, l.n1, l.n2};
  const int st[3] = {1, l.pitch, l.plane};
  bool lowghost = false;
#pragma unroll
  for (int d = 0; d < D; ++d) lowghost |= (c[d] == 0) && !(d == 2 && a.per.zlo);
  const float* ui = a.ua + l.cstride * ci;
  float r = 0.f;
  if (!lowghost) {
#pragma unroll
    for (int dj = 0; dj < D; ++dj) {
      if (c[dj] > N[dj] - 2 && !(dj == 2 && a.per.zhi)) continue;   // upper ghost in j: no j-contribution
      const float* uj = a.ua + l.cstride * dj;
      const bool pj = a.per.p[dj] != 0;
      int cl = c[dj], Nl = N[dj], cu = c[dj] + 1, Nu = N[dj];
      zface_eff(a.per, dj, cl, Nl);
      zface_eff(a.per, dj, cu, Nu);
      r += face_flux<S>(ui, uj, o, st[ci], st[dj], cl, Nl, pj, a.nu);
      if (c[dj] == N[dj] - 2 && pj)                  // Φ[CIj(j,I,2)] Flow.jl:62
        r -= face_flux<S>(ui, uj, o - st[dj] * (N[dj] - 3), st[ci], st[dj], 1, N[dj], true, a.nu);
      else
        r -= face_flux<S>(ui, uj, o + st[dj], st[ci], st[dj], cu, Nu, pj, a.nu);
    }
  }
  const long long oc = l.cstride * ci + o;
  // accelerate! adds g + ∂ₜU as ONE term (Flow.jl:73): uniform part and sampled field are summed first
  if (ACCF) r += a.acc.v[ci] + accf[oc];
  else r += a.acc.v[ci];
  return a.u0[oc] + a.dt * r - a.V[oc];
}

template <int D, int S>
__global__ void __launch_bounds__(256) k_mom_rhs(const MomArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;   // lane 31 of block 0 is i==0; i==1 starts a line
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  const int o = l.off(i, j, k);
#pragma unroll
  for (int ci = 0; ci < D; ++ci) a.f[l.cstride * ci + o] = rhs_f<D, S>(a, ci, i, j, k);
}

template <int D, int S>
__global__ void __launch_bounds__(256) k_mom_rhs_acc(const MomArgs a, const float* accf) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  const int o = l.off(i, j, k);
#pragma unroll
  for (int ci = 0; ci < D; ++ci) a.f[l.cstride * ci + o] = rhs_f<D, S, true>(a, ci, i, j, k, accf);
}

struct BdimArgs {
  Lay l;
  float* u;
  const float* f;
  const float* V;
  const float* mu0;
  const float* mu1;
  int corrector;     // 0: u = X ; 1: u = (u + X)·½   (scale_u! Flow.jl:209,211-214)
};

// K-BDIM: second loop of BDIM! (Flow.jl:179) over inside_u, fused with scale_u!.
template <int D>
__global__ void __launch_bounds__(256) k_bdim(const BdimArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z + 1 : 0;
  if (i > l.n0 - 2 || j > l.n1 - 2 || (D == 3 && k > l.n2 - 2)) return;
  const int o = l.off(i, j, k);
  const int st[3] = {1, l.pitch, l.plane};
#pragma unroll
  for (int ci = 0; ci < D; ++ci) {
    const long long oc = l.cstride * ci + o;
    const float* fc = a.f + l.cstride * ci;
    float s = 0.f;
#pragma unroll
    for (int dj = 0; dj < D; ++dj) {
      const float m = a.mu1[l.cstride * (ci + D * dj) + o];
      s += m * (fc[o + st[dj]] - fc[o - st[dj]]);       // μddn Flow.jl:20-26
    }
    const float X = s / 2.f + a.V[oc] + a.mu0[oc] * fc[o];
    a.u[oc] = a.corrector ? (a.u[oc] + X) * 0.5f : X;
  }
}

// ---------------------------------------------------------------- BC! (core.jl:200-219), fused.
// The reference fills ghosts with D² sequential passes (component i outer, direction j inner);
// corner/edge values depend on that order.  Here every written cell resolves its final value in
// one kernel by walking the passes backwards: the value a pass-j ghost receives is the value its
// source cell had after passes < j.  The walk always ends either at a Dirichlet constant or at a
// cell no pass writes, so reading in place is race-free.
struct BcArgs {
  Lay l;
  float* a;
  Per per;
  Vec3 U;
  int saveexit;
};

template <int D>
__device__ __forceinline__ float bc_resolve(const BcArgs& a, int ci, int c0, int c1, int c2) {
  int c[3] = {c0, c1, c2};
  const int N[3] = {a.l.n0, a.l.n1, a.l.n2};
#pragma unroll
  for (int dj = D - 1; dj >= 0; --dj) {
    const int x = c[dj], n = N[dj];
    const bool lo_if = dj == 2 && a.per.zlo, hi_if = dj == 2 && a.per.zhi;   // slab interface: no pass writes there
    if (a.per.p[dj]) {
      if (x == 0) c[dj] = n - 2; else if (x == n - 1) c[dj] = 1;
    } else if (ci == dj) {
      if ((x <= 1 && !lo_if) || (x == n - 1 && !hi_if && !(a.saveexit && ci == 0))) return a.U.v[ci];
    } else {
      if (x == 0 && !lo_if) c[dj] = 1; else if (x == n - 1 && !hi_if) c[dj] = n - 2;
    }
  }
  return a.a[a.l.cstride * ci + a.l.off(c[0], c[1], c[2])];
}

// grid.z enumerates (direction dj, which plane: 0 -> index 0, 1 -> index 1, 2 -> index n-1);
// threads cover the plane's other two coordinates (fastest remaining dim on threadIdx.x).
template <int D>
__global__ void k_bc_vec(const BcArgs a) {
  pdl_entry();
  const int dj = blockIdx.z / 3, which = blockIdx.z % 3;
  const Lay& l = a.l;
  const int N[3] = {l.n0, l.n1, l.n2};
  if (dj == 2 && (which == 2 ? a.per.zhi : a.per.zlo)) return;   // slab interface: filled by the halo exchange
  const int da = dj == 0 ? 1 : 0;                 // fast in-plane dim
  const int db = dj == 2 ? 1 : 2;                 // slow in-plane dim (unused in 2D)
  const int pa = blockIdx.x * blockDim.x + threadIdx.x;
  const int pb = blockIdx.y * blockDim.y + threadIdx.y;
  if (pa >= N[da] || (D == 3 ? pb >= N[db] : pb > 0)) return;
  int c[3] = {0, 0, 0};
  c[dj] = which == 0 ? 0 : which == 1 ? 1 : N[dj] - 1;
  c[da] = pa;
  if (D == 3) c[db] = pb;
#pragma unroll
  for (int ci = 0; ci < D; ++ci) {
    if (which == 1 && (ci != dj || a.per.p[dj])) continue;         // index 1 is only written by Dirichlet
    if (which == 2 && ci == dj && !a.per.p[dj] && a.saveexit && ci == 0) {
      // exit plane kept (saveexit): still rewritten by later passes on its edges
      bool edge = false;
      for (int d = 0; d < D; ++d) if (d != dj && (c[d] == 0 || c[d] == N[d] - 1)) edge = true;
      if (!edge) continue;
    }
    const float v = bc_resolve<D>(a, ci, c[0], c[1], c[2]);
    a.a[l.cstride * ci + l.off(c[0], c[1], c[2])] = v;
  }
}

// BC! with a Function uBC(i,x,t) (core.jl:201-219) sampled by the caller into tables (wl_set_bc_tables): per
// (component ci, direction dj) three planes over the other dims (ghosts included, lowest remaining dim fastest):
//   ci == dj: Dirichlet values at index 0, 1 and n-1;   ci != dj: uBC(I)-uBC(I±δ) at index 0, unused, at index n-1.
// Same backward walk as bc_resolve; the tangential differences met on the way are added afterwards in pass order
// (lowest direction first), which is what the reference's sequential passes compute.
struct BcTabArgs {
  Lay l;
  float* a;
  Per per;
  int saveexit;
  const float* tab;
  int off[9];      // start of the 3 planes of pair ci*D+dj
  int psz[3];      // plane size of direction dj
};

template <int D>
__device__ __forceinline__ float bc_resolve_tab(const BcTabArgs& a, int ci, int c0, int c1, int c2) {
  int c[3] = {c0, c1, c2};
  const int N[3] = {a.l.n0, a.l.n1, a.l.n2};
  float ds[2] = {0.f, 0.f};
  int nd = 0;
  float base = 0.f;
  bool have = false;
#pragma unroll
  for (int dj = D - 1; dj >= 0; --dj) {
    if (have) continue;
    const int x = c[dj], n = N[dj];
    const bool lo_if = dj == 2 && a.per.zlo, hi_if = dj == 2 && a.per.zhi;   // slab interface: no pass writes there
    if (a.per.p[dj]) {
      if (x == 0) c[dj] = n - 2; else if (x == n - 1) c[dj] = 1;
      continue;
    }
    const int ia = dj == 0 ? 1 : 0, ib = dj == 2 ? 1 : 2;       // in-plane dims, fast and slow
    const int idx = c[ia] + (D == 3 ? N[ia] * c[ib] : 0);
    const float* T = a.tab + a.off[ci * D + dj];
    const int ps = a.psz[dj];
    if (ci == dj) {
      if (x <= 1 && !lo_if) { base = T[(x == 0 ? 0 : ps) + idx]; have = true; }
      else if (x == n - 1 && !hi_if && !(a.saveexit && ci == 0)) { base = T[2 * ps + idx]; have = true; }
    } else {
      if (x == 0 && !lo_if) { ds[nd++] = T[idx]; c[dj] = 1; }
      else if (x == n - 1 && !hi_if) { ds[nd++] = T[2 * ps + idx]; c[dj] = n - 2; }
    }
  }
  if (!have) base = a.a[a.l.cstride * ci + a.l.off(c[0], c[1], c[2])];
  if (nd == 2) base = base + ds[1];
  if (nd >= 1) base = base + ds[0];
  return base;
}

// launch shape and written-cell rules of k_bc_vec
template <int D>
__global__ void k_bc_vec_tab(const BcTabArgs a) {
  pdl_entry();
  const int dj = blockIdx.z / 3, which = blockIdx.z % 3;
  const Lay& l = a.l;
  const int N[3] = {l.n0, l.n1, l.n2};
  if (dj == 2 && (which == 2 ? a.per.zhi : a.per.zlo)) return;   // slab interface: filled by the halo exchange
  const int da = dj == 0 ? 1 : 0;
  const int db = dj == 2 ? 1 : 2;
  const int pa = blockIdx.x * blockDim.x + threadIdx.x;
  const int pb = blockIdx.y * blockDim.y + threadIdx.y;
  if (pa >= N[da] || (D == 3 ? pb >= N[db] : pb > 0)) return;
  int c[3] = {0, 0, 0};
  c[dj] = which == 0 ? 0 : which == 1 ? 1 : N[dj] - 1;
  c[da] = pa;
  if (D == 3) c[db] = pb;
#pragma unroll
  for (int ci = 0; ci < D; ++ci) {
    if (which == 1 && (ci != dj || a.per.p[dj])) continue;
    if (which == 2 && ci == dj && !a.per.p[dj] && a.saveexit && ci == 0) {
      bool edge = false;
      for (int d = 0; d < D; ++d) if (d != dj && (c[d] == 0 || c[d] == N[d] - 1)) edge = true;
      if (!edge) continue;
    }
    const float v = bc_resolve_tab<D>(a, ci, c[0], c[1], c[2]);
    a.a[l.cstride * ci + l.off(c[0], c[1], c[2])] = v;
  }
}

// perBC! (core.jl:239-243) for one direction; launched once per periodic direction, in order.
struct PerArgs { Lay l; float* a; int dj; };
template <int D>
__global__ void k_perbc(const PerArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int N[3] = {l.n0, l.n1, l.n2};
  const int dj = a.dj;
  const int da = dj == 0 ? 1 : 0, db = dj == 2 ? 1 : 2;
  const int pa = blockIdx.x * blockDim.x + threadIdx.x;
  const int pb = blockIdx.y * blockDim.y + threadIdx.y;
  if (pa >= N[da] || (D == 3 ? pb >= N[db] : pb > 0)) return;
  int c[3] = {0, 0, 0};
  c[da] = pa;
  if (D == 3) c[db] = pb;
  const int st = l.stride(dj), n = N[dj];
  const int o = l.off(c[0], c[1], c[2]);
  a.a[o] = a.a[o + st * (n - 2)];
  a.a[o + st * (n - 1)] = a.a[o + st];
}

// ---------------------------------------------------------------- exitBC! (core.jl:226-233)
// One block: plane mean, 1-D convection, plane mean, flux correction (sums in double, fixed order).
struct ExitArgs { Lay l; float* u; const float* u0; float dt; };
template <int D>
__global__ void __launch_bounds__(1024) k_exitbc(const ExitArgs a) {
  pdl_entry();
  __shared__ double sh[32];
  __shared__ float bcast;
  const Lay& l = a.l;
  const int m1 = l.n1 - 2, m2 = D == 3 ? l.n2 - 2 : 1;
  const int len = m1 * m2;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  auto block_sum = [&](double v) -> double {
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[wid] = v;
    __syncthreads();
    double t = 0;
    if (wid == 0) { t = lane < (blockDim.x >> 5) ? sh[lane] : 0.0; t = warp_sum(t); }
    return t;   // valid in thread 0
  };
  double s = 0;
  for (int q = tid; q < len; q += blockDim.x) {
    const int j = 1 + q % m1, k = D == 3 ? 1 + q / m1 : 0;
    s += (double)a.u[l.off(1, j, k)];
  }
  s = block_sum(s);
  if (tid == 0) bcast = (float)s / (float)len;
  __syncthreads();
  const float U = bcast;
  s = 0;
  for (int q = tid; q < len; q += blockDim.x) {
    const int j = 1 + q % m1, k = D == 3 ? 1 + q / m1 : 0;
    const int o = l.off(l.n0 - 1, j, k);
    const float v = a.u0[o] - U * a.dt * (a.u0[o] - a.u0[o - 1]);
    a.u[o] = v;
    s += (double)v;
  }
  s = block_sum(s);
  __syncthreads();
  if (tid == 0) bcast = (float)s / (float)len - U;
  __syncthreads();
  const float fl = bcast;
  for (int q = tid; q < len; q += blockDim.x) {
    const int j = 1 + q % m1, k = D == 3 ? 1 + q / m1 : 0;
    a.u[l.off(l.n0 - 1, j, k)] -= fl;
  }
}

// exitBC! across slabs: every rank owns a z-range of both x-planes, so the two plane means become
// all-gathered partial sums (folded in rank order) and the kernel splits in three.
// stage 0: Σ u[2,J] -> out;  stage 1: convect the exit plane with U from in0, Σ -> out;
// stage 2: subtract the flux defect (in0: inflow sums, in1: exit sums).  `len` = cells of the GLOBAL plane.
struct ExitDistArgs { Lay l; float* u; const float* u0; float dt; int stage; long long len; RedIn in0, in1; Red* out; };
__global__ void __launch_bounds__(1024) k_exitbc_dist(const ExitDistArgs a) {
  pdl_entry();
  __shared__ double sh[32];
  const Lay& l = a.l;
  const int m1 = l.n1 - 2, m2 = l.n2 - 2;
  const int len = m1 * m2;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const float U = a.stage > 0 ? (float)redin_sum(a.in0) / (float)a.len : 0.f;
  const float fl = a.stage == 2 ? (float)redin_sum(a.in1) / (float)a.len - U : 0.f;
  double s = 0;
  for (int q = tid; q < len; q += blockDim.x) {
    const int j = 1 + q % m1, k = 1 + q / m1;
    if (a.stage == 0) {
      s += (double)a.u[l.off(1, j, k)];
    } else if (a.stage == 1) {
      const int o = l.off(l.n0 - 1, j, k);
      const float v = a.u0[o] - U * a.dt * (a.u0[o] - a.u0[o - 1]);
      a.u[o] = v;
      s += (double)v;
    } else {
      a.u[l.off(l.n0 - 1, j, k)] -= fl;
    }
  }
  if (a.stage == 2) return;
  s = warp_sum(s);
  if (lane == 0) sh[wid] = s;
  __syncthreads();
  if (wid == 0) {
    double t = lane < (blockDim.x >> 5) ? sh[lane] : 0.0;
    t = warp_sum(t);
    if (lane == 0) { a.out->sum = t; a.out->maxv = -INFINITY; }
  }
}

// ---------------------------------------------------------------- projection pieces (Flow.jl:223-232)
struct ProjArgs { Lay l; float* u; float* x; float* z; const float* L; float dt; };

// z = div(u) inside; x *= dt everywhere (Flow.jl:225)
template <int D>
__global__ void __launch_bounds__(256) k_div_scale(const ProjArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  const int o = l.off(i, j, k);
  a.x[o] *= a.dt;
  const bool in = i >= 1 && i <= l.n0 - 2 && j >= 1 && j <= l.n1 - 2 && (D == 2 || (k >= 1 && k <= l.n2 - 2));
  if (!in) return;
  const int st[3] = {1, l.pitch, l.plane};
  float s = 0.f;
#pragma unroll
  for (int d = 0; d < D; ++d) { const float* ud = a.u + l.cstride * d; s += ud[o + st[d]] - ud[o]; }
  a.z[o] = s;
}
// u[I,i] -= L[I,i]·∂ᵢx inside (Flow.jl:227-229)
template <int D>
__global__ void __launch_bounds__(256) k_project(const ProjArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z + 1 : 0;
  if (i > l.n0 - 2 || j > l.n1 - 2 || (D == 3 && k > l.n2 - 2)) return;
  const int o = l.off(i, j, k);
  const int st[3] = {1, l.pitch, l.plane};
  const float xc = a.x[o];
#pragma unroll
  for (int d = 0; d < D; ++d) {
    const long long oc = l.cstride * d + o;
    a.u[oc] -= a.L[oc] * (xc - a.x[o - st[d]]);
  }
}
// x ./= dt everywhere (Flow.jl:230)
template <int D>
__global__ void __launch_bounds__(256) k_unscale(const ProjArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  a.x[l.off(i, j, k)] /= a.dt;
}

// ---------------------------------------------------------------- CFL (Flow.jl:234-244)
// σ = flux_out inside; max over the WHOLE σ array (ghosts keep the stale conv-diff Φ values
// k_ghost_phi wrote, as in the reference).
struct CflArgs { Lay l; const float* u; float* sigma; RedWork w; Red* out; int zlo, zhi; };
template <int D>
__global__ void __launch_bounds__(256) k_cfl(const CflArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  float m = -INFINITY;
  // 1-D persistent grid over ALL cells, four tiles per trip with their loads issued together (see
  // k_shift_norms): the loop is otherwise latency-bound
  const int nbx = (l.n0 + LPAD + blockDim.x - 1) / blockDim.x, nby = (l.n1 + blockDim.y - 1) / blockDim.y;
  const int nbz = D == 3 ? (l.n2 + blockDim.z - 1) / blockDim.z : 1;
  const int nt = nbx * nby * nbz;
  const int st[3] = {1, l.pitch, l.plane};
  for (int t0 = blockIdx.x; t0 < nt; t0 += 4 * gridDim.x) {
    int o[4]; int cls[4];          // 0 skip, 1 interior, 2 ghost that takes part in the max
    float lo[4][3], hi[4][3], sg[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int t = t0 + u * gridDim.x;
      const int tx = t % nbx, q = t / nbx, ty = q % nby, tz = q / nby;
      const int i = tx * blockDim.x + threadIdx.x - LPAD, j = ty * blockDim.y + threadIdx.y;
      const int k = D == 3 ? tz * blockDim.z + threadIdx.z : 0;
      const bool have = t < nt && i >= 0 && i < l.n0 && j < l.n1 && k < l.n2;
      const bool in = i >= 1 && i <= l.n0 - 2 && j >= 1 && j <= l.n1 - 2 && (D == 2 || (k >= 1 && k <= l.n2 - 2));
      // slab interfaces: the planes beyond them belong to the neighbour
      const bool mine = !(D == 3 && ((a.zlo && k == 0) || (a.zhi && k == l.n2 - 1)));
      cls[u] = !have ? 0 : (in ? 1 : (mine ? 2 : 0));
      o[u] = have ? l.off(i, j, k) : l.off(1, 1, D == 3 ? 1 : 0);
      const int oi = cls[u] == 1 ? o[u] : l.off(1, 1, D == 3 ? 1 : 0);
#pragma unroll
      for (int d = 0; d < D; ++d) {
        const float* ud = a.u + l.cstride * d;
        lo[u][d] = ud[oi]; hi[u][d] = ud[oi + st[d]];
      }
      sg[u] = cls[u] == 2 ? a.sigma[o[u]] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (cls[u] == 1) {
        float s = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) s += fmaxf(0.f, hi[u][d]) + fmaxf(0.f, -lo[u][d]);
        a.sigma[o[u]] = s;
        m = fmaxf(m, s);
      } else if (cls[u] == 2) {
        m = fmaxf(m, sg[u]);
      }
    }
  }
  block_reduce_finish(0.0, m, a.w, a.out);
}

// Stale Φ left in σ's upper-ghost cells by the corrector's conv_diff! (Flow.jl:47, ranges core.jl:55-57):
// the last pass (component D, the largest direction j whose range covers the cell) wins.
struct GhostPhiArgs { Lay l; const float* u; float* sigma; Per per; float nu; };
template <int D, int S>
__global__ void k_ghost_phi(const GhostPhiArgs a) {
  pdl_entry();
  const int dg = blockIdx.z;                       // plane: coordinate dg == N-1
  const Lay& l = a.l;
  const int N[3] = {l.n0, l.n1, l.n2};
  const int da = dg == 0 ? 1 : 0, db = dg == 2 ? 1 : 2;
  const int pa = blockIdx.x * blockDim.x + threadIdx.x;
  const int pb = blockIdx.y * blockDim.y + threadIdx.y;
  if (pa >= N[da] || (D == 3 ? pb >= N[db] : pb > 0)) return;
  int c[3] = {0, 0, 0};
  c[dg] = N[dg] - 1; c[da] = pa; if (D == 3) c[db] = pb;
  if (D == 3 && dg == 2 && a.per.zhi) return;                  // not a ghost plane of the global grid
  for (int d = 0; d < D; ++d) if (c[d] == 0) return;           // lower ghosts are never written (interface planes: the neighbour's)
  if (D == 3 && a.per.zhi && c[2] == N[2] - 1) return;
  const int st[3] = {1, l.pitch, l.plane};
  const int ci = D - 1;
  for (int dj = D - 1; dj >= 0; --dj) {
    const int lo = (a.per.p[dj] || (dj == 2 && a.per.zlo)) ? 1 : 2;
    if (c[dj] < lo || c[dj] > N[dj] - 2) continue;
    const int o = l.off(c[0], c[1], c[2]);
    int cj = c[dj], Nj = N[dj];
    zface_eff(a.per, dj, cj, Nj);
    a.sigma[o] = face_flux<S>(a.u + l.cstride * ci, a.u + l.cstride * dj, o, st[ci], st[dj], cj, Nj,
                              a.per.p[dj] != 0, a.nu);
    return;
  }
}

}  // namespace wl
