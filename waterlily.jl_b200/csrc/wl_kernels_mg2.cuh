// wl_kernels_mg2.cuh -- fused pressure-side kernels (v2).
//
//  * Periodic directions are handled by WRAP-READS: a kernel that the reference precedes with
//    perBC!(a) reads a's wrapped interior value instead of a ghost, so the ghost-fill launches
//    disappear.  Where the reference's ghosts are observable or deliberately stale (ϵ during the
//    red-black sweeps, the final pressure) the producing thread also stores the periodic images.
//  * K-DIVRES  = `@inside z=div(u)`, `x.*=dt`, perBC!(x), residual! loop 1, Σr   (Flow.jl:225, Poisson.jl:92-95)
//  * K-DOWN    = Jacobi! (ϵ=r·iD, perBC!, increment! with ω=1) + restrict! + fill!(coarse.x,0)
//                (Poisson.jl:100-114, MultiLevelPoisson.jl:92-94); one thread per COARSE cell
//  * K-UP      = prolongate! + perBC! + increment!(ω)                          (MultiLevelPoisson.jl:99-100)
//  * K-PROJ    = u -= L·∂x for every component, x./=dt, perBC!(x)              (Flow.jl:227-230, MLP.jl:126)
//  * k_coarse  = every level below a size threshold in ONE single-block launch (the V-cycle recursion
//                from that level down and back up, plus its smoother), because those levels are pure
//                launch latency.
// All arithmetic keeps the reference's operation order: results are bit-identical to the v1 kernels.
#pragma once
#include "wl_kernels_mg.cuh"

namespace wl {

struct LevDev {
  Lay l;
  const float* L;
  const float* iD;
  float* x;
  float* r;
  float* eps;
  int c[3];          // coarsening mask from the next finer level to this one
  TileFlags uf;      // uniform-coefficient tiles of this level (4-cells-per-thread kernels)
  int pf;            // 4-cells-per-thread kernels: planes ahead whose lines are prefetched into L2 (0 = off), see pf4()
};

// offsets of the lower/upper neighbour in direction d for an INTERIOR cell, wrapped when periodic.
// EDGE=false: the caller knows that no cell of its tile sits next to a periodic seam, so the
// offsets are the plain strides (no compares/selects: these kernels are issue-bound, not HBM-bound).
template <int D, bool EDGE = true>
__device__ __forceinline__ void nbr_off(const Lay& l, const Per& per, int i, int j, int k, int* om, int* op) {
  const int c[3] = {i, j, k};
  const int N[3] = {l.n0, l.n1, l.n2};
  const int st[3] = {1, l.pitch, l.plane};
#pragma unroll
  for (int d = 0; d < D; ++d) {
    om[d] = (EDGE && per.p[d] && c[d] == 1) ? (N[d] - 3) * st[d] : -st[d];
    op[d] = (EDGE && per.p[d] && c[d] == N[d] - 2) ? -(N[d] - 3) * st[d] : st[d];
  }
}
// does the tile of blockDim cells whose first interior cell is (i0,j0,k0) touch a periodic seam?
template <int D>
__device__ __forceinline__ bool tile_on_seam(const Lay& l, const Per& per, int i0, int j0, int k0) {
  bool e = false;
  if (per.p[0]) e |= i0 == 1 || i0 + (int)blockDim.x > l.n0 - 2;
  if (per.p[1]) e |= j0 == 1 || j0 + (int)blockDim.y > l.n1 - 2;
  if (D == 3 && per.p[2]) e |= k0 == 1 || k0 + (int)blockDim.z > l.n2 - 2;
  return e;
}
// for_tiles_inside with the seam decision made once per tile: f(i,j,k,EdgeYes/EdgeNo)
template <int D, class F>
__device__ __forceinline__ void for_tiles_seam(const Lay& l, const Per& per, F f);
struct EdgeYes { static constexpr bool value = true; };
struct EdgeNo { static constexpr bool value = false; };
template <int D, class F>
__device__ __forceinline__ void for_tiles_seam(const Lay& l, const Per& per, F f) {
  const int nbx = (l.n0 - 2 + blockDim.x - 1) / blockDim.x, nby = (l.n1 - 2 + blockDim.y - 1) / blockDim.y;
  const int nbz = D == 3 ? (l.n2 - 2 + blockDim.z - 1) / blockDim.z : 1;
  const int nt = nbx * nby * nbz;
  for (int t = blockIdx.x; t < nt; t += gridDim.x) {
    const int tx = t % nbx, q = t / nbx, ty = q % nby, tz = q / nby;
    const int i0 = tx * blockDim.x + 1, j0 = ty * blockDim.y + 1, k0 = D == 3 ? tz * blockDim.z + 1 : 0;
    const int i = i0 + threadIdx.x, j = j0 + threadIdx.y, k = D == 3 ? k0 + threadIdx.z : 0;
    const bool in = i <= l.n0 - 2 && j <= l.n1 - 2 && (D == 2 || k <= l.n2 - 2);
    if (tile_on_seam<D>(l, per, i0, j0, k0)) { if (in) f(i, j, k, EdgeYes{}); }
    else if (in) f(i, j, k, EdgeNo{});
  }
}
// store v into every periodic image (ghost copy) of interior cell (i,j,k): what the sequential
// perBC! passes (core.jl:239-243) leave in the ghosts, edges and corners included.
template <int D>
__device__ __forceinline__ void store_images(float* a, const Lay& l, const Per& per, int i, int j, int k, float v) {
  const int c[3] = {i, j, k};
  const int N[3] = {l.n0, l.n1, l.n2};
  int img[3];
  bool any = false;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    img[d] = -1;
    if (d < D && per.p[d]) {
      if (c[d] == 1) img[d] = N[d] - 1;
      else if (c[d] == N[d] - 2) img[d] = 0;
      if (N[d] == 3 && c[d] == 1) img[d] = -2;      // degenerate single-cell direction: both ghosts
    }
    any |= img[d] != -1;
  }
  if (!any) return;
  for (int m = 1; m < 8; ++m) {
    int q[3];
    bool ok = true;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      if (m & (1 << d)) { if (img[d] < 0) ok = false; q[d] = img[d]; } else q[d] = c[d];
    }
    if (ok) a[l.off(q[0], q[1], q[2])] = v;
  }
}
// `mult` (Poisson.jl:70-76) at an interior cell with wrapped neighbour offsets
template <int D>
__device__ __forceinline__ float multA_w(const Lay& l, const float* __restrict__ L, const float* __restrict__ x, int o,
                                         const int* om, const int* op) {
  const int st[3] = {1, l.pitch, l.plane};
  float s = x[o] * diagL<D>(l, L, o);
#pragma unroll
  for (int d = 0; d < D; ++d) {
    const float* Ld = L + l.cstride * d;
    s += x[o + om[d]] * Ld[o] + x[o + op[d]] * Ld[o + st[d]];
  }
  return s;
}

#define WL_IJK_FROM_BLOCK(D, l)                                                 \
  const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;                              \
  const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;                      \
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z + 1 : 0;         \
  const bool inb = !(i > l.n0 - 2 || j > l.n1 - 2 || (D == 3 && k > l.n2 - 2));

// ---------------------------------------------------------------- K-DIVRES
// x·dt is written to a second buffer `xs` (the level's idle ϵ array; the host then swaps the two
// pointers) because neighbours still read the unscaled x.  Non-periodic ghosts of every scalar work
// array are zero for the whole run, periodic ghosts are rewritten by whoever needs them, so the
// buffers are interchangeable.
struct DivResArgs { Lay l; const float* u; const float* x; float* xs; const float* L; float* r; Per per; float dt; RedWork w; Red* out; TileFlags uf; int pf; };
template <int D>
__global__ void __launch_bounds__(256) k_divres(const DivResArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  double s = 0.0;
  for_tiles_seam<D>(l, a.per, [&](int i, int j, int k, auto edge) {
    constexpr bool EDGE = decltype(edge)::value;
    const int o = l.off(i, j, k);
    const int st[3] = {1, l.pitch, l.plane};
    int om[3], op[3];
    nbr_off<D, EDGE>(l, a.per, i, j, k, om, op);
    float u0[3], u1[3], L0[3], L1[3], xm[3], xp[3];
    const float x0 = __ldg(a.x + o);
#pragma unroll
    for (int d = 0; d < D; ++d) {                    // every load before the first use
      const float* ud = a.u + l.cstride * d;
      const float* Ld = a.L + l.cstride * d;
      u0[d] = __ldg(ud + o); u1[d] = __ldg(ud + o + st[d]);
      L0[d] = __ldg(Ld + o); L1[d] = __ldg(Ld + o + st[d]);
      xm[d] = __ldg(a.x + o + om[d]); xp[d] = __ldg(a.x + o + op[d]);
    }
    float z = 0.f, dg = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) { z += u1[d] - u0[d]; dg -= L0[d] + L1[d]; }
    const float xc = x0 * a.dt;
    float r = 0.f;
    if (dg != 0.f) {
      float ax = xc * dg;
#pragma unroll
      for (int d = 0; d < D; ++d) ax += (xm[d] * a.dt) * L0[d] + (xp[d] * a.dt) * L1[d];
      r = z - ax;
    }
    a.r[o] = r;
    a.xs[o] = xc;
    s += (double)r;
  });
  block_reduce_finish(s, -INFINITY, a.w, a.out);
}

// ---------------------------------------------------------------- K-DOWN
// first != 0 (level 1 of the first V-cycle of a solve): the residual still lacks the mean removal of
// residual! (Poisson.jl:95-97).  The mean Σr/N arrives in `in`; the kernel subtracts it on the fly
// from every r it reads and reports L₁/L∞ of the shifted residual (MultiLevelPoisson.jl:111) in
// `out`, which saves the separate pass over r and one host round trip per solve.
struct DownArgs { LevDev f, c; Per per; int first; RedIn in; long long ncells; RedWork w; Red* out; };
// One thread per FINE cell (coalesced); the children of a coarse cell sit in one block (block
// extents are even and blocks start on odd interior indices), so the restriction gathers them from
// shared memory and adds them in the reference's order (first dimension fastest).
template <int D>
__global__ void __launch_bounds__(256) k_down(const DownArgs a) {
  pdl_entry();
  __shared__ float tile[256];
  const Lay& lc = a.c.l;
  const Lay& l = a.f.l;
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  double s1 = 0.0; float mx = -INFINITY;
  float sh = 0.f;                                                      // residual mean to remove (first cycle only)
  if (a.first) {
    const float mean = (float)redin_sum(a.in) / (float)a.ncells;
    if (fabsf(mean) > 2.f * 1.1920929e-07f) sh = mean;
  }
  const bool shift = sh != 0.f;
  // 1-D grid over the tiles of blockDim cells: one tile per block, or (first cycle, where the block
  // ends in a grid-wide reduction) a persistent loop of a few blocks per SM
  const int nbx = (l.n0 - 2 + blockDim.x - 1) / blockDim.x, nby = (l.n1 - 2 + blockDim.y - 1) / blockDim.y;
  const int nbz = D == 3 ? (l.n2 - 2 + blockDim.z - 1) / blockDim.z : 1;
  const int nt = nbx * nby * nbz;
  for (int tq = blockIdx.x; tq < nt; tq += gridDim.x) {
  const int tbx = tq % nbx, tqq = tq / nbx, tby = tqq % nby, tbz = tqq / nby;
  const int i0 = tbx * blockDim.x + 1, j0 = tby * blockDim.y + 1, k0 = D == 3 ? tbz * blockDim.z + 1 : 0;
  const int i = i0 + threadIdx.x, j = j0 + threadIdx.y, k = D == 3 ? k0 + threadIdx.z : 0;
  const bool inb = !(i > l.n0 - 2 || j > l.n1 - 2 || (D == 3 && k > l.n2 - 2));
  float rn = 0.f;
  auto body = [&](auto edge) {
    constexpr bool EDGE = decltype(edge)::value;
    const int o = l.off(i, j, k);
    const int st[3] = {1, l.pitch, l.plane};
    int om[3], op[3];
    nbr_off<D, EDGE>(l, a.per, i, j, k, om, op);
    float L0[3], L1[3], rm[3], rp[3], dm[3], dp[3];
    float r0 = a.f.r[o];
    const float x0 = a.f.x[o], d0 = __ldg(a.f.iD + o);                 // all reads before the first store
#pragma unroll
    for (int d = 0; d < D; ++d) {
      const float* Ld = a.f.L + l.cstride * d;
      L0[d] = __ldg(Ld + o); L1[d] = __ldg(Ld + o + st[d]);
      rm[d] = a.f.r[o + om[d]]; dm[d] = __ldg(a.f.iD + o + om[d]);
      rp[d] = a.f.r[o + op[d]]; dp[d] = __ldg(a.f.iD + o + op[d]);
    }
    if (shift) {
      // ghost neighbours of non-periodic directions are not shifted by the reference, but their iD is
      // 0 and so is the face coefficient: the product contributes a (signed) zero either way
      r0 = r0 - sh;
#pragma unroll
      for (int d = 0; d < D; ++d) { rm[d] = rm[d] - sh; rp[d] = rp[d] - sh; }
    }
    if (a.first) { s1 += (double)fabsf(r0); mx = fmaxf(mx, fabsf(r0)); }
    const float e0 = r0 * d0;                                          // ϵ = r·iD  Poisson.jl:112
    float dg = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) dg -= L0[d] + L1[d];
    float s = e0 * dg;                                                 // mult(I,L,D,ϵ)
#pragma unroll
    for (int d = 0; d < D; ++d) s += (rm[d] * dm[d]) * L0[d] + (rp[d] * dp[d]) * L1[d];
    rn = r0 - s;                                                       // increment!, ω = 1  Poisson.jl:102
    // r is still being read by the neighbours' ϵ: the new residual goes to the idle ϵ array and
    // the host swaps the two pointers afterwards
    a.f.eps[o] = rn;
    a.f.x[o] = x0 + e0;                                                // Poisson.jl:103
  };
  if (inb) {
    if (tile_on_seam<D>(l, a.per, i0, j0, k0)) body(EdgeYes{});
    else body(EdgeNo{});
  }
  tile[tid] = rn;
  __syncthreads();
  const int c[3] = {i, j, k};
  const int t[3] = {(int)threadIdx.x, (int)threadIdx.y, (int)threadIdx.z};
  int cc[3], n[3];
  bool parent = inb;                                                   // first child of its coarse cell
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const bool co = d < D && a.c.c[d];
    if (co && (t[d] & 1)) parent = false;
    cc[d] = co ? (c[d] + 1) / 2 : c[d];
    n[d] = co ? 2 : 1;
  }
  if (parent) {
    float sum = 0.f;                                                   // restrict  MLP.jl:13-19
    for (int kk = 0; kk < n[2]; ++kk)
      for (int jj = 0; jj < n[1]; ++jj)
        for (int ii = 0; ii < n[0]; ++ii) sum += tile[tid + ii + (int)blockDim.x * (jj + (int)blockDim.y * kk)];
    const int oc = lc.off(cc[0], cc[1], cc[2]);
    a.c.r[oc] = sum;
    a.c.x[oc] = 0.f;                                                   // fill!(coarse.x,0)  MLP.jl:94
  }
  __syncthreads();                                                     // the tile buffer is reused by the next trip
  }
  if (a.first) block_reduce_partial(s1, mx, a.w);                      // folded by k_fold_partials
}

// ---------------------------------------------------------------- K-UP
// write_eps: the smoother of this level always follows (MultiLevelPoisson.jl:97-100,115), so K-UP also
// leaves its first statement ϵ = r·iD (+ periodic images, Poisson.jl:142-143) behind
struct UpArgs { LevDev f, c; Per per; float omega; RedWork w; Red* out; int want_norms; int write_eps; };
template <int D>
__global__ void __launch_bounds__(256) k_up(const UpArgs a) {
  pdl_entry();
  const Lay& l = a.f.l;
  WL_IJK_FROM_BLOCK(D, l)
  double s1 = 0.0; float m = -INFINITY;
  auto body = [&](auto edge) {
    constexpr bool EDGE = decltype(edge)::value;
    const int o = l.off(i, j, k);
    const int st[3] = {1, l.pitch, l.plane};
    const int c[3] = {i, j, k};
    const int N[3] = {l.n0, l.n1, l.n2};
    const int cst[3] = {1, a.c.l.pitch, a.c.l.plane};
    // coarse cell of this cell and of its (periodically wrapped) neighbours: down(I,c) MLP.jl:7.
    // Away from a seam the lower neighbour of an odd index and the upper neighbour of an even index
    // share the cell's own parent.
    int oc = 0, dm[3], dp[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      if (d >= D) { dm[d] = dp[d] = 0; continue; }
      const bool co = a.c.c[d] != 0;
      const int cc = co ? (c[d] + 1) >> 1 : c[d];
      oc += cc * cst[d];
      if (EDGE) {
        const int xm = (a.per.p[d] && c[d] == 1) ? N[d] - 2 : c[d] - 1;
        const int xp = (a.per.p[d] && c[d] == N[d] - 2) ? 1 : c[d] + 1;
        dm[d] = ((co ? (xm + 1) >> 1 : xm) - cc) * cst[d];
        dp[d] = ((co ? (xp + 1) >> 1 : xp) - cc) * cst[d];
      } else {
        dm[d] = co ? -(c[d] & 1) * cst[d] : -cst[d];
        dp[d] = co ? ((c[d] & 1) ^ 1) * cst[d] : cst[d];
      }
    }
    float L0[3], L1[3], em[3], ep[3];
    const float e0 = __ldg(a.c.x + oc);
    const float r0 = a.f.r[o], x0 = a.f.x[o];            // all reads before the first store
    const float d0 = a.write_eps ? __ldg(a.f.iD + o) : 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) {
      const float* Ld = a.f.L + l.cstride * d;
      L0[d] = __ldg(Ld + o); L1[d] = __ldg(Ld + o + st[d]);
      em[d] = __ldg(a.c.x + oc + dm[d]); ep[d] = __ldg(a.c.x + oc + dp[d]);
    }
    float dg = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) dg -= L0[d] + L1[d];
    float s = e0 * dg;
#pragma unroll
    for (int d = 0; d < D; ++d) s += em[d] * L0[d] + ep[d] * L1[d];
    const float r = r0 - a.omega * s;
    a.f.r[o] = r;
    a.f.x[o] = x0 + a.omega * e0;
    if (a.write_eps) {
      const float e = r * d0;
      a.f.eps[o] = e;
      if (EDGE) store_images<D>(a.f.eps, l, a.per, i, j, k, e);
    }
    s1 = (double)fabsf(r); m = fabsf(r);
  };
  if (inb) {
    if (tile_on_seam<D>(l, a.per, (int)(blockIdx.x * blockDim.x) + 1, (int)(blockIdx.y * blockDim.y) + 1,
                        D == 3 ? (int)(blockIdx.z * blockDim.z) + 1 : 0)) body(EdgeYes{});
    else body(EdgeNo{});
  }
  if (a.want_norms) block_reduce_finish(s1, m, a.w, a.out);
}

// ---------------------------------------------------------------- smoother pieces
// ϵ = r·iD inside + perBC!(ϵ) images  (GaussSeidelRB! Poisson.jl:142-143).  The sweeps then read
// these ghosts as they are (never refreshed between sweeps: the reference's stale-ghost behaviour).
struct SmArgs {
  LevDev p; Per per; float omega; int k0; RedWork w; Red* out; int want_norms;
  // k_incr4 after fused half sweeps (k_gs_half_cx): arrival flags the boundary slabs wait for
  unsigned int* wflag; unsigned int wait_seq_lo, wait_seq_hi; int wait_lo, wait_hi;
};
template <int D>
__global__ void __launch_bounds__(256) k_eps_init_g(const SmArgs a) {
  pdl_entry();
  const Lay& l = a.p.l;
  WL_IJK_FROM_BLOCK(D, l)
  if (!inb) return;
  const int o = l.off(i, j, k);
  const float e = a.p.r[o] * __ldg(a.p.iD + o);
  a.p.eps[o] = e;
  if (tile_on_seam<D>(l, a.per, (int)(blockIdx.x * blockDim.x) + 1, (int)(blockIdx.y * blockDim.y) + 1,
                      D == 3 ? (int)(blockIdx.z * blockDim.z) + 1 : 0))
    store_images<D>(a.p.eps, l, a.per, i, j, k, e);
}
// increment! (Poisson.jl:100-104) with the perBC!(ϵ) folded into wrap-reads; optional L₁/L∞ of the new r
template <int D>
__global__ void __launch_bounds__(256) k_increment_w(const SmArgs a) {
  pdl_entry();
  const Lay& l = a.p.l;
  double s1 = 0.0; float m = -INFINITY;
  // 1-D persistent grid (see for_tiles_inside); every read is issued before the first store so the
  // thread pays ONE memory round trip (r, x, ϵ may alias as far as the compiler knows)
  for_tiles_seam<D>(l, a.per, [&](int i, int j, int k, auto edge) {
    constexpr bool EDGE = decltype(edge)::value;
    const int o = l.off(i, j, k);
    const int st[3] = {1, l.pitch, l.plane};
    int om[3], op[3];
    nbr_off<D, EDGE>(l, a.per, i, j, k, om, op);
    float L0[3], L1[3], em[3], ep[3];
    const float r0 = a.p.r[o], x0 = a.p.x[o], e0 = a.p.eps[o];
#pragma unroll
    for (int d = 0; d < D; ++d) {
      const float* Ld = a.p.L + l.cstride * d;
      L0[d] = __ldg(Ld + o); L1[d] = __ldg(Ld + o + st[d]);
      em[d] = a.p.eps[o + om[d]]; ep[d] = a.p.eps[o + op[d]];
    }
    float dg = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) dg -= L0[d] + L1[d];
    float s = e0 * dg;                                       // mult(I,L,D,ϵ)  Poisson.jl:70-76
#pragma unroll
    for (int d = 0; d < D; ++d) s += em[d] * L0[d] + ep[d] * L1[d];
    const float r = r0 - a.omega * s;
    a.p.r[o] = r;
    a.p.x[o] = x0 + a.omega * e0;
    s1 += (double)fabsf(r); m = fmaxf(m, fabsf(r));
  });
  if (a.want_norms) block_reduce_finish(s1, m, a.w, a.out);
}

// ---------------------------------------------------------------- K-PROJ
// u[I,i] -= L[I,i]·(x[I]-x[I-δᵢ]) for all i (after the solver's final perBC!(x): wrap-read), and the
// unscaled pressure p = x/dt with its periodic images, written to the spare buffer (host swaps).
struct ProjArgs2 { Lay l; float* u; const float* x; float* pn; const float* L; Per per; float dt; TileFlags uf; };
template <int D>
__global__ void __launch_bounds__(256) k_proj(const ProjArgs2 a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_IJK_FROM_BLOCK(D, l)
  if (!inb) return;
  const int o = l.off(i, j, k);
  const bool edge = tile_on_seam<D>(l, a.per, (int)(blockIdx.x * blockDim.x) + 1, (int)(blockIdx.y * blockDim.y) + 1,
                                    D == 3 ? (int)(blockIdx.z * blockDim.z) + 1 : 0);
  int om[3], op[3];
  if (edge) nbr_off<D, true>(l, a.per, i, j, k, om, op);
  else nbr_off<D, false>(l, a.per, i, j, k, om, op);
  const float xc = __ldg(a.x + o);
  float uv[3], Lv[3], xm[3];
#pragma unroll
  for (int d = 0; d < D; ++d) {                              // all reads first: one memory round trip
    const long long oc = l.cstride * d + o;
    uv[d] = a.u[oc]; Lv[d] = __ldg(a.L + oc); xm[d] = __ldg(a.x + o + om[d]);
  }
#pragma unroll
  for (int d = 0; d < D; ++d) a.u[l.cstride * d + o] = uv[d] - Lv[d] * (xc - xm[d]);
  const float p = xc / a.dt;
  a.pn[o] = p;
  if (edge) store_images<D>(a.pn, l, a.per, i, j, k, p);
  if (D == 3) {                                              // slab interfaces: see k_proj4
    if (a.per.zlo && k == 1) a.pn[o - l.plane] = xm[2] / a.dt;
    if (a.per.zhi && k == l.n2 - 2) a.pn[o + l.plane] = __ldg(a.x + o + l.plane) / a.dt;
  }
}

// ---------------------------------------------------------------- all small levels in one block
constexpr int WL_MAXLEV = 12;
struct CoarseArgs {
  LevDev lv[WL_MAXLEV];
  int first, nlev;     // levels first..nlev-1 are processed here
  Per per;
  float omega;
};
template <int D, class F>
__device__ __forceinline__ void for_inside(const Lay& l, F f) {
  const int m0 = l.n0 - 2, m1 = l.n1 - 2, m2 = D == 3 ? l.n2 - 2 : 1;
  const int tot = m0 * m1 * m2;
  // q < 2^21: (q+0.5)*(1/m) truncates to q/m exactly, and costs 3 instructions instead of ~40
  const float im0 = 1.0f / (float)m0, im1 = 1.0f / (float)m1;
  for (int q = threadIdx.x; q < tot; q += blockDim.x) {
    const int t = (int)(((float)q + 0.5f) * im0), i = 1 + q - t * m0;
    const int tk = D == 3 ? (int)(((float)t + 0.5f) * im1) : 0, j = 1 + t - tk * m1, k = D == 3 ? 1 + tk : 0;
    f(i, j, k, l.off(i, j, k));
  }
}
template <int D>
__device__ __forceinline__ void blk_smooth(const LevDev& p, const Per& per, float w) {
  const Lay& l = p.l;
  for_inside<D>(l, [&](int i, int j, int k, int o) {
    const float e = p.r[o] * p.iD[o];
    p.eps[o] = e;
    store_images<D>(p.eps, l, per, i, j, k, e);
  });
  __syncthreads();
  const int st[3] = {1, l.pitch, l.plane};
  const int nl = D == 3 ? l.n2 : l.n1;
  for (int k0 = 1; k0 <= 4; ++k0) {
    for_inside<D>(l, [&](int i, int j, int k, int o) {
      const int cl = D == 3 ? k : j;
      if (cl > 2 * (nl / 2) - 2) return;
      const int jsum = i + j + (D == 3 ? k : 0) + D;
      if (((jsum ^ (k0 + 1)) & 1) != 0) return;
      float s = p.r[o];
#pragma unroll
      for (int q = 0; q < D; ++q) {
        const float* Lq = p.L + l.cstride * q;
        s -= p.eps[o - st[q]] * Lq[o] + p.eps[o + st[q]] * Lq[o + st[q]];
      }
      p.eps[o] = s * p.iD[o];
    });
    __syncthreads();
  }
  for_inside<D>(l, [&](int i, int j, int k, int o) {
    int om[3], op[3];
    nbr_off<D>(l, per, i, j, k, om, op);
    p.r[o] = p.r[o] - w * multA_w<D>(l, p.L, p.eps, o, om, op);
    p.x[o] = p.x[o] + w * p.eps[o];
  });
  __syncthreads();
}
// Vcycle!(ml;l=first,ω) followed by smooth!(levels[first];ω)   (MultiLevelPoisson.jl:88-101,106)
template <int D>
__device__ __forceinline__ void coarse_body(const LevDev* lv, int first, int nlev, const Per& per, float omega) {
  const int last = nlev - 1;
  for (int lev = first; lev < last; ++lev) {              // way down: Jacobi!, restrict!, fill!(x,0)
    const LevDev f = lv[lev];
    const LevDev c = lv[lev + 1];
    const Lay& l = f.l;
    for_inside<D>(l, [&](int i, int j, int k, int o) { f.eps[o] = f.r[o] * f.iD[o]; });
    __syncthreads();
    for_inside<D>(l, [&](int i, int j, int k, int o) {
      int om[3], op[3];
      nbr_off<D>(l, per, i, j, k, om, op);
      f.r[o] = f.r[o] - multA_w<D>(l, f.L, f.eps, o, om, op);
      f.x[o] = f.x[o] + f.eps[o];
    });
    __syncthreads();
    for_inside<D>(c.l, [&](int ci, int cj, int ck, int oc) {
      const int cc[3] = {ci, cj, ck};
      int lo[3], hi[3];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        if (d >= D) { lo[d] = hi[d] = 0; }
        else if (!c.c[d]) { lo[d] = hi[d] = cc[d]; }
        else { lo[d] = 2 * cc[d] - 1; hi[d] = 2 * cc[d]; }
      }
      float s = 0.f;
      for (int kk = lo[2]; kk <= hi[2]; ++kk)
        for (int jj = lo[1]; jj <= hi[1]; ++jj)
          for (int ii = lo[0]; ii <= hi[0]; ++ii) s += f.r[l.off(ii, jj, kk)];
      c.r[oc] = s;
      c.x[oc] = 0.f;
    });
    __syncthreads();
  }
  for (int lev = last - 1; lev >= first; --lev) {         // way up: smooth!(coarse), prolongate!, increment!
    const LevDev f = lv[lev];
    const LevDev c = lv[lev + 1];
    blk_smooth<D>(c, per, omega);
    const Lay& l = f.l;
    for_inside<D>(l, [&](int i, int j, int k, int o) {
      const int ci = c.c[0] ? (i + 1) / 2 : i, cj = c.c[1] ? (j + 1) / 2 : j, ck = (D == 3 && c.c[2]) ? (k + 1) / 2 : k;
      f.eps[o] = c.x[c.l.off(ci, cj, ck)];
    });
    __syncthreads();
    for_inside<D>(l, [&](int i, int j, int k, int o) {
      int om[3], op[3];
      nbr_off<D>(l, per, i, j, k, om, op);
      f.r[o] = f.r[o] - omega * multA_w<D>(l, f.L, f.eps, o, om, op);
      f.x[o] = f.x[o] + omega * f.eps[o];
    });
    __syncthreads();
  }
  { const LevDev f = lv[first]; blk_smooth<D>(f, per, omega); }
}
template <int D>
__global__ void __launch_bounds__(1024) k_coarse(const CoarseArgs a) {
  pdl_entry();
  coarse_body<D>(a.lv, a.first, a.nlev, a.per, a.omega);
}

}  // namespace wl
