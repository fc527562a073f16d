// wl_kernels_mg4.cuh -- x-vectorised pressure kernels: one thread owns FOUR consecutive cells of a row.
//
// ncu on the one-cell-per-thread kernels (profiles/r1_v9_ncu_full_sphere512.tsv) shows them issue-bound,
// not HBM-bound: 120-400 instructions per cell, most of them 64-bit address arithmetic around scalar
// loads (K-UP: 80 % issue-active at 52 % DRAM).  With four cells per thread the rows above/below and
// the planes in front/behind arrive as one 128-bit load each, the index arithmetic is paid once per
// four cells, and the kernels move towards the HBM roofline.  The arithmetic per cell and its order
// are those of the scalar kernels (wl_kernels_mg2.cuh), so results stay bit-identical.
//
// Layout facts used: interior cell i==1 of every row is 128-byte aligned, pitch/plane/cstride are
// multiples of 32 floats, so a thread whose first cell is i = 1+4m loads aligned float4s.  The
// kernels require (n0-2) % 4 == 0; other levels use the scalar kernels.
#pragma once
#include <type_traits>

#include "wl_kernels_mg2.cuh"

// Resident blocks per SM the register allocation is held to (second __launch_bounds__ argument).  These kernels live
// for one memory round trip, so what they gain from more warps in flight outweighs a few spilled bytes -- measured per
// kernel on the sphere grid (profiles/r2_c34_*, r2_c35_*, r2_c36_*; level-1 launch, us):
//   k_proj4   3 blocks (70 regs) 206 | 4 (64) 180 | 5 (48, 8 B spilled) 170      -> 5
//   k_up4     4 (64) 168 | 5 (48, 32 B spilled) 161                              -> 5
//   k_down4<1> (lean instantiation)  4 (56) 158 | 5 (43) 138 | 6 (40) 127        -> 6
//   k_divres4<1>                     4 (56) 158 | 5 (48, 32 B spilled) 174       -> 4
//   k_incr4   4 (64) 148 | 5 (48, 72 B spilled) 150                              -> 4
// The general instantiations (k_down4<0/2>, k_divres4<0/2>) need 77-80 registers and spill when held to 64: 3 blocks.
#ifndef WL_MG4_MINB_A
#define WL_MG4_MINB_A 4
#endif
#ifndef WL_MG4_MINB_B
#define WL_MG4_MINB_B 3
#endif

namespace wl {

struct F4 { float v[4]; };
__device__ __forceinline__ F4 ld4g(const float* p) {           // read-only data
  const float4 t = __ldg(reinterpret_cast<const float4*>(p));
  return F4{{t.x, t.y, t.z, t.w}};
}
__device__ __forceinline__ F4 ld4(const float* p) {            // data other threads of this kernel may write elsewhere
  const float4 t = *reinterpret_cast<const float4*>(p);
  return F4{{t.x, t.y, t.z, t.w}};
}
__device__ __forceinline__ void st4(float* p, const F4& a) {
  *reinterpret_cast<float4*>(p) = make_float4(a.v[0], a.v[1], a.v[2], a.v[3]);
}
// ---- face coefficients of a quad by tile class (TileFlags, wl_common.cuh): 0 loaded, 1 the level's constants, 2 the
// constants with the boundary faces zeroed by index compares.  ly*/lz* are the same for the quad's four cells in
// classes 1 and 2, so the compiler keeps scalars there.
// (class 2: a row that touches no y/z boundary -- warp-uniform -- has the plain constants everywhere except the quad's
//  first lower / last upper x face at the ends of the row: `xq` = 1 + (first face zero) + 2·(last face zero), 0 = general)
struct QuadL { F4 lx, ly0, ly1, lz0, lz1; float lx4; int xq; };
template <int CLS>
__device__ __forceinline__ QuadL quad_coeffs(const float* L, const Lay& l, const Per& per, const TileFlags& t, int o, int i, int j, int k) {
  QuadL q;
  q.xq = 0;
  if (CLS == 0) {
    const float* Lx = L; const float* Ly = L + l.cstride; const float* Lz = L + 2 * l.cstride;
    q.lx = ld4g(Lx + o); q.lx4 = __ldg(Lx + o + 4);
    q.ly0 = ld4g(Ly + o); q.ly1 = ld4g(Ly + o + l.pitch); q.lz0 = ld4g(Lz + o); q.lz1 = ld4g(Lz + o + l.plane);
  } else if (CLS == 3) {
    // a tile whose only boundary is an x end (the split kernels' first launch): two selects, constants otherwise
    const bool z0 = pattern_zero(per, 0, i, l.n0), z4 = pattern_zero(per, 0, i + 4, l.n0);
    q.xq = 1 + (z0 ? 1 : 0) + (z4 ? 2 : 0);
    q.lx = F4{{z0 ? 0.f : t.lu[0], t.lu[0], t.lu[0], t.lu[0]}};
    q.lx4 = z4 ? 0.f : t.lu[0];
    q.ly0 = q.ly1 = F4{{t.lu[1], t.lu[1], t.lu[1], t.lu[1]}};
    q.lz0 = q.lz1 = F4{{t.lu[2], t.lu[2], t.lu[2], t.lu[2]}};
  } else if (CLS == 2 && !(pattern_zero(per, 1, j, l.n1) || pattern_zero(per, 1, j + 1, l.n1) ||
                           pattern_zero(per, 2, k, l.n2) || pattern_zero(per, 2, k + 1, l.n2))) {
    const bool z0 = pattern_zero(per, 0, i, l.n0), z4 = pattern_zero(per, 0, i + 4, l.n0);
    q.xq = 1 + (z0 ? 1 : 0) + (z4 ? 2 : 0);
    q.lx = F4{{z0 ? 0.f : t.lu[0], t.lu[0], t.lu[0], t.lu[0]}};
    q.lx4 = z4 ? 0.f : t.lu[0];
    q.ly0 = q.ly1 = F4{{t.lu[1], t.lu[1], t.lu[1], t.lu[1]}};
    q.lz0 = q.lz1 = F4{{t.lu[2], t.lu[2], t.lu[2], t.lu[2]}};
  } else {
    auto f = [&](int d, int c, int n) { return (CLS >= 2 && pattern_zero(per, d, c, n)) ? 0.f : t.lu[d]; };
#pragma unroll
    for (int c = 0; c < 4; ++c) q.lx.v[c] = f(0, i + c, l.n0);
    q.lx4 = f(0, i + 4, l.n0);
    const float a0 = f(1, j, l.n1), a1 = f(1, j + 1, l.n1), b0 = f(2, k, l.n2), b1 = f(2, k + 1, l.n2);
    q.ly0 = F4{{a0, a0, a0, a0}}; q.ly1 = F4{{a1, a1, a1, a1}}; q.lz0 = F4{{b0, b0, b0, b0}}; q.lz1 = F4{{b1, b1, b1, b1}};
  }
  return q;
}
// iD of the quad's own cells: loaded, the level's constant, or inv_or_zero(diag) of the synthesised faces (set_diag!'s
// own expression, Poisson.jl:43-55: the same operands in the same order, so the same bits as the stored value)
template <int CLS>
__device__ __forceinline__ F4 quad_iD(const float* iD, const QuadL& q, const TileFlags& t, int o) {
  if (CLS == 0) return ld4g(iD + o);
  if (CLS == 1) return F4{{t.idu, t.idu, t.idu, t.idu}};
  if (CLS == 3 || q.xq) return F4{{((q.xq - 1) & 1) ? t.idx1 : t.idu, t.idu, t.idu, ((q.xq - 1) & 2) ? t.idx1 : t.idu}};   // xq-1 = z0 + 2·z4
  F4 d;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float dg = 0.f;
    dg -= q.lx.v[c] + (c < 3 ? q.lx.v[c < 3 ? c + 1 : 3] : q.lx4); dg -= q.ly0.v[c] + q.ly1.v[c]; dg -= q.lz0.v[c] + q.lz1.v[c];
    d.v[c] = inv_or_zero(dg);
  }
  return d;
}
#define WL_CLASS_DISPATCH(cls, body)                                              \
  do {                                                                            \
    if ((cls) == 1) body(std::integral_constant<int, 1>{});                       \
    else if ((cls) >= 2) body(std::integral_constant<int, 2>{});                  \
    else body(std::integral_constant<int, 0>{});                                  \
  } while (0)
inline bool vec4_ok(const Lay& l) { return l.D == 3 && (l.n0 - 2) % 4 == 0 && l.n0 - 2 >= 8; }

// These kernels live for one memory round trip and blocks are issued in z order: a thread asks for the 16 bytes its
// counterpart `pf` planes further on (a block ~1.5 waves later) will load, so that those loads hit L2 (a warp's requests
// fall on four lines).  Measured on the half sweep: 97 -> 91 us (profiles/r2_c38_*).
__device__ __forceinline__ void pf4(const float* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// thread -> first cell of its quad; blockDim = (32,4,2) covers 128 x 4 x 2 cells
#define WL_QUAD_FROM_BLOCK(l)                                                   \
  const int i = 1 + 4 * (blockIdx.x * blockDim.x + threadIdx.x);                \
  const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;                      \
  const int k = blockIdx.z * blockDim.z + threadIdx.z + 1;                      \
  const bool inb = !(i > l.n0 - 2 || j > l.n1 - 2 || k > l.n2 - 2);

// The split kernels (K-DIVRES, K-DOWN).  MODE 0: one launch over the grid of tiles, every class.  MODE 1: the same grid,
// compiled for classes 1 and 3 only (constants + x-end selects: no coefficient or iD arrays, 56 instead of 77-80
// registers, one more resident block per SM); blocks of class 0 / 2 tiles leave at once.  MODE 2: a 1-D launch over
// TileFlags::list, the class 0 / 2 tiles, with the general code.  (Serving the y/z boundary tiles from the lean
// instantiation too -- iD from a 27-entry table -- was measured: it slows the common rows more than the shorter list
// saves, profiles/r2_c28_*.)  A tile's reduction partial is stored at the tile's linear index in every mode, so the
// folded sums do not depend on the split.
template <int MODE>
__device__ __forceinline__ bool split_tile(const TileFlags& uf, int& bx, int& by, int& bz, int& cls) {
  if (MODE == 2) {
    const int t = uf.list[blockIdx.x];
    bx = t % uf.nx; by = (t / uf.nx) % uf.ny; bz = t / (uf.nx * uf.ny);
  } else {
    bx = blockIdx.x; by = blockIdx.y; bz = blockIdx.z;
  }
  cls = tile_class(uf, bx, by, bz);
  return !(MODE == 1 && (cls == 0 || cls == 2));
}
#define WL_QUAD_FROM_TILE(l, bx, by, bz)                                        \
  const int i = 1 + 4 * ((bx) * blockDim.x + threadIdx.x);                      \
  const int j = (by) * blockDim.y + threadIdx.y + 1;                            \
  const int k = (bz) * blockDim.z + threadIdx.z + 1;                            \
  const bool inb = !(i > l.n0 - 2 || j > l.n1 - 2 || k > l.n2 - 2);
#define WL_SPLIT_DISPATCH(MODE, cls, body)                                        \
  do {                                                                            \
    if (MODE == 1) body(std::integral_constant<int, 3>{});                        \
    else if (MODE == 2) { if ((cls) == 2) body(std::integral_constant<int, 2>{}); else body(std::integral_constant<int, 0>{}); } \
    else WL_CLASS_DISPATCH(cls, body);                                            \
  } while (0)

// y/z neighbour offsets of the quad's row (wrapped across periodic seams) and the x offsets of the
// cell left of the quad / right of it
__device__ __forceinline__ void quad_nbr(const Lay& l, const Per& per, int i, int j, int k, int& oxm, int& oxp,
                                         int& oym, int& oyp, int& ozm, int& ozp) {
  oxm = (per.p[0] && i == 1) ? l.n0 - 3 : -1;                            // relative to the quad's first cell
  oxp = (per.p[0] && i + 3 == l.n0 - 2) ? 3 - (l.n0 - 3) : 4;
  oym = (per.p[1] && j == 1) ? (l.n1 - 3) * l.pitch : -l.pitch;
  oyp = (per.p[1] && j == l.n1 - 2) ? -(l.n1 - 3) * l.pitch : l.pitch;
  ozm = (per.p[2] && k == 1) ? (l.n2 - 3) * l.plane : -l.plane;
  ozp = (per.p[2] && k == l.n2 - 2) ? -(l.n2 - 3) * l.plane : l.plane;
}

// ---------------------------------------------------------------- increment! (Poisson.jl:100-104), 4 cells per thread
// perBC!(ϵ) folded into wrap-reads as in k_increment_w; L₁/L∞ partials per block (k_fold_partials folds them)
__global__ void __launch_bounds__(256, WL_MG4_MINB_A) k_incr4(const SmArgs a) {
  pdl_entry();
  const Lay& l = a.p.l;
  if (a.wait_lo | a.wait_hi) {                         // ϵ ghost planes pushed by the neighbours' last half sweep
    const bool wlo = a.wait_lo && blockIdx.z == 0, whi = a.wait_hi && blockIdx.z == gridDim.z - 1;
    if (wlo || whi) {
      if (threadIdx.x + threadIdx.y + threadIdx.z == 0) {
        if (wlo) wait_flag(a.wflag + 10, a.wait_seq_lo, a.wflag + 5);
        if (whi) wait_flag(a.wflag + 11, a.wait_seq_hi, a.wflag + 5);
      }
      __syncthreads();
    }
  }
  WL_QUAD_FROM_BLOCK(l)
  double s1 = 0.0; float m = -INFINITY;
  // the tile class is fetched first and the loads every class needs are issued BEFORE the branch on it: a thread that
  // waits for the class byte and only then starts its loads pays two memory latencies in a row (ncu source page of
  // K-DOWN, profiles/r2_c15_*: 10 % of the stall samples sat on the compare of the class)
  const int cls = tile_class(a.p.uf, blockIdx.x, blockIdx.y, blockIdx.z);
  if (inb) {
  const int o = l.off(i, j, k);
  int oxm, oxp, oym, oyp, ozm, ozp;
  quad_nbr(l, a.per, i, j, k, oxm, oxp, oym, oyp, ozm, ozp);
  const F4 e = ld4(a.p.eps + o), eym = ld4(a.p.eps + o + oym), eyp = ld4(a.p.eps + o + oyp),
           ezm = ld4(a.p.eps + o + ozm), ezp = ld4(a.p.eps + o + ozp);
  const float exm = a.p.eps[o + oxm], exp_ = a.p.eps[o + oxp];
  const F4 r0 = ld4(a.p.r + o), x0 = ld4(a.p.x + o);
  if (a.p.pf > 0 && k + a.p.pf <= l.n2 - 2) { const int q = o + a.p.pf * l.plane; pf4(a.p.eps + q); pf4(a.p.r + q); pf4(a.p.x + q); }
  auto body = [&](auto tag) {
    constexpr int CLS = decltype(tag)::value;
    const QuadL ql = quad_coeffs<CLS>(a.p.L, l, a.per, a.p.uf, o, i, j, k);
    const F4 &lx = ql.lx, &ly0 = ql.ly0, &ly1 = ql.ly1, &lz0 = ql.lz0, &lz1 = ql.lz1; const float lx4 = ql.lx4;
    F4 rn, xn;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float lxl = lx.v[q], lxu = q < 3 ? lx.v[q < 3 ? q + 1 : 3] : lx4;
      const float el = q == 0 ? exm : e.v[q > 0 ? q - 1 : 0], er = q == 3 ? exp_ : e.v[q < 3 ? q + 1 : 3];
      float dg = 0.f;                                              // diag(I,L)  Poisson.jl:49-55
      dg -= lxl + lxu; dg -= ly0.v[q] + ly1.v[q]; dg -= lz0.v[q] + lz1.v[q];
      float s = e.v[q] * dg;                                       // mult(I,L,D,ϵ)  Poisson.jl:70-76
      s += el * lxl + er * lxu;
      s += eym.v[q] * ly0.v[q] + eyp.v[q] * ly1.v[q];
      s += ezm.v[q] * lz0.v[q] + ezp.v[q] * lz1.v[q];
      const float r = r0.v[q] - a.omega * s;
      rn.v[q] = r;
      xn.v[q] = x0.v[q] + a.omega * e.v[q];
      s1 += (double)fabsf(r); m = fmaxf(m, fabsf(r));
    }
    st4(a.p.r + o, rn);
    st4(a.p.x + o, xn);
  };
  WL_CLASS_DISPATCH(cls, body);
  }
  if (a.want_norms) block_reduce_partial(s1, m, a.w);           // folded by k_fold_partials
}

// ---------------------------------------------------------------- CFL (Flow.jl:234-244), 4 cells per thread
// Same cells, same per-cell expression and the same "whole-array maximum" as k_cfl: rows and planes
// 0..N-1 are all visited, a thread on a ghost row takes the maximum of the σ already stored there (zeros
// below, the corrector's stale Φ above -- k_ghost_phi), a thread on an interior row computes its quad and
// the first / last thread of the row also folds in the row's two x ghosts.  The maximum does not depend
// on the order, so Δt is bit-identical to the scalar kernel's.
__global__ void __launch_bounds__(256) k_cfl4(const CflArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = 1 + 4 * (blockIdx.x * blockDim.x + threadIdx.x);
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = blockIdx.z * blockDim.z + threadIdx.z;
  float m = -INFINITY;
  // slab interfaces: the planes beyond them belong to the neighbour
  const bool mine = !((a.zlo && k == 0) || (a.zhi && k == l.n2 - 1));
  if (i <= l.n0 - 2 && j < l.n1 && k < l.n2 && mine) {
    const int o = l.off(i, j, k);
    const bool in = j >= 1 && j <= l.n1 - 2 && k >= 1 && k <= l.n2 - 2;
    const float g0 = i == 1 ? a.sigma[o - 1] : -INFINITY, g1 = i + 3 == l.n0 - 2 ? a.sigma[o + 4] : -INFINITY;
    m = fmaxf(g0, g1);
    if (in) {
      const float* ux = a.u; const float* uy = a.u + l.cstride; const float* uz = a.u + 2 * l.cstride;
      const F4 u0 = ld4g(ux + o); const float u04 = __ldg(ux + o + 4);
      const F4 v0 = ld4g(uy + o), v1 = ld4g(uy + o + l.pitch), w0 = ld4g(uz + o), w1 = ld4g(uz + o + l.plane);
      F4 sg;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float uxu = q < 3 ? u0.v[q < 3 ? q + 1 : 3] : u04;
        float s = 0.f;                                             // flux_out, Flow.jl:238-244
        s += fmaxf(0.f, uxu) + fmaxf(0.f, -u0.v[q]);
        s += fmaxf(0.f, v1.v[q]) + fmaxf(0.f, -v0.v[q]);
        s += fmaxf(0.f, w1.v[q]) + fmaxf(0.f, -w0.v[q]);
        sg.v[q] = s;
        m = fmaxf(m, s);
      }
      st4(a.sigma + o, sg);
    } else {
      const F4 sg = ld4(a.sigma + o);
#pragma unroll
      for (int q = 0; q < 4; ++q) m = fmaxf(m, sg.v[q]);
    }
  }
  block_reduce_partial(0.0, m, a.w);
}

// ---------------------------------------------------------------- K-DIVRES, 4 cells per thread (see k_divres)
template <int MODE>
__global__ void __launch_bounds__(256, MODE == 1 ? WL_MG4_MINB_A : WL_MG4_MINB_B) k_divres4(const DivResArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  int bx, by, bz, cls;                                                    // class first: see k_incr4
  if (!split_tile<MODE>(a.uf, bx, by, bz, cls)) return;
  WL_QUAD_FROM_TILE(l, bx, by, bz)
  double s = 0.0;
  if (inb) {
  const int o = l.off(i, j, k);
  int oxm, oxp, oym, oyp, ozm, ozp;
  quad_nbr(l, a.per, i, j, k, oxm, oxp, oym, oyp, ozm, ozp);
  const float* ux = a.u; const float* uy = a.u + l.cstride; const float* uz = a.u + 2 * l.cstride;
  const F4 u0 = ld4g(ux + o); const float u04 = __ldg(ux + o + 4);
  const F4 v0 = ld4g(uy + o), v1 = ld4g(uy + o + l.pitch), w0 = ld4g(uz + o), w1 = ld4g(uz + o + l.plane);
  const F4 x = ld4g(a.x + o), xym = ld4g(a.x + o + oym), xyp = ld4g(a.x + o + oyp),
           xzm = ld4g(a.x + o + ozm), xzp = ld4g(a.x + o + ozp);
  const float xxm = __ldg(a.x + o + oxm), xxp = __ldg(a.x + o + oxp);
  if (a.pf > 0 && k + a.pf <= l.n2 - 2) { const int q = o + a.pf * l.plane; pf4(ux + q); pf4(uy + q); pf4(uz + q); pf4(a.x + q); }
  auto body = [&](auto tag) {
    constexpr int CLS = decltype(tag)::value;
    const QuadL ql = quad_coeffs<CLS>(a.L, l, a.per, a.uf, o, i, j, k);
    const F4 &lx = ql.lx, &ly0 = ql.ly0, &ly1 = ql.ly1, &lz0 = ql.lz0, &lz1 = ql.lz1; const float lx4 = ql.lx4;
    F4 rr, xs;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float uxu = q < 3 ? u0.v[q < 3 ? q + 1 : 3] : u04;
      const float lxl = lx.v[q], lxu = q < 3 ? lx.v[q < 3 ? q + 1 : 3] : lx4;
      const float xl = q == 0 ? xxm : x.v[q > 0 ? q - 1 : 0], xr = q == 3 ? xxp : x.v[q < 3 ? q + 1 : 3];
      float z = 0.f, dg = 0.f;
      z += uxu - u0.v[q]; z += v1.v[q] - v0.v[q]; z += w1.v[q] - w0.v[q];
      dg -= lxl + lxu; dg -= ly0.v[q] + ly1.v[q]; dg -= lz0.v[q] + lz1.v[q];
      const float xc = x.v[q] * a.dt;
      float r = 0.f;
      if (dg != 0.f) {
        float ax = xc * dg;
        ax += (xl * a.dt) * lxl + (xr * a.dt) * lxu;
        ax += (xym.v[q] * a.dt) * ly0.v[q] + (xyp.v[q] * a.dt) * ly1.v[q];
        ax += (xzm.v[q] * a.dt) * lz0.v[q] + (xzp.v[q] * a.dt) * lz1.v[q];
        r = z - ax;
      }
      rr.v[q] = r; xs.v[q] = xc;
      s += (double)r;
    }
    st4(a.r + o, rr);
    st4(a.xs + o, xs);
  };
  WL_SPLIT_DISPATCH(MODE, cls, body);
  }
  block_reduce_partial(s, -INFINITY, a.w, MODE == 0 ? -1 : bx + a.uf.nx * (by + a.uf.ny * bz));
}

// ---------------------------------------------------------------- K-PROJ, 4 cells per thread (see k_proj)
__global__ void __launch_bounds__(256, 5) k_proj4(const ProjArgs2 a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_QUAD_FROM_BLOCK(l)
  if (!inb) return;
  const int o = l.off(i, j, k);
  int oxm, oxp, oym, oyp, ozm, ozp;
  quad_nbr(l, a.per, i, j, k, oxm, oxp, oym, oyp, ozm, ozp);
  const F4 x = ld4g(a.x + o), xym = ld4g(a.x + o + oym), xzm = ld4g(a.x + o + ozm);
  const float xxm = __ldg(a.x + o + oxm);
  float* ux = a.u; float* uy = a.u + l.cstride; float* uz = a.u + 2 * l.cstride;
  const F4 u0 = ld4(ux + o), v0 = ld4(uy + o), w0 = ld4(uz + o);
  // (no pf4 here: K-PROJ streams 16 B/cell in and 16 B/cell out at 97 % of the bandwidth; prefetching its four inputs made
  //  it 7 % slower, profiles/r2_c39_*)
  F4 lx, ly, lz;                                        // lower faces only (Flow.jl:228)
  const int cls = tile_class(a.uf, blockIdx.x, blockIdx.y, blockIdx.z);
  if (cls == 0) { lx = ld4g(a.L + o); ly = ld4g(a.L + l.cstride + o); lz = ld4g(a.L + 2 * l.cstride + o); }
  else {
    auto f = [&](int d, int c, int n) { return (cls >= 2 && pattern_zero(a.per, d, c, n)) ? 0.f : a.uf.lu[d]; };
#pragma unroll
    for (int c = 0; c < 4; ++c) lx.v[c] = f(0, i + c, l.n0);
    const float a0 = f(1, j, l.n1), b0 = f(2, k, l.n2);
    ly = F4{{a0, a0, a0, a0}}; lz = F4{{b0, b0, b0, b0}};
  }
  F4 un, vn, wn, pn;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float xl = q == 0 ? xxm : x.v[q > 0 ? q - 1 : 0];
    un.v[q] = u0.v[q] - lx.v[q] * (x.v[q] - xl);
    vn.v[q] = v0.v[q] - ly.v[q] * (x.v[q] - xym.v[q]);
    wn.v[q] = w0.v[q] - lz.v[q] * (x.v[q] - xzm.v[q]);
    pn.v[q] = x.v[q] / a.dt;
  }
  st4(ux + o, un); st4(uy + o, vn); st4(uz + o, wn);
  st4(a.pn + o, pn);
  const bool seam = (a.per.p[0] && (i == 1 || i + 3 == l.n0 - 2)) || (a.per.p[1] && (j == 1 || j == l.n1 - 2)) ||
                    (a.per.p[2] && (k == 1 || k == l.n2 - 2));
  if (seam) {
#pragma unroll
    for (int q = 0; q < 4; ++q) store_images<3>(a.pn, l, a.per, i + q, j, k, pn.v[q]);
  }
  // slab interfaces: the neighbour's p on the ghost plane is x/Δt of the exchanged x ghost -- the same
  // operation on the same operands, so it is written here instead of being exchanged
  if (a.per.zlo && k == 1) {
    F4 g;
#pragma unroll
    for (int q = 0; q < 4; ++q) g.v[q] = xzm.v[q] / a.dt;
    st4(a.pn + o - l.plane, g);
  }
  if (a.per.zhi && k == l.n2 - 2) {
    const F4 xzp = ld4g(a.x + o + l.plane);
    F4 g;
#pragma unroll
    for (int q = 0; q < 4; ++q) g.v[q] = xzp.v[q] / a.dt;
    st4(a.pn + o + l.plane, g);
  }
}

// ---------------------------------------------------------------- K-UP, 4 cells per thread (see k_up)
// prolongate! + perBC! + increment!(ω) (+ ϵ = r·iD for the smoother that follows).  Any coarsening mask.
__global__ void __launch_bounds__(256, 5) k_up4(const UpArgs a) {
  pdl_entry();
  const Lay& l = a.f.l;
  const Lay& lc = a.c.l;
  WL_QUAD_FROM_BLOCK(l)
  if (!inb) return;
  const int o = l.off(i, j, k);
  const bool cx = a.c.c[0] != 0, cy = a.c.c[1] != 0, cz = a.c.c[2] != 0;
  // coarse coordinates of the row and of its (periodically wrapped) neighbours: down(I,c) MLP.jl:7
  const int jm = (a.per.p[1] && j == 1) ? l.n1 - 2 : j - 1, jp = (a.per.p[1] && j == l.n1 - 2) ? 1 : j + 1;
  const int km = (a.per.p[2] && k == 1) ? l.n2 - 2 : k - 1, kp = (a.per.p[2] && k == l.n2 - 2) ? 1 : k + 1;
  const int im = (a.per.p[0] && i == 1) ? l.n0 - 2 : i - 1, ip = (a.per.p[0] && i + 3 == l.n0 - 2) ? 1 : i + 4;
  auto dn = [](bool co, int c) { return co ? (c + 1) >> 1 : c; };
  const int cj = dn(cy, j), ck = dn(cz, k);
  const int c0 = dn(cx, i);
  auto row = [&](int jj, int kk) {                              // coarse values under the quad's four cells
    const float* p = a.c.x + lc.pitch * jj + lc.plane * kk;
    if (cx) { const float v0 = __ldg(p + c0), v1 = __ldg(p + c0 + 1); return F4{{v0, v0, v1, v1}}; }
    return ld4g(p + i);
  };
  const F4 e = row(cj, ck), eym = row(dn(cy, jm), ck), eyp = row(dn(cy, jp), ck), ezm = row(cj, dn(cz, km)), ezp = row(cj, dn(cz, kp));
  const float* pc = a.c.x + lc.pitch * cj + lc.plane * ck;
  const float exm = __ldg(pc + dn(cx, im)), exp_ = __ldg(pc + dn(cx, ip));
  if (a.f.pf > 0 && k + a.f.pf <= l.n2 - 2) { const int q = o + a.f.pf * l.plane; pf4(a.f.r + q); pf4(a.f.x + q); }
  auto body = [&](auto tag) {
  constexpr int CLS = decltype(tag)::value;
  const QuadL ql = quad_coeffs<CLS>(a.f.L, l, a.per, a.f.uf, o, i, j, k);
  const F4 &lx = ql.lx, &ly0 = ql.ly0, &ly1 = ql.ly1, &lz0 = ql.lz0, &lz1 = ql.lz1; const float lx4 = ql.lx4;
  const F4 r0 = ld4(a.f.r + o), x0 = ld4(a.f.x + o);
  F4 d0{{0.f, 0.f, 0.f, 0.f}};
  if (a.write_eps) d0 = quad_iD<CLS>(a.f.iD, ql, a.f.uf, o);
  F4 rn, xn, en;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float lxl = lx.v[q], lxu = q < 3 ? lx.v[q < 3 ? q + 1 : 3] : lx4;
    const float el = q == 0 ? exm : e.v[q > 0 ? q - 1 : 0], er = q == 3 ? exp_ : e.v[q < 3 ? q + 1 : 3];
    float dg = 0.f;
    dg -= lxl + lxu; dg -= ly0.v[q] + ly1.v[q]; dg -= lz0.v[q] + lz1.v[q];
    float s = e.v[q] * dg;
    s += el * lxl + er * lxu;
    s += eym.v[q] * ly0.v[q] + eyp.v[q] * ly1.v[q];
    s += ezm.v[q] * lz0.v[q] + ezp.v[q] * lz1.v[q];
    const float r = r0.v[q] - a.omega * s;
    rn.v[q] = r;
    xn.v[q] = x0.v[q] + a.omega * e.v[q];
    en.v[q] = r * d0.v[q];
  }
  st4(a.f.r + o, rn);
  st4(a.f.x + o, xn);
  if (a.write_eps) {
    st4(a.f.eps + o, en);
    const bool seam = (a.per.p[0] && (i == 1 || i + 3 == l.n0 - 2)) || (a.per.p[1] && (j == 1 || j == l.n1 - 2)) ||
                      (a.per.p[2] && (k == 1 || k == l.n2 - 2));
    if (seam) {
#pragma unroll
      for (int q = 0; q < 4; ++q) store_images<3>(a.f.eps, l, a.per, i + q, j, k, en.v[q]);
    }
  }
  };
  WL_CLASS_DISPATCH(tile_class(a.f.uf, blockIdx.x, blockIdx.y, blockIdx.z), body);
}

// ---------------------------------------------------------------- K-DOWN, 4 cells per thread (see k_down)
// Jacobi! (ϵ=r·iD, perBC!, increment! with ω=1) + restrict! + fill!(coarse.x,0), for levels coarsened
// in every direction.  blockDim (32,4,2): the block's 128x4x2 cells are the children of 64x2x1 coarse cells.
template <int MODE>
__global__ void __launch_bounds__(256, MODE == 1 ? 6 : WL_MG4_MINB_B) k_down4(const DownArgs a) {
  pdl_entry();
  __shared__ float tile[2][4][128];
  const Lay& l = a.f.l;
  const Lay& lc = a.c.l;
  int bx, by, bz, cls;
  if (!split_tile<MODE>(a.f.uf, bx, by, bz, cls)) return;              // (block-uniform: no thread is left at the barrier)
  WL_QUAD_FROM_TILE(l, bx, by, bz)
  double s1 = 0.0; float mx = -INFINITY;
  float sh = 0.f;                                                      // residual mean to remove (first cycle only)
  if (a.first) {
    const float mean = (float)redin_sum(a.in) / (float)a.ncells;
    if (fabsf(mean) > 2.f * 1.1920929e-07f) sh = mean;
  }
  F4 rn{{0.f, 0.f, 0.f, 0.f}};
  auto body = [&](auto tag) {
    constexpr int CLS = decltype(tag)::value;
    const int o = l.off(i, j, k);
    int oxm, oxp, oym, oyp, ozm, ozp;
    quad_nbr(l, a.per, i, j, k, oxm, oxp, oym, oyp, ozm, ozp);
    const QuadL ql = quad_coeffs<CLS>(a.f.L, l, a.per, a.f.uf, o, i, j, k);
    const F4 &lx = ql.lx, &ly0 = ql.ly0, &ly1 = ql.ly1, &lz0 = ql.lz0, &lz1 = ql.lz1; const float lx4 = ql.lx4;
    F4 r = ld4g(a.f.r + o), rym = ld4g(a.f.r + o + oym), ryp = ld4g(a.f.r + o + oyp), rzm = ld4g(a.f.r + o + ozm),
       rzp = ld4g(a.f.r + o + ozp);
    float rxm = __ldg(a.f.r + o + oxm), rxp = __ldg(a.f.r + o + oxp);
    // iD of the quad and of its six neighbours: class 1 -- the tile AND its neighbouring tiles hold the constants and no
    // boundary is near, so every one of them is idu; class 3 -- an x end is the only boundary near, see below; class 2
    // (y/z boundary or interface tiles) loads the neighbours'
    const F4 d = quad_iD<CLS>(a.f.iD, ql, a.f.uf, o);
    F4 dym = d, dyp = d, dzm = d, dzp = d;
    float dxm = a.f.uf.idu, dxp = a.f.uf.idu;
    if (CLS == 3) {
      // a neighbour beyond a boundary face is a ghost cell (iD 0: set_diag! never writes it); any other neighbour row has
      // the quad's x pattern and its own count of y/z boundary faces
      // the y/z neighbours sit at the same x as the quad's cells and have their iD (dym..dzp were initialised with d); the
      // x neighbours are a plain cell or the ghost cell beyond the boundary face (iD 0: set_diag! never writes it)
      if ((ql.xq - 1) & 1) dxm = 0.f;
      if ((ql.xq - 1) & 2) dxp = 0.f;
    } else if (CLS != 1) {
      dym = ld4g(a.f.iD + o + oym); dyp = ld4g(a.f.iD + o + oyp); dzm = ld4g(a.f.iD + o + ozm); dzp = ld4g(a.f.iD + o + ozp);
      dxm = __ldg(a.f.iD + o + oxm); dxp = __ldg(a.f.iD + o + oxp);
    }
    const F4 x0 = ld4(a.f.x + o);
    if (sh != 0.f) {
      // ghost neighbours of non-periodic directions are not shifted by the reference, but their iD and
      // the face coefficient towards them are 0: they contribute a (signed) zero either way
      rxm = rxm - sh; rxp = rxp - sh;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        r.v[q] = r.v[q] - sh; rym.v[q] = rym.v[q] - sh; ryp.v[q] = ryp.v[q] - sh; rzm.v[q] = rzm.v[q] - sh; rzp.v[q] = rzp.v[q] - sh;
      }
    }
    F4 xn;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float lxl = lx.v[q], lxu = q < 3 ? lx.v[q < 3 ? q + 1 : 3] : lx4;
      const float el = q == 0 ? rxm * dxm : r.v[q > 0 ? q - 1 : 0] * d.v[q > 0 ? q - 1 : 0];
      const float er = q == 3 ? rxp * dxp : r.v[q < 3 ? q + 1 : 3] * d.v[q < 3 ? q + 1 : 3];
      const float e0 = r.v[q] * d.v[q];                                // ϵ = r·iD  Poisson.jl:112
      float dg = 0.f;
      dg -= lxl + lxu; dg -= ly0.v[q] + ly1.v[q]; dg -= lz0.v[q] + lz1.v[q];
      float s = e0 * dg;                                               // mult(I,L,D,ϵ)
      s += el * lxl + er * lxu;
      s += (rym.v[q] * dym.v[q]) * ly0.v[q] + (ryp.v[q] * dyp.v[q]) * ly1.v[q];
      s += (rzm.v[q] * dzm.v[q]) * lz0.v[q] + (rzp.v[q] * dzp.v[q]) * lz1.v[q];
      rn.v[q] = r.v[q] - s;                                            // increment!, ω = 1  Poisson.jl:102
      xn.v[q] = x0.v[q] + e0;                                          // Poisson.jl:103
      if (a.first) { s1 += (double)fabsf(r.v[q]); mx = fmaxf(mx, fabsf(r.v[q])); }
    }
    st4(a.f.eps + o, rn);          // neighbours still read r: the new residual is staged in ϵ (host swaps the pointers)
    st4(a.f.x + o, xn);
  };
  if (inb && a.f.pf > 0 && k + a.f.pf <= l.n2 - 2) { const int q = l.off(i, j, k) + a.f.pf * l.plane; pf4(a.f.r + q); pf4(a.f.x + q); }
  if (inb) WL_SPLIT_DISPATCH(MODE, cls, body);
  {
    float* t = &tile[threadIdx.z][threadIdx.y][4 * threadIdx.x];
    t[0] = rn.v[0]; t[1] = rn.v[1]; t[2] = rn.v[2]; t[3] = rn.v[3];
  }
  __syncthreads();
  if (inb && !(threadIdx.y & 1) && !(threadIdx.z & 1)) {             // first child in y and z: two coarse cells
    const int cj = (j + 1) >> 1, ck = (k + 1) >> 1, c0 = (i + 1) >> 1;
    const int oc = lc.off(c0, cj, ck);
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      float sum = 0.f;                                                 // restrict  MLP.jl:13-19, first dimension fastest
#pragma unroll
      for (int kk = 0; kk < 2; ++kk)
#pragma unroll
        for (int jj = 0; jj < 2; ++jj)
#pragma unroll
          for (int ii = 0; ii < 2; ++ii) sum += tile[threadIdx.z + kk][threadIdx.y + jj][4 * threadIdx.x + 2 * p + ii];
      a.c.r[oc + p] = sum;
      a.c.x[oc + p] = 0.f;                                             // fill!(coarse.x,0)  MLP.jl:94
    }
  }
  if (a.first) {
    block_reduce_partial(s1, mx, a.w, MODE == 0 ? -1 : bx + a.f.uf.nx * (by + a.f.uf.ny * bz));
  }
}

}  // namespace wl
