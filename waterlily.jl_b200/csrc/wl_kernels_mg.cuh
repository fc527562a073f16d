// wl_kernels_mg.cuh -- pressure-side kernels: operator, residual!, Jacobi!/GaussSeidelRB!/increment!,
// restrict!/prolongate!, restrictL!/set_diag!.  Reference: src/Poisson.jl, src/MultiLevelPoisson.jl.
// D and iD are never read on the hot path: they are recomputed from L in the reference's own
// summation order (`diag`, Poisson.jl:49-55), which is bit-identical and saves 4-8 B/cell.
#pragma once
#include "wl_common.cuh"

namespace wl {

struct LevArgs {
  Lay l;
  const float* L;
  float* x;
  float* r;
  float* eps;
  const float* z;
  float omega;
  int k0;            // GS colour pass 1..4
  RedWork w;
  Red* out;
  RedIn in;          // residual sum input (k_shift_norms)
  long long ncells;  // interior cells of the GLOBAL level (mean of the residual)
  int want_norms;
  TileFlags uf;      // pattern tiles (k_gs_half_c, k_gs_half_cx)
  int pf;            // planes ahead to prefetch into L2 (k_gs_half_c), 0 = off
  Per per;           // ... and the level's boundary kinds, for their class-2 path
};

#define WL_INTERIOR_IJK(D, l)                                                   \
  const int i = blockIdx.x * blockDim.x + threadIdx.x + 1;                              \
  const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;                      \
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z + 1 : 0;         \
  const bool inb = !(i > l.n0 - 2 || j > l.n1 - 2 || (D == 3 && k > l.n2 - 2));

// `diag(I,L)` Poisson.jl:49-55
template <int D>
__device__ __forceinline__ float diagL(const Lay& l, const float* __restrict__ L, int o) {
  const int st[3] = {1, l.pitch, l.plane};
  float s = 0.f;
#pragma unroll
  for (int d = 0; d < D; ++d) { const float* Ld = L + l.cstride * d; s -= Ld[o] + Ld[o + st[d]]; }
  return s;
}
__device__ __forceinline__ float inv_or_zero(float d) { return d == 0.f ? d : 1.f / d; }   // Poisson.jl:45
// `mult(I,L,D,x)` Poisson.jl:70-76
template <int D>
__device__ __forceinline__ float multA(const Lay& l, const float* __restrict__ L, const float* __restrict__ x, int o) {
  const int st[3] = {1, l.pitch, l.plane};
  float s = x[o] * diagL<D>(l, L, o);
#pragma unroll
  for (int d = 0; d < D; ++d) {
    const float* Ld = L + l.cstride * d;
    s += x[o - st[d]] * Ld[o] + x[o + st[d]] * Ld[o + st[d]];
  }
  return s;
}

// set_diag! (Poisson.jl:43-46): materialises D and iD for the data surface only.
struct DiagArgs { Lay l; const float* L; float* Dg; float* iD; };
template <int D>
__global__ void __launch_bounds__(256) k_set_diag(const DiagArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  if (!inb) return;
  const int o = l.off(i, j, k);
  const float d = diagL<D>(l, a.L, o);
  a.Dg[o] = d;
  a.iD[o] = inv_or_zero(d);
}

// residual! first loop (Poisson.jl:94) + Σr partials.  perBC!(x) has been applied by the caller.
template <int D>
__global__ void __launch_bounds__(256) k_residual(const LevArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  double s = 0.0;
  if (inb) {
    const int o = l.off(i, j, k);
    const float d = diagL<D>(l, a.L, o);
    float r = 0.f;
    if (d != 0.f) {
      const int st[3] = {1, l.pitch, l.plane};
      float ax = a.x[o] * d;
#pragma unroll
      for (int q = 0; q < D; ++q) {
        const float* Lq = a.L + l.cstride * q;
        ax += a.x[o - st[q]] * Lq[o] + a.x[o + st[q]] * Lq[o + st[q]];
      }
      r = a.z[o] - ax;
    }
    a.r[o] = r;
    s = (double)r;
  }
  block_reduce_finish(s, -INFINITY, a.w, a.out);
}

// residual! second loop (Poisson.jl:95-97) fused with the initial L₁/L∞ (MultiLevelPoisson.jl:111):
// s = Σr/N in Float32; if |s| > 2eps: r -= s inside.  Norms of the resulting r.
template <int D>
__global__ void __launch_bounds__(256) k_shift_norms(const LevArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const float s = (float)redin_sum(a.in) / (float)a.ncells;
  const bool shift = fabsf(s) > 2.f * 1.1920929e-07f;
  double s1 = 0.0; float m = -INFINITY;
  // 1-D persistent grid, four tiles per trip with their loads issued together: one load in flight
  // per thread leaves the kernel latency-bound at ~1.5 TB/s (Little's law), four reach the HBM rate
  const int nbx = (l.n0 - 2 + blockDim.x - 1) / blockDim.x, nby = (l.n1 - 2 + blockDim.y - 1) / blockDim.y;
  const int nbz = D == 3 ? (l.n2 - 2 + blockDim.z - 1) / blockDim.z : 1;
  const int nt = nbx * nby * nbz;
  for (int t0 = blockIdx.x; t0 < nt; t0 += 4 * gridDim.x) {
    int o[4]; bool ok[4]; float r[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int t = t0 + u * gridDim.x;
      const int tx = t % nbx, q = t / nbx, ty = q % nby, tz = q / nby;
      const int i = tx * blockDim.x + threadIdx.x + 1, j = ty * blockDim.y + threadIdx.y + 1;
      const int k = D == 3 ? tz * blockDim.z + threadIdx.z + 1 : 0;
      ok[u] = t < nt && i <= l.n0 - 2 && j <= l.n1 - 2 && (D == 2 || k <= l.n2 - 2);
      o[u] = ok[u] ? l.off(i, j, k) : l.off(1, 1, D == 3 ? 1 : 0);
      r[u] = a.r[o[u]];
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (!ok[u]) continue;
      float rr = r[u];
      if (shift) { rr = rr - s; a.r[o[u]] = rr; }
      s1 += (double)fabsf(rr); m = fmaxf(m, fabsf(rr));
    }
  }
  block_reduce_finish(s1, m, a.w, a.out);
}

// ϵ = r·iD inside (Jacobi! Poisson.jl:112, GaussSeidelRB! :142)
template <int D>
__global__ void __launch_bounds__(256) k_eps_init(const LevArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  if (!inb) return;
  const int o = l.off(i, j, k);
  a.eps[o] = a.r[o] * inv_or_zero(diagL<D>(l, a.L, o));
}

// increment! (Poisson.jl:100-104) after the caller's perBC!(ϵ); optional L₁/L∞ of the new r.
template <int D>
__global__ void __launch_bounds__(256) k_increment(const LevArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  double s1 = 0.0; float m = -INFINITY;
  if (inb) {
    const int o = l.off(i, j, k);
    const float r = a.r[o] - a.omega * multA<D>(l, a.L, a.eps, o);
    a.r[o] = r;
    a.x[o] = a.x[o] + a.omega * a.eps[o];
    s1 = (double)fabsf(r); m = fabsf(r);
  }
  if (a.want_norms) block_reduce_finish(s1, m, a.w, a.out);
}

// One red-black half sweep (gauss_rb Poisson.jl:124-128 over half_rangek :130-132).
// Sweep k₀ updates the cells whose Julia index sum has parity k₀+1, in the last dimension only up
// to Julia index 2·(size÷2)−1 (an odd-sized last dimension leaves its last interior cell unswept).
template <int D>
__global__ void __launch_bounds__(256) k_gs_half(const LevArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  if (!inb) return;
  const int nl = D == 3 ? l.n2 : l.n1;
  const int cl = D == 3 ? k : j;
  if (cl > 2 * (nl / 2) - 2) return;
  const int jsum = i + j + (D == 3 ? k : 0) + D;          // Julia index sum
  if (((jsum ^ (a.k0 + 1)) & 1) != 0) return;
  const int o = l.off(i, j, k);
  const int st[3] = {1, l.pitch, l.plane};
  float s = a.r[o];
  float dg = 0.f;
#pragma unroll
  for (int q = 0; q < D; ++q) {
    const float* Lq = a.L + l.cstride * q;
    const float lo = Lq[o], hi = Lq[o + st[q]];
    dg -= lo + hi;
    s -= a.eps[o - st[q]] * lo + a.eps[o + st[q]] * hi;
  }
  a.eps[o] = s * inv_or_zero(dg);
}

// The update of one cell, ϵ = (r - Σ L·ϵ_nb)·iD (Poisson.jl:142-147), by tile class.  The arithmetic on ϵ is the same in
// every branch; what differs is where L and iD = inv_or_zero(-ΣL) come from: loaded (class 0), the level's constants
// (class 1: idu is that expression evaluated once on the host), or constants with the boundary faces zeroed (class 2).
// In class 2 a row that touches no y/z boundary -- a warp-uniform test -- needs only two selects for the x ends, and its
// iD is idu or idx1 (one zero x face; 0 + lu == lu + 0 == lu, so one constant serves both ends).
template <int D>
__device__ __forceinline__ float gs_update(const LevArgs& a, int cls, int i, int j, int k, int o, float r0,
                                           const float (&em)[3], const float (&ep)[3]) {
  const Lay& l = a.l;
  float lo[3], hi[3];
  if (D == 3 && cls == 1) {
    float s = r0;
#pragma unroll
    for (int q = 0; q < 3; ++q) s -= em[q] * a.uf.lu[q] + ep[q] * a.uf.lu[q];
    return s * a.uf.idu;
  }
  const int c3[3] = {i, j, k}, N3[3] = {l.n0, l.n1, l.n2};
  if (D == 3 && cls >= 2) {
    const bool yz_plain = !(pattern_zero(a.per, 1, j, l.n1) || pattern_zero(a.per, 1, j + 1, l.n1) ||
                            pattern_zero(a.per, 2, k, l.n2) || pattern_zero(a.per, 2, k + 1, l.n2));
    if (yz_plain) {
      const bool z0 = pattern_zero(a.per, 0, i, l.n0), z1 = pattern_zero(a.per, 0, i + 1, l.n0);
      float s = r0;
      s -= em[0] * (z0 ? 0.f : a.uf.lu[0]) + ep[0] * (z1 ? 0.f : a.uf.lu[0]);
      s -= em[1] * a.uf.lu[1] + ep[1] * a.uf.lu[1];
      s -= em[2] * a.uf.lu[2] + ep[2] * a.uf.lu[2];
      return s * ((z0 || z1) ? a.uf.idx1 : a.uf.idu);
    }
#pragma unroll
    for (int q = 0; q < D; ++q) {
      lo[q] = pattern_zero(a.per, q, c3[q], N3[q]) ? 0.f : a.uf.lu[q];
      hi[q] = pattern_zero(a.per, q, c3[q] + 1, N3[q]) ? 0.f : a.uf.lu[q];
    }
  } else {
    const int st[3] = {1, l.pitch, l.plane};
#pragma unroll
    for (int q = 0; q < D; ++q) { const float* Lq = a.L + l.cstride * q; lo[q] = Lq[o]; hi[q] = Lq[o + st[q]]; }
  }
  float s = r0, dg = 0.f;
#pragma unroll
  for (int q = 0; q < D; ++q) {
    dg -= lo[q] + hi[q];
    s -= em[q] * lo[q] + ep[q] * hi[q];
  }
  return s * inv_or_zero(dg);
}

// The same half sweep with the threads COMPACTED onto the colour being updated: thread tx owns the
// x-pair (2tx+1, 2tx+2) of its row and updates the member of the sweep's colour, so every lane of a
// warp does work (the one-thread-per-cell version idles half of them and is issue-bound).
// gridDim.x = ceil((n0-2)/64).  Same arithmetic, same order: bit-identical.
template <int D>
__global__ void __launch_bounds__(256) k_gs_half_c(const LevArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  // a block covers 64 x 4 x 2 cells: half a tile in x.  Pattern tiles: the coefficients are known without loading them.
  // The class byte is fetched first and the loads every class needs are issued before the branch on it (one memory
  // round trip, not two in a row).
  const int cls = D == 3 ? tile_class(a.uf, blockIdx.x >> 1, blockIdx.y, blockIdx.z) : 0;
  const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z + 1 : 0;
  const int i = 1 + 2 * (blockIdx.x * blockDim.x + threadIdx.x) + ((j + k + D + a.k0) & 1);
  if (i > l.n0 - 2 || j > l.n1 - 2 || (D == 3 && k > l.n2 - 2)) return;
  const int nl = D == 3 ? l.n2 : l.n1;
  const int cl = D == 3 ? k : j;
  if (cl > 2 * (nl / 2) - 2) return;
  const int o = l.off(i, j, k);
  const int st[3] = {1, l.pitch, l.plane};
  float em[3] = {0.f, 0.f, 0.f}, ep[3] = {0.f, 0.f, 0.f};
  const float r0 = a.r[o];
#pragma unroll
  for (int q = 0; q < D; ++q) { em[q] = a.eps[o - st[q]]; ep[q] = a.eps[o + st[q]]; }
  // The sweep is bound by memory latency (ncu: 57 % of its stall samples on the first use of a load, DRAM at 50 %), and
  // blocks are issued in z order: pull the lines of the block that will run a few block layers later into L2 now
  if (D == 3 && a.pf > 0 && k + a.pf <= l.n2 - 2) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(a.r + o + a.pf * l.plane));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(a.eps + o + a.pf * l.plane));
  }
  a.eps[o] = gs_update<D>(a, cls, i, j, k, o, r0, em, ep);
}

// The half sweep with TWO cells per thread: rows j and j+2 of one tile (same colour, same x, same tile class), so that
// indices, bounds, colour and the class byte -- 72 of the one-cell kernel's ~120 instructions, ncu profiles/r2_c47_* -- are
// paid once per two updates.  With the L2 prefetch in place the sweep is bound by instruction issue (80 %), not by
// latency.  Block (32,2,4) = 64 x 4 x 4 cells: half a tile in x, two tile layers in z; a warp is one row pair.
// Same per-cell arithmetic (gs_update): bit-identical.  3D, non-distributed levels.  Level-1 launch on the sphere grid:
// 91.7 -> 83.0 us; four cells per thread (rows j, j+2 on planes k, k+2) measured 85.3 us (profiles/r2_c48_*).
__global__ void __launch_bounds__(256) k_gs_half_c2(const LevArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int k = 1 + 4 * blockIdx.z + threadIdx.z;
  const int cls = tile_class(a.uf, blockIdx.x >> 1, blockIdx.y, (k - 1) >> 1);
  const int j = 1 + 4 * blockIdx.y + threadIdx.y;                    // second row: j + 2
  const int i = 1 + 2 * (blockIdx.x * blockDim.x + threadIdx.x) + ((j + k + 3 + a.k0) & 1);
  if (i > l.n0 - 2 || j > l.n1 - 2 || k > l.n2 - 2 || k > 2 * (l.n2 / 2) - 2) return;
  const bool two = j + 2 <= l.n1 - 2;
  const int o = l.off(i, j, k), o2 = o + 2 * l.pitch;
  const float* e = a.eps + o;
  const float* r = a.r + o;
  float em[3], ep[3], fm[3], fp[3];
  const float r0 = r[0];
  em[0] = e[-1]; ep[0] = e[1]; em[1] = e[-l.pitch]; ep[1] = e[l.pitch]; em[2] = e[-l.plane]; ep[2] = e[l.plane];
  float r1 = 0.f;
  if (two) {
    const float* e2 = e + 2 * l.pitch;
    r1 = r[2 * l.pitch];
    fm[0] = e2[-1]; fp[0] = e2[1]; fm[1] = ep[1]; fp[1] = e2[l.pitch]; fm[2] = e2[-l.plane]; fp[2] = e2[l.plane];
  }
  const int pfd = a.pf;
  if (pfd > 0 && k + pfd <= l.n2 - 2) {
    const float* pe = e + (long long)pfd * l.plane;
    const float* pr = r + (long long)pfd * l.plane;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(pe)); asm volatile("prefetch.global.L2 [%0];" ::"l"(pr));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(pe + 2 * l.pitch)); asm volatile("prefetch.global.L2 [%0];" ::"l"(pr + 2 * l.pitch));
  }
  a.eps[o] = gs_update<3>(a, cls, i, j, k, o, r0, em, ep);
  if (two) a.eps[o2] = gs_update<3>(a, cls, i, j + 2, k, o2, r1, fm, fp);
}

// ---- pattern-tile classification (see TileFlags in wl_common.cuh).  Stage 1: core[t] = every cell of tile t -- its
// interior cells, plus the ghost / halo row, column or plane next to a tile at the end of a direction, which the
// kernels read as upper faces -- holds the pattern value in all three components (and the same for its first / last 32
// columns alone).  Stage 2: class[t] = 0 unless t and its 26 neighbours (wrapped across periodic directions, absent beyond
// the end of the others; of a neighbour to the left or right only the 32 columns next to t) hold the pattern;
// then 2 if the tile touches the end of a non-periodic direction (physical boundary or slab interface), else 1.
struct TileArgs { Lay l; const float* L; unsigned char* core; unsigned char* cls; int nx, ny, nz; Per per; float lu[3]; int* list; int* count; };
__global__ void __launch_bounds__(256) k_tile_core(const TileArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int N[3] = {l.n0, l.n1, l.n2};
  const int t3[3] = {(int)blockIdx.x, (int)blockIdx.y, (int)blockIdx.z}, nt[3] = {a.nx, a.ny, a.nz}, ts[3] = {UT_X, UT_Y, UT_Z};
  int lo[3], hi[3];                                    // index range of the tile, widened to the UPPER ghost at the end of a
#pragma unroll                                         // direction (read as the upper face of the last cell; nothing reads a
  for (int d = 0; d < 3; ++d) {                        // lower ghost or a ghost's tangential components)
    lo[d] = 1 + ts[d] * t3[d]; hi[d] = min(lo[d] + ts[d] - 1, N[d] - 2);
    if (t3[d] == nt[d] - 1) hi[d] = N[d] - 1;
  }
  const int ex = hi[0] - lo[0] + 1, ey = hi[1] - lo[1] + 1, ez = hi[2] - lo[2] + 1;
  int bad = 0;
  for (int q = threadIdx.x; q < ex * ey * ez; q += blockDim.x) {
    const int c[3] = {lo[0] + q % ex, lo[1] + (q / ex) % ey, lo[2] + q / (ex * ey)};
    const int o = l.off(c[0], c[1], c[2]);
    int ng = 0, g = -1;
#pragma unroll
    for (int d = 0; d < 3; ++d) if (c[d] == N[d] - 1) { ++ng; g = d; }
    if (ng > 1) continue;                              // edges and corners of the array: never read
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      if (ng == 1 && d != g) continue;                 // upper ghost: only its normal component is an upper face
      const float want = pattern_zero(a.per, d, c[d], N[d]) ? 0.f : a.lu[d];
      if (a.L[l.cstride * d + o] != want) bad |= 1 | (c[0] < lo[0] + 32 ? 2 : 0) | (c[0] > hi[0] - 32 ? 4 : 0);
    }
  }
  // bit 0: the whole tile holds the pattern; bits 1, 2: its first / last 32 columns do -- what an x neighbour needs of it
  // (the kernels reach one cell, and that cell's faces, beyond a tile: a 128-cell-wide halo in x would be wasteful)
  const int b0 = __syncthreads_or(bad & 1), b1 = __syncthreads_or(bad & 2), b2 = __syncthreads_or(bad & 4);
  if (threadIdx.x == 0) a.core[blockIdx.x + a.nx * (blockIdx.y + a.ny * blockIdx.z)] = (b0 ? 0 : 1) | (b1 ? 0 : 2) | (b2 ? 0 : 4);
}
__global__ void __launch_bounds__(256) k_tile_class(const TileArgs a) {
  pdl_entry();
  const int nt = a.nx * a.ny * a.nz;
  const int n[3] = {a.nx, a.ny, a.nz};
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < nt; t += gridDim.x * blockDim.x) {
    const int c[3] = {t % a.nx, (t / a.nx) % a.ny, t / (a.nx * a.ny)};
    int ok = 1, edge = 0, edge_yz = 0;
    // a tile is an edge tile if it holds a boundary-adjacent cell (index 1 or N-2: the first and the last tile of a
    // direction) or has one as a neighbour: the tile before the last when the last one is a single cell thick
    const int N3[3] = {a.l.n0, a.l.n1, a.l.n2}, ts[3] = {UT_X, UT_Y, UT_Z};
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const int last_thick = (N3[d] - 2) - ts[d] * (n[d] - 1);
      const int e = !a.per.p[d] && (c[d] == 0 || c[d] == n[d] - 1 || (c[d] == n[d] - 2 && last_thick == 1));
      edge |= e;
      if (d > 0) edge_yz |= e;
    }
    for (int dz = -1; dz <= 1 && ok; ++dz)
      for (int dy = -1; dy <= 1 && ok; ++dy)
        for (int dx = -1; dx <= 1 && ok; ++dx) {
          int q[3] = {c[0] + dx, c[1] + dy, c[2] + dz};
          bool absent = false;
#pragma unroll
          for (int d = 0; d < 3; ++d) {
            if (q[d] < 0 || q[d] >= n[d]) {
              if (a.per.p[d]) q[d] = q[d] < 0 ? n[d] - 1 : 0; else absent = true;
            }
          }
          if (!absent) ok = a.core[q[0] + a.nx * (q[1] + a.ny * q[2])] & (dx == 0 ? 1 : (dx < 0 ? 4 : 2));
        }
    a.cls[t] = (unsigned char)(ok ? (edge ? (edge_yz ? 2 : 3) : 1) : 0);
  }
}

// The class-0/2 tiles in ascending order (the second launch of the split kernels walks them: neighbours in the list are
// neighbours in memory).  One block: every thread counts its contiguous chunk, the block scans the counts, every thread
// writes its chunk's ids from its offset.
__global__ void __launch_bounds__(1024) k_tile_list(const TileArgs a) {
  pdl_entry();
  __shared__ int cnt[1024];
  const int nt = a.nx * a.ny * a.nz, per = (nt + 1023) / 1024;
  const int t0 = min(nt, (int)threadIdx.x * per), t1 = min(nt, t0 + per);
  int c = 0;
  for (int t = t0; t < t1; ++t) c += (a.cls[t] == 0 || a.cls[t] == 2);
  cnt[threadIdx.x] = c;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {                       // inclusive scan
    const int v = threadIdx.x >= off ? cnt[threadIdx.x - off] : 0;
    __syncthreads();
    cnt[threadIdx.x] += v;
    __syncthreads();
  }
  int o = cnt[threadIdx.x] - c;
  for (int t = t0; t < t1; ++t) if (a.cls[t] == 0 || a.cls[t] == 2) a.list[o++] = t;
  if (threadIdx.x == 1023) *a.count = cnt[1023];
}

// The half sweep with the slab-interface exchange FUSED in (multi-GPU, peer-mapped arenas; host side in
// smooth2).  Blocks of the two boundary slabs run first.  Producer: a thread that updates a cell of plane 1
// (plane n2-2) also stores it into the lower (upper) neighbour's ghost plane over NVLink; the last block of
// each boundary slab raises that neighbour's arrival flag.  Consumer: the boundary-slab blocks of the NEXT
// sweep (or of the increment) wait for their arrival flag before they read the ghost plane; everything
// else of that kernel overlaps the transfer.  No separate exchange kernel, no ready handshake: having seen
// arrival(s-1) from a neighbour proves that its boundary blocks finished sweep s-1, the only readers of the
// ghost cells sweep s overwrites.
struct GsLink {
  float* peer_lo; float* peer_hi;                    // the neighbours' copies of ϵ (same arena offset)
  unsigned int* my; unsigned int* nb_lo; unsigned int* nb_hi;   // arena headers: [8],[9] tickets, [10] arrival
                                                     // from the lower neighbour, [11] from the upper
  unsigned int push_seq_lo, push_seq_hi, wait_seq_lo, wait_seq_hi;
  int push_lo, push_hi, wait_lo, wait_hi;
};
// TWO: a thread takes rows j and j+2 of its tile, as in k_gs_half_c2 (block (32,2,2) on the same grid of 64 x 4 x 2-cell
// blocks, so the layer order, the tickets and the flags are those of the one-row form).
template <int D, bool TWO = false>
__global__ void __launch_bounds__(256) k_gs_half_cx(const LevArgs a, const GsLink x) {
  pdl_entry();
  const Lay& l = a.l;
  const int gz = gridDim.z;
  const int zb = blockIdx.z == 0 ? 0 : (blockIdx.z == 1 ? gz - 1 : (int)blockIdx.z - 1);   // boundary slabs first
  const bool lo_blk = zb == 0, hi_blk = zb == gz - 1;
  const int tid = threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
  if ((lo_blk && x.wait_lo) || (hi_blk && x.wait_hi)) {
    if (tid == 0) {
      if (lo_blk && x.wait_lo) wait_flag(x.my + 10, x.wait_seq_lo, x.my + 5);
      if (hi_blk && x.wait_hi) wait_flag(x.my + 11, x.wait_seq_hi, x.my + 5);
    }
    __syncthreads();
  }
  constexpr int NR = TWO ? 2 : 1;
  const int jb = (TWO ? 4 * blockIdx.y : blockIdx.y * blockDim.y) + threadIdx.y + 1;
  const int k = zb * blockDim.z + threadIdx.z + 1;
  const int i = 1 + 2 * (blockIdx.x * blockDim.x + threadIdx.x) + ((jb + k + D + a.k0) & 1);   // rows jb, jb+2: same colour
  if (i <= l.n0 - 2 && jb <= l.n1 - 2 && k <= l.n2 - 2 && k <= 2 * (l.n2 / 2) - 2) {
    const int st[3] = {1, l.pitch, l.plane};
    const int cls = tile_class(a.uf, blockIdx.x >> 1, TWO ? blockIdx.y : (jb - 1) >> 2, zb);
    int o[NR]; bool on[NR]; float r0[NR], em[NR][3], ep[NR][3];
#pragma unroll
    for (int q = 0; q < NR; ++q) {
      on[q] = jb + 2 * q <= l.n1 - 2;
      o[q] = l.off(i, jb + 2 * q, k);
      if (!on[q]) continue;
      r0[q] = a.r[o[q]];
#pragma unroll
      for (int d = 0; d < D; ++d) { em[q][d] = a.eps[o[q] - st[d]]; ep[q][d] = a.eps[o[q] + st[d]]; }   // before the branch on cls
    }
    if (a.pf > 0 && k + a.pf <= l.n2 - 2) {                // as in k_gs_half_c (after the two boundary layers z runs upwards)
#pragma unroll
      for (int q = 0; q < NR; ++q) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.r + o[q] + a.pf * l.plane));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.eps + o[q] + a.pf * l.plane));
      }
    }
#pragma unroll
    for (int q = 0; q < NR; ++q) {
      if (!on[q]) continue;
      const float v = gs_update<D>(a, cls, i, jb + 2 * q, k, o[q], r0[q], em[q], ep[q]);
      a.eps[o[q]] = v;
      if (x.push_lo && k == 1) x.peer_lo[o[q] + l.plane * (l.n2 - 2)] = v;
      if (x.push_hi && k == l.n2 - 2) x.peer_hi[o[q] - l.plane * (l.n2 - 2)] = v;
    }
  }
  const bool sig_lo = lo_blk && x.push_lo, sig_hi = hi_blk && x.push_hi;
  if (sig_lo || sig_hi) {
    __syncthreads();
    if (tid == 0) {
      __threadfence_system();
      const unsigned int nside = gridDim.x * gridDim.y;
      if (sig_lo && atomicAdd(x.my + 8, 1u) == nside - 1) {
        x.my[8] = 0u; __threadfence_system(); st_relaxed_sys(x.nb_lo + 11, x.push_seq_lo);
      }
      if (sig_hi && atomicAdd(x.my + 9, 1u) == nside - 1) {
        x.my[9] = 0u; __threadfence_system(); st_relaxed_sys(x.nb_hi + 10, x.push_seq_hi);
      }
    }
  }
}

// restrict! (MultiLevelPoisson.jl:13-19,49): coarse r = Σ fine children, first dim fastest.
struct XferArgs { Lay lf, lc; const float* src; float* dst; int c[3]; };
template <int D>
__global__ void __launch_bounds__(256) k_restrict(const XferArgs a) {
  pdl_entry();
  const Lay& l = a.lc;
  WL_INTERIOR_IJK(D, l)
  if (!inb) return;
  const int cc[3] = {i, j, k};
  int lo[3], hi[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    if (d >= D) { lo[d] = hi[d] = 0; }
    else if (!a.c[d]) { lo[d] = hi[d] = cc[d]; }
    else { lo[d] = 2 * cc[d] - 1; hi[d] = 2 * cc[d]; }
  }
  float s = 0.f;
  for (int kk = lo[2]; kk <= hi[2]; ++kk)
    for (int jj = lo[1]; jj <= hi[1]; ++jj)
      for (int ii = lo[0]; ii <= hi[0]; ++ii) s += a.src[a.lf.off(ii, jj, kk)];
  a.dst[l.off(i, j, k)] = s;
}
// prolongate! (MultiLevelPoisson.jl:7,50): fine ϵ[I] = coarse x[down(I,c)]
template <int D>
__global__ void __launch_bounds__(256) k_prolongate(const XferArgs a) {
  pdl_entry();
  const Lay& l = a.lf;
  WL_INTERIOR_IJK(D, l)
  if (!inb) return;
  const int ci = a.c[0] ? (i + 1) / 2 : i, cj = a.c[1] ? (j + 1) / 2 : j, ck = (D == 3 && a.c[2]) ? (k + 1) / 2 : k;
  a.dst[l.off(i, j, k)] = a.src[a.lc.off(ci, cj, ck)];
}
// restrictL! interior loop (MultiLevelPoisson.jl:20-26,42-46); BC!(a,0) is applied by the caller.
struct XferLArgs { Lay lf, lc; const float* Lf; float* Lc; int c[3]; };
template <int D>
__global__ void __launch_bounds__(256) k_restrictL(const XferLArgs a) {
  pdl_entry();
  const Lay& l = a.lc;
  WL_INTERIOR_IJK(D, l)
  if (!inb) return;
  const int cc[3] = {i, j, k};
#pragma unroll
  for (int ci = 0; ci < D; ++ci) {
    int lo[3], hi[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      if (d >= D) { lo[d] = hi[d] = 0; }
      else if (!a.c[d]) { lo[d] = hi[d] = cc[d]; }
      else if (d == ci) { lo[d] = hi[d] = 2 * cc[d] - 1; }
      else { lo[d] = 2 * cc[d] - 1; hi[d] = 2 * cc[d]; }
    }
    const float* src = a.Lf + a.lf.cstride * ci;
    float s = 0.f;
    for (int kk = lo[2]; kk <= hi[2]; ++kk)
      for (int jj = lo[1]; jj <= hi[1]; ++jj)
        for (int ii = lo[0]; ii <= hi[0]; ++ii) s += src[a.lf.off(ii, jj, kk)];
    a.Lc[l.cstride * ci + l.off(i, j, k)] = a.c[ci] ? s / 2.f : s;
  }
}

}  // namespace wl
