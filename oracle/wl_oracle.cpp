// wl_oracle.cpp -- TEST INFRASTRUCTURE ONLY (not the product path).
//
// CPU restatement of the WaterLily.jl time-step hot path (TzuYaoHuang/WaterLily.jl v1.8.0),
// written loop-for-loop after the reference's `@loop` call sites so that the CUDA path in
// waterlily.jl_b200/csrc can be checked against it.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load this library.
//
// Parity status: the reference is Julia and Julia is not installed in this image, so the
// reference itself cannot be run here.  This restatement is pinned against every
// known-answer test the reference's own test-suite holds for the path
// (tests/test_oracle_*.py cite them); field-level parity is "pinned by restatement".
//
// Conventions: arrays are column-major exactly as in Julia, one ghost layer per side.
// 0-based index i here == Julia index i+1.  2D grids are stored as n[2]==1 with no ghost
// in the third direction.  Element type `real` (float by default, -DWL_REAL=double for
// the Float64 build used by a few reference tests).
//
// Each function cites the reference file:line it follows (paths relative to the
// reference repository root).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

#ifndef WL_REAL
#define WL_REAL float
#endif
typedef WL_REAL real;

namespace {

struct Grid {
  int D = 3;
  int n[3] = {1, 1, 1};  // sizes including ghosts (n[2]==1 for D==2)
  long s[3] = {1, 1, 1}; // strides
  long cells = 1;
  void init(int D_, const int* ng) {
    D = D_;
    for (int d = 0; d < 3; ++d) n[d] = d < D ? ng[d] : 1;
    s[0] = 1; s[1] = n[0]; s[2] = (long)n[0] * n[1];
    cells = (long)n[0] * n[1] * n[2];
  }
  long idx(int i, int j, int k) const { return i + s[1] * j + s[2] * k; }
  long ninside() const { long c = 1; for (int d = 0; d < D; ++d) c *= n[d] - 2; return c; }
};

struct Box { int lo[3], hi[3]; }; // inclusive, 0-based

// `inside(a)` core.jl:47
Box inside(const Grid& g) {
  Box b;
  for (int d = 0; d < 3; ++d) { b.lo[d] = d < g.D ? 1 : 0; b.hi[d] = d < g.D ? g.n[d] - 2 : 0; }
  return b;
}
Box whole(const Grid& g) {
  Box b;
  for (int d = 0; d < 3; ++d) { b.lo[d] = 0; b.hi[d] = g.n[d] - 1; }
  return b;
}
// `slice(dims,i,j,low)` core.jl:188 (i is 0-based here, low is 0-based)
Box slice(const Grid& g, int i, int j, int low = 0) {
  Box b;
  for (int d = 0; d < 3; ++d) {
    if (d >= g.D) { b.lo[d] = b.hi[d] = 0; }
    else if (d == j) { b.lo[d] = b.hi[d] = i; }
    else { b.lo[d] = low; b.hi[d] = g.n[d] - 1; }
  }
  return b;
}
// `inside_u(dims,j)` core.jl:55-57 : dim j 3:N-1 (Julia) others 2:N
Box inside_u(const Grid& g, int j) {
  Box b;
  for (int d = 0; d < 3; ++d) {
    if (d >= g.D) { b.lo[d] = b.hi[d] = 0; }
    else if (d == j) { b.lo[d] = 2; b.hi[d] = g.n[d] - 2; }
    else { b.lo[d] = 1; b.hi[d] = g.n[d] - 1; }
  }
  return b;
}

// One `@loop` == one call of this (core.jl:125-156).  The body must be independent per I.
template <class F>
inline void loop(const Grid& g, const Box& b, F f) {
  if (b.hi[0] < b.lo[0] || b.hi[1] < b.lo[1] || b.hi[2] < b.lo[2]) return;
  const long work = (long)(b.hi[0] - b.lo[0] + 1) * (b.hi[1] - b.lo[1] + 1) * (b.hi[2] - b.lo[2] + 1);
#pragma omp parallel for collapse(2) schedule(static) if (work > 16384)
  for (int k = b.lo[2]; k <= b.hi[2]; ++k)
    for (int j = b.lo[1]; j <= b.hi[1]; ++j) {
      long base = g.idx(0, j, k);
      for (int i = b.lo[0]; i <= b.hi[0]; ++i) f(base + i, i, j, k);
    }
}

// Reductions: the reference uses Base `sum`/`maximum` in T (pairwise on CPU).  We accumulate
// in double and round once, i.e. return the correctly rounded value of the exact sum.
template <class F>
inline double sum_box(const Grid& g, const Box& b, F f) {
  double tot = 0;
#pragma omp parallel for collapse(2) reduction(+ : tot) schedule(static)
  for (int k = b.lo[2]; k <= b.hi[2]; ++k)
    for (int j = b.lo[1]; j <= b.hi[1]; ++j) {
      long base = g.idx(0, j, k);
      double s = 0;
      for (int i = b.lo[0]; i <= b.hi[0]; ++i) s += (double)f(base + i);
      tot += s;
    }
  return tot;
}
template <class F>
inline real max_box(const Grid& g, const Box& b, F f) {
  real m = -std::numeric_limits<real>::infinity();
#pragma omp parallel for collapse(2) reduction(max : m) schedule(static)
  for (int k = b.lo[2]; k <= b.hi[2]; ++k)
    for (int j = b.lo[1]; j <= b.hi[1]; ++j) {
      long base = g.idx(0, j, k);
      for (int i = b.lo[0]; i <= b.hi[0]; ++i) m = std::max(m, (real)f(base + i));
    }
  return m;
}

inline int coord(int d, int i, int j, int k) { return d == 0 ? i : d == 1 ? j : k; }

// ---------------------------------------------------------------- boundary conditions

// `BC!(a,U,saveexit,perdir,t)` for a constant tuple U, core.jl:200-219.
// For tuple BCs the tangential branch `U+a[I±δ]-U` is an exact copy (pinned by
// test/test_core.jl:24-27, which compares with ==).
void BC(const Grid& g, real* a, const real* U, bool saveexit, const int* per) {
  const int D = g.D;
  for (int i = 0; i < D; ++i) {
    real* ai = a + g.cells * i;
    for (int j = 0; j < D; ++j) {
      const long sj = g.s[j];
      const int N = g.n[j];
      if (per[j]) {
        loop(g, slice(g, 0, j), [&](long I, int, int, int) { ai[I] = ai[I + sj * (N - 2)]; });
        loop(g, slice(g, N - 1, j), [&](long I, int, int, int) { ai[I] = ai[I - sj * (N - 2)]; });
      } else if (i == j) {
        for (int s = 0; s < 2; ++s)
          loop(g, slice(g, s, j), [&](long I, int, int, int) { ai[I] = U[i]; });
        if (!saveexit || i > 0)
          loop(g, slice(g, N - 1, j), [&](long I, int, int, int) { ai[I] = U[i]; });
      } else {
        loop(g, slice(g, 0, j), [&](long I, int, int, int) { ai[I] = ai[I + sj]; });
        loop(g, slice(g, N - 1, j), [&](long I, int, int, int) { ai[I] = ai[I - sj]; });
      }
    }
  }
}

// `BC!(a,uBC::Function,saveexit,perdir,t)` core.jl:201-219: Dirichlet uBC(i,loc(i,I),t) on the normal component,
// `uBC(I) + a[I±δ] − uBC(I±δ)` on the tangential ones.  `i` is passed 1-based, like the reference's closures.
// The reference evaluates that sum under @fastmath (core.jl:139,150), which lets LLVM fold it to a plain copy when
// uBC does not vary along j (its own test demands `==` there, test_core.jl:24-27); the association used here,
// a[I±δ] + (uBC(I) − uBC(I±δ)), has that property exactly and is one of the orders @fastmath permits otherwise.
typedef real (*wlo_fn)(int i, const real* x, real t);
inline real at_loc(wlo_fn f, int i, int a, int b, int c, real t) {      // f(i, loc(i,I), t), 0-based (a,b,c)
  real x[3] = {(real)a - (real)0.5, (real)b - (real)0.5, (real)c - (real)0.5};
  x[i] -= (real)0.5;
  return f(i + 1, x, t);
}
void BC_fn(const Grid& g, real* a, wlo_fn ub, bool saveexit, const int* per, real t) {
  const int D = g.D;
  for (int i = 0; i < D; ++i) {
    real* ai = a + g.cells * i;
    for (int j = 0; j < D; ++j) {
      const long sj = g.s[j];
      const int N = g.n[j];
      const int dj[3] = {j == 0, j == 1, j == 2};
      if (per[j]) {
        loop(g, slice(g, 0, j), [&](long I, int, int, int) { ai[I] = ai[I + sj * (N - 2)]; });
        loop(g, slice(g, N - 1, j), [&](long I, int, int, int) { ai[I] = ai[I - sj * (N - 2)]; });
      } else if (i == j) {
        for (int s = 0; s < 2; ++s)
          loop(g, slice(g, s, j), [&](long I, int x, int y, int z) { ai[I] = at_loc(ub, i, x, y, z, t); });
        if (!saveexit || i > 0)
          loop(g, slice(g, N - 1, j), [&](long I, int x, int y, int z) { ai[I] = at_loc(ub, i, x, y, z, t); });
      } else {
        loop(g, slice(g, 0, j), [&](long I, int x, int y, int z) {
          ai[I] = ai[I + sj] + (at_loc(ub, i, x, y, z, t) - at_loc(ub, i, x + dj[0], y + dj[1], z + dj[2], t)); });
        loop(g, slice(g, N - 1, j), [&](long I, int x, int y, int z) {
          ai[I] = ai[I - sj] + (at_loc(ub, i, x, y, z, t) - at_loc(ub, i, x - dj[0], y - dj[1], z - dj[2], t)); });
      }
    }
  }
}
// either form of flow.uBC at time t
struct BCSpec { const real* U; wlo_fn fn; real t; };
inline void BC(const Grid& g, real* a, const BCSpec& b, bool saveexit, const int* per) {
  if (b.fn) BC_fn(g, a, b.fn, saveexit, per, b.t); else BC(g, a, b.U, saveexit, per);
}

// `exitBC!(u,u⁰,Δt)` core.jl:226-233
void exitBC(const Grid& g, real* u, const real* u0, real dt) {
  Box ex;   // slice(N.-1, N[1], 1, 2): dim 1 at the upper ghost, others interior
  Box in;   // slice(N.-1, 2, 1, 2): first interior face plane
  for (int d = 0; d < 3; ++d) {
    if (d >= g.D) { ex.lo[d] = ex.hi[d] = in.lo[d] = in.hi[d] = 0; }
    else if (d == 0) { ex.lo[d] = ex.hi[d] = g.n[0] - 1; in.lo[d] = in.hi[d] = 1; }
    else { ex.lo[d] = in.lo[d] = 1; ex.hi[d] = in.hi[d] = g.n[d] - 2; }
  }
  long len = 1;
  for (int d = 1; d < g.D; ++d) len *= g.n[d] - 2;
  real U = (real)sum_box(g, in, [&](long I) { return u[I]; }) / (real)len;
  loop(g, ex, [&](long I, int, int, int) { u[I] = u0[I] - U * dt * (u0[I] - u0[I - 1]); });
  real flux = (real)sum_box(g, ex, [&](long I) { return u[I]; }) / (real)len - U;
  loop(g, ex, [&](long I, int, int, int) { u[I] -= flux; });
}

// `perBC!(a,perdir)` core.jl:239-243
void perBC(const Grid& g, real* a, const int* per) {
  for (int j = 0; j < g.D; ++j) {
    if (!per[j]) continue;
    const long sj = g.s[j];
    const int N = g.n[j];
    loop(g, slice(g, 0, j), [&](long I, int, int, int) { a[I] = a[I + sj * (N - 2)]; });
    loop(g, slice(g, N - 1, j), [&](long I, int, int, int) { a[I] = a[I - sj * (N - 2)]; });
  }
}

// ---------------------------------------------------------------- momentum side

// `median(a,b,c)` Flow.jl:27-36
inline real median3(real a, real b, real c) {
  if (a > b) { if (b >= c) return b; if (a > c) return c; }
  else       { if (b <= c) return b; if (a < c) return c; }
  return a;
}
// schemes Flow.jl:4-6
inline real quick(real u, real c, real d) {
  return median3((5 * c + 2 * d - u) / 6, c, median3(10 * c - 9 * u, c, d));
}
inline real vanLeer(real u, real c, real d) {
  return (c <= std::min(u, d) || c >= std::max(u, d)) ? c : c + (d - c) * (c - u) / (d - u);
}
inline real cds(real, real c, real d) { return (c + d) / 2; }
inline real lam(int scheme, real u, real c, real d) {
  return scheme == 0 ? quick(u, c, d) : scheme == 1 ? cds(u, c, d) : vanLeer(u, c, d);
}

// `conv_diff!(r,u,Φ,λ;ν,perdir)` Flow.jl:38-62
void conv_diff(const Grid& g, real* r, const real* u, real* Phi, int scheme, real nu, const int* per) {
  const int D = g.D;
  std::fill(r, r + g.cells * D, (real)0);                           // Flow.jl:39
  for (int i = 0; i < D; ++i) {
    const real* ui = u + g.cells * i;
    real* ri = r + g.cells * i;
    const long si = g.s[i];
    for (int j = 0; j < D; ++j) {
      const real* uj = u + g.cells * j;
      const long sj = g.s[j];
      const int N = g.n[j];
      auto uadv = [&](long I) { return (uj[I] + uj[I - si]) / 2; };  // ϕ(i,CI(I,j),u) Flow.jl:3
      auto dif = [&](long I) { return nu * (ui[I] - ui[I - sj]); };   // ν*∂(j,CI(I,i),u) Flow.jl:1
      auto up = [&](long I) { return lam(scheme, ui[I + sj], ui[I], ui[I - sj]); };
      // lower boundary Flow.jl:56,60-61
      if (!per[j]) {
        loop(g, slice(g, 1, j, 1), [&](long I, int, int, int) {
          real ua = uadv(I);
          real f = ua > 0 ? ua * ((ui[I] + ui[I - sj]) / 2) : ua * up(I);     // ϕuL Flow.jl:10
          ri[I] += f - dif(I);
        });
      } else {
        loop(g, slice(g, 1, j, 1), [&](long I, int, int, int) {
          real ua = uadv(I);
          long Ip = I + sj * (N - 4);   // CIj(j,·,N[j]-2): Julia index N-2 == 0-based N-3; here I_j==1
          real f = ua > 0 ? ua * lam(scheme, ui[Ip], ui[I - sj], ui[I]) : ua * up(I);   // ϕuP Flow.jl:9
          Phi[I] = f - dif(I);
          ri[I] += Phi[I];
        });
      }
      // inner cells Flow.jl:47-49 (two launches)
      Box in = inside_u(g, j);
      loop(g, in, [&](long I, int, int, int) {
        real ua = uadv(I);
        real f = ua > 0 ? ua * lam(scheme, ui[I - 2 * sj], ui[I - sj], ui[I]) : ua * up(I); // ϕu Flow.jl:8
        Phi[I] = f - dif(I);
        ri[I] += Phi[I];
      });
      loop(g, in, [&](long I, int, int, int) { ri[I - sj] -= Phi[I]; });
      // upper boundary Flow.jl:57,62
      if (!per[j]) {
        loop(g, slice(g, N - 1, j, 1), [&](long I, int, int, int) {
          real ua = uadv(I);
          real f = ua < 0 ? ua * ((ui[I] + ui[I - sj]) / 2)
                          : ua * lam(scheme, ui[I - 2 * sj], ui[I - sj], ui[I]);          // ϕuR Flow.jl:11
          ri[I - sj] += -f + dif(I);
        });
      } else {
        loop(g, slice(g, N - 1, j, 1), [&](long I, int, int, int) {
          ri[I - sj] -= Phi[I - sj * (N - 2)];   // Φ[CIj(j,I,2)]
        });
      }
    }
  }
}

// `BDIM!(a)` Flow.jl:176-180 with `μddn` Flow.jl:20-26
void BDIM(const Grid& g, real* u, const real* u0, real* f, const real* V, const real* mu0,
          const real* mu1, real dt) {
  const int D = g.D;
  const long tot = g.cells * D;
#pragma omp parallel for schedule(static)
  for (long I = 0; I < tot; ++I) f[I] = u0[I] + dt * f[I] - V[I];
  Box in = inside(g);
  for (int i = 0; i < D; ++i) {
    const long o = g.cells * i;
    loop(g, in, [&](long I, int, int, int) {
      real s = 0;
      for (int j = 0; j < D; ++j)
        s += mu1[I + g.cells * (i + (long)D * j)] * (f[o + I + g.s[j]] - f[o + I - g.s[j]]);
      u[o + I] += s / 2 + V[o + I] + mu0[o + I] * f[o + I];
    });
  }
}

// `scale_u!(a,scale)` Flow.jl:211-214
void scale_u(const Grid& g, real* u, real s) {
  Box in = inside(g);
  for (int i = 0; i < g.D; ++i) {
    real* ui = u + g.cells * i;
    loop(g, in, [&](long I, int, int, int) { ui[I] *= s; });
  }
}

// `div(I,u)` Flow.jl:13-19
inline real divu(const Grid& g, const real* u, long I) {
  real s = 0;
  for (int i = 0; i < g.D; ++i) s += u[g.cells * i + I + g.s[i]] - u[g.cells * i + I];
  return s;
}
// `flux_out(I,u)` Flow.jl:238-244
inline real flux_out(const Grid& g, const real* u, long I) {
  real s = 0;
  for (int i = 0; i < g.D; ++i)
    s += std::max((real)0, u[g.cells * i + I + g.s[i]]) + std::max((real)0, -u[g.cells * i + I]);
  return s;
}

// ---------------------------------------------------------------- pressure side

// `Poisson` struct Poisson.jl:22-39.  Level 1 aliases the Flow's p, μ₀, σ (WaterLily.jl:97).
struct Level {
  Grid g;
  real *L = nullptr, *x = nullptr, *z = nullptr;     // possibly aliased
  std::vector<real> D, iD, eps, r;                   // owned
  std::vector<real> ownL, ownx, ownz;                // storage for coarse levels
  int per[3] = {0, 0, 0};
  int c[3] = {0, 0, 0};                              // coarsening mask from the finer level to this one
};

// `diag(I,L)` Poisson.jl:49-55
inline real diagL(const Grid& g, const real* L, long I) {
  real s = 0;
  for (int i = 0; i < g.D; ++i) s -= L[g.cells * i + I] + L[g.cells * i + I + g.s[i]];
  return s;
}
// `set_diag!(D,iD,L)` Poisson.jl:43-46
void set_diag(Level& p) {
  const Grid& g = p.g;
  loop(g, inside(g), [&](long I, int, int, int) { p.D[I] = diagL(g, p.L, I); });
  loop(g, inside(g), [&](long I, int, int, int) { p.iD[I] = p.D[I] == 0 ? p.D[I] : 1 / p.D[I]; });
}
// `mult(I,L,D,x)` Poisson.jl:70-76
inline real mult(const Grid& g, const real* L, const real* Dg, const real* x, long I) {
  real s = x[I] * Dg[I];
  for (int i = 0; i < g.D; ++i)
    s += x[I - g.s[i]] * L[g.cells * i + I] + x[I + g.s[i]] * L[g.cells * i + I + g.s[i]];
  return s;
}
// `mult!(p,x)` Poisson.jl:63-69
void mult_into_z(Level& p, real* x) {
  const Grid& g = p.g;
  perBC(g, x, p.per);
  std::fill(p.z, p.z + g.cells, (real)0);
  loop(g, inside(g), [&](long I, int, int, int) { p.z[I] = mult(g, p.L, p.D.data(), x, I); });
}
// `residual!(p)` Poisson.jl:92-98
void residual(Level& p) {
  const Grid& g = p.g;
  perBC(g, p.x, p.per);
  loop(g, inside(g), [&](long I, int, int, int) {
    p.r[I] = p.iD[I] == 0 ? (real)0 : p.z[I] - mult(g, p.L, p.D.data(), p.x, I);
  });
  real s = (real)sum_box(g, whole(g), [&](long I) { return p.r[I]; }) / (real)g.ninside();
  if (std::fabs(s) <= 2 * std::numeric_limits<real>::epsilon()) return;
  loop(g, inside(g), [&](long I, int, int, int) { p.r[I] = p.r[I] - s; });
}
// `increment!(p;ω)` Poisson.jl:100-104
void increment(Level& p, real w) {
  const Grid& g = p.g;
  perBC(g, p.eps.data(), p.per);
  loop(g, inside(g), [&](long I, int, int, int) {
    p.r[I] = p.r[I] - w * mult(g, p.L, p.D.data(), p.eps.data(), I);
    p.x[I] = p.x[I] + w * p.eps[I];
  });
}
// `Jacobi!(p;it=1,ω=1)` Poisson.jl:111-114
void Jacobi(Level& p, int it = 1, real w = 1) {
  const Grid& g = p.g;
  for (int n = 0; n < it; ++n) {
    loop(g, inside(g), [&](long I, int, int, int) { p.eps[I] = p.r[I] * p.iD[I]; });
    increment(p, w);
  }
}
// `GaussSeidelRB!(p;it=4,ω)` Poisson.jl:116-148
void GaussSeidelRB(Level& p, int it, real w) {
  const Grid& g = p.g;
  const int D = g.D;
  real* e = p.eps.data();
  loop(g, inside(g), [&](long I, int, int, int) { e[I] = p.r[I] * p.iD[I]; });
  perBC(g, e, p.per);
  // half_rangek Poisson.jl:130-132: last dim Iv ∈ 2:size÷2 (Julia), k = 2Iv-1-(ΣIv_front+k₀)%2
  Box hb = inside(g);
  const int nl = g.n[D - 1];
  hb.lo[D - 1] = 1; hb.hi[D - 1] = nl / 2 - 1;   // 0-based Iv-1 ∈ 1 .. size÷2-1
  for (int k0 = 1; k0 <= it; ++k0) {
    loop(g, hb, [&](long, int i, int j, int k) {
      int c3[3] = {i, j, k};
      int Iv = c3[D - 1] + 1;                    // Julia value of the half-range index
      int front = 0;
      for (int d = 0; d < D - 1; ++d) front += c3[d] + 1;
      int kk = 2 * Iv - 1 - (front + k0) % 2;    // Julia index in the last dim
      c3[D - 1] = kk - 1;
      long I = g.idx(c3[0], c3[1], c3[2]);
      real s = p.r[I];                           // gauss Poisson.jl:116-122
      for (int d = 0; d < D; ++d)
        s -= e[I - g.s[d]] * p.L[g.cells * d + I] + e[I + g.s[d]] * p.L[g.cells * d + I + g.s[d]];
      e[I] = s * p.iD[I];
    });
  }
  increment(p, w);
}
// `pcg!(p;it=6)` Poisson.jl:166-186 ("next" row; kept so the plain-Poisson golden tests run)
void pcg(Level& p, int it = 6) {
  const Grid& g = p.g;
  real *x = p.x, *r = p.r.data(), *e = p.eps.data(), *z = p.z;
  const real epsT = std::numeric_limits<real>::epsilon();
  Box in = inside(g), wh = whole(g);
  loop(g, in, [&](long I, int, int, int) { z[I] = e[I] = r[I] * p.iD[I]; });
  real rho = (real)sum_box(g, wh, [&](long I) { return r[I] * z[I]; });
  if (std::fabs(rho) < 10 * epsT) return;
  bool anyper = p.per[0] || p.per[1] || p.per[2];
  for (int i = 1; i <= it; ++i) {
    perBC(g, e, p.per);
    loop(g, in, [&](long I, int, int, int) { z[I] = mult(g, p.L, p.D.data(), e, I); });
    real alpha = rho / (real)sum_box(g, anyper ? in : wh, [&](long I) { return z[I] * e[I]; });
    if (std::fabs(alpha) < 1e-2 || std::fabs(alpha) > 1e2) return;
    loop(g, in, [&](long I, int, int, int) { x[I] += alpha * e[I]; r[I] -= alpha * z[I]; });
    if (i == it) return;
    loop(g, in, [&](long I, int, int, int) { z[I] = r[I] * p.iD[I]; });
    real rho2 = (real)sum_box(g, wh, [&](long I) { return r[I] * z[I]; });
    if (std::fabs(rho2) < 10 * epsT) return;
    real beta = rho2 / rho;
    loop(g, in, [&](long I, int, int, int) { e[I] = beta * e[I] + z[I]; });
    rho = rho2;
  }
}
// norms Poisson.jl:190-191 (whole array; ghosts of r are 0)
real L1(const Level& p) { return (real)sum_box(p.g, whole(p.g), [&](long I) { return std::fabs(p.r[I]); }); }
real Linf(const Level& p) { return max_box(p.g, whole(p.g), [&](long I) { return std::fabs(p.r[I]); }); }

// `divisible` MultiLevelPoisson.jl:52
inline bool divisible(int N) { return N % 2 == 0 && N > 4; }

// `restrictL!(a,b,c;perdir)` MultiLevelPoisson.jl:42-48 with `restrictL`/`upL` :9-11,20-26
void restrictL(Level& co, const Level& fi) {
  const Grid &gc = co.g, &gf = fi.g;
  const int D = gc.D;
  for (int i = 0; i < D; ++i) {
    loop(gc, inside(gc), [&](long I, int ci, int cj, int ck) {
      int cc[3] = {ci, cj, ck}, lo[3], hi[3];
      for (int d = 0; d < 3; ++d) {
        if (d >= D) { lo[d] = hi[d] = 0; }
        else if (!co.c[d]) { lo[d] = hi[d] = cc[d]; }
        else if (d == i) { lo[d] = hi[d] = 2 * cc[d] - 1; }          // Julia 2I-2
        else { lo[d] = 2 * cc[d] - 1; hi[d] = 2 * cc[d]; }            // Julia 2I-2 : 2I-1
      }
      real s = 0;
      for (int k = lo[2]; k <= hi[2]; ++k)
        for (int j = lo[1]; j <= hi[1]; ++j)
          for (int ii = lo[0]; ii <= hi[0]; ++ii) s += fi.L[gf.cells * i + gf.idx(ii, j, k)];
      co.L[gc.cells * i + I] = co.c[i] ? s / 2 : s;
    });
  }
  real zero[3] = {0, 0, 0};
  BC(gc, co.L, zero, false, co.per);
}
// `restrict!(a,b,c)` MultiLevelPoisson.jl:13-19,49
void restrict_r(Level& co, const Level& fi) {
  const Grid &gc = co.g, &gf = fi.g;
  const int D = gc.D;
  loop(gc, inside(gc), [&](long I, int ci, int cj, int ck) {
    int cc[3] = {ci, cj, ck}, lo[3], hi[3];
    for (int d = 0; d < 3; ++d) {
      if (d >= D) { lo[d] = hi[d] = 0; }
      else if (!co.c[d]) { lo[d] = hi[d] = cc[d]; }
      else { lo[d] = 2 * cc[d] - 1; hi[d] = 2 * cc[d]; }
    }
    real s = 0;
    for (int k = lo[2]; k <= hi[2]; ++k)
      for (int j = lo[1]; j <= hi[1]; ++j)
        for (int ii = lo[0]; ii <= hi[0]; ++ii) s += fi.r[gf.idx(ii, j, k)];
    co.r[I] = s;
  });
}
// `prolongate!(a,b,c)` MultiLevelPoisson.jl:7,50 : fine ϵ[I] = coarse x[down(I,c)]
void prolongate(Level& fi, const Level& co) {
  const Grid &gc = co.g, &gf = fi.g;
  loop(gf, inside(gf), [&](long I, int i, int j, int k) {
    int f3[3] = {i, j, k}, c3[3];
    for (int d = 0; d < 3; ++d) c3[d] = (d < gf.D && co.c[d]) ? (f3[d] + 1) / 2 : f3[d];
    fi.eps[I] = co.x[gc.idx(c3[0], c3[1], c3[2])];
  });
}

struct MLP {
  std::vector<Level> lv;
  std::vector<int16_t> n;
};

void alloc_level_work(Level& p) {
  p.D.assign(p.g.cells, 0); p.iD.assign(p.g.cells, 0);
  p.eps.assign(p.g.cells, 0); p.r.assign(p.g.cells, 0);
}
// `MultiLevelPoisson(x,L,z;maxlevels=10,perdir)` MultiLevelPoisson.jl:68-76 + `restrictML` :33-41
// returns false when fewer than 3 levels can be built (the reference asserts).
bool mlp_build(MLP& ml, const Grid& g, real* x, real* L, real* z, const int* per, int maxlevels) {
  ml.lv.clear();
  ml.lv.reserve(maxlevels + 2);
  ml.lv.emplace_back();
  Level& l0 = ml.lv.back();
  l0.g = g; l0.x = x; l0.L = L; l0.z = z;
  for (int d = 0; d < 3; ++d) l0.per[d] = per[d];
  alloc_level_work(l0);
  set_diag(l0);
  auto anydiv = [](const Level& l) { for (int d = 0; d < l.g.D; ++d) if (divisible(l.g.n[d])) return true; return false; };
  while (anydiv(ml.lv.back()) && (int)ml.lv.size() <= maxlevels) {
    const Level& f = ml.lv.back();
    Level c;
    int ng[3];
    for (int d = 0; d < 3; ++d) {
      c.c[d] = d < f.g.D && divisible(f.g.n[d]);
      ng[d] = c.c[d] ? 1 + f.g.n[d] / 2 : f.g.n[d];
      c.per[d] = f.per[d];
    }
    c.g.init(f.g.D, ng);
    ml.lv.push_back(std::move(c));
    Level& co = ml.lv.back();
    co.ownL.assign(co.g.cells * co.g.D, 0); co.ownx.assign(co.g.cells, 0); co.ownz.assign(co.g.cells, 0);
    co.L = co.ownL.data(); co.x = co.ownx.data(); co.z = co.ownz.data();
    alloc_level_work(co);
    restrictL(co, ml.lv[ml.lv.size() - 2]);
    set_diag(co);
  }
  return ml.lv.size() > 2;
}
// `update!(ml)` MultiLevelPoisson.jl:79-86
void mlp_update(MLP& ml) {
  set_diag(ml.lv[0]);
  for (size_t l = 1; l < ml.lv.size(); ++l) { restrictL(ml.lv[l], ml.lv[l - 1]); set_diag(ml.lv[l]); }
}
// `Vcycle!(ml;l,ω)` MultiLevelPoisson.jl:88-101
void Vcycle(MLP& ml, size_t l, real w) {
  Level &fine = ml.lv[l], &coarse = ml.lv[l + 1];
  Jacobi(fine);                                     // it=1, ω=1: ω is NOT forwarded
  restrict_r(coarse, fine);
  std::fill(coarse.x, coarse.x + coarse.g.cells, (real)0);
  if (l + 2 < ml.lv.size()) Vcycle(ml, l + 1, w);
  GaussSeidelRB(coarse, 4, w);
  prolongate(fine, coarse);
  increment(fine, w);
}
struct Trace { int n; float rinf[64], r1[64], omega[64]; };
// `solver!(ml;tol=2e-3,itmx=32)` MultiLevelPoisson.jl:108-128
int mlp_solver(MLP& ml, double tol, int itmx, Trace* tr) {
  Level& p = ml.lv[0];
  double r1tol = (tol / 10) * (double)p.g.ninside();   // l1n_tol Poisson.jl:194
  double rinftol = tol;
  residual(p);
  real r1 = L1(p), rinf = Linf(p), w = 1;
  int np = 0;
  if (tr) { tr->rinf[0] = rinf; tr->r1[0] = r1; tr->omega[0] = w; }
  while (np < itmx) {
    Vcycle(ml, 0, w);
    GaussSeidelRB(p, 4, w);
    real rnew = L1(p); rinf = Linf(p); ++np;
    if (tr && np < 64) { tr->rinf[np] = rinf; tr->r1[np] = rnew; tr->omega[np] = w; }
    if (rnew >= r1) w = (real)std::max(0.2, 0.9 * (double)w);
    else if (rnew < r1) w = (real)std::min(1.0, 1.02 * (double)w);
    r1 = rnew;
    if ((double)r1 < r1tol && (double)rinf < rinftol) break;
  }
  perBC(p.g, p.x, p.per);
  ml.n.push_back((int16_t)np);
  if (tr) tr->n = np;
  return np;
}
// stand-alone `solver!(p::Poisson;tol=2e-3,itmx=1e3)` Poisson.jl:212-223
int poisson_solver(Level& p, double tol, int itmx) {
  double r1tol = (tol / 10) * (double)p.g.ninside(), rinftol = tol;
  residual(p);
  real r1 = L1(p), rinf = Linf(p);
  int np = 0;
  while (np < itmx) {
    pcg(p); r1 = L1(p); rinf = Linf(p); ++np;
    if ((double)r1 < r1tol && (double)rinf < rinftol) break;
  }
  perBC(p.g, p.x, p.per);
  return np;
}

// ---------------------------------------------------------------- body (parametric AutoBody)

}  // namespace

extern "C" {
// Parametric stand-in for `AutoBody(sdf,map)` with map(x,t) = x - centre(t),
// centre(t) = center + vel*t + acc*t^2 + amp*sin(omega*t).  (AutoBody.jl:1-37)
struct wlo_body {
  int kind;          // 0 NoBody, 1 sphere/circle, 2 cylinder (infinite along `axis`), 3 torus (ring normal `axis`)
  int axis;
  float center[3];
  float radius;      // sphere / cylinder radius, torus major radius
  float minor;       // torus minor radius
  float vel[3], acc[3], amp[3];
  float omega;
};
// One entry of a postfix CSG program over `AutoBody(sdf, RigidMap(x₀,θ;xₚ,V,ω))` leaves (Body.jl:91-107, RigidMap.jl:28-50).
struct wlo_node {
  int op;            // 0 leaf (push), 1 ∪/+ = min, 2 ∩ = max, 3 unary − (complement)
  int kind, axis;    // leaf: 0 NoBody, 1 sphere/circle, 2 cylinder, 3 torus, 4 capsule (segment ±minor along `axis`, radius)
  float radius, minor;
  float x0[3], theta[3], xp[3], V[3], omega[3];   // 2D: θ and ω are theta[2], omega[2]
};
struct wlo_config {
  int D;
  int N[3];          // interior cells per direction
  int perdir[3];
  int exitBC;
  int scheme;        // 0 quick, 1 cds, 2 vanLeer
  float nu;
  float dt0;
  float uBC[3];
  int maxlevels;
  float tol;
  int itmx;
};
}

namespace {

struct Meas { real d; real n[3]; real V[3]; };

// `measure(body::AutoBody,x,t;fastd²)` AutoBody.jl:29-37 with closed-form gradient (J = I)
Meas measure(const wlo_body& b, int D, const real* x, real t, real fastd2) {
  Meas m; m.d = std::numeric_limits<real>::infinity();
  for (int k = 0; k < 3; ++k) m.n[k] = m.V[k] = 0;
  if (b.kind == 0) return m;                                   // NoBody Body.jl:82
  real xi[3] = {0, 0, 0}, g[3] = {0, 0, 0};
  for (int k = 0; k < D; ++k)
    xi[k] = x[k] - ((real)b.center[k] + (real)b.vel[k] * t + (real)b.acc[k] * t * t +
                    (real)b.amp[k] * std::sin((real)b.omega * t));
  real d;
  if (b.kind == 1) {
    real s = 0; for (int k = 0; k < D; ++k) s += xi[k] * xi[k];
    real rt = std::sqrt(s); d = rt - (real)b.radius;
    for (int k = 0; k < D; ++k) g[k] = xi[k] / rt;
  } else if (b.kind == 2) {
    real s = 0; for (int k = 0; k < D; ++k) if (k != b.axis) s += xi[k] * xi[k];
    real rt = std::sqrt(s); d = rt - (real)b.radius;
    for (int k = 0; k < D; ++k) g[k] = k == b.axis ? (real)0 : xi[k] / rt;
  } else {
    real s = 0; for (int k = 0; k < D; ++k) if (k != b.axis) s += xi[k] * xi[k];
    real rho = std::sqrt(s), q = rho - (real)b.radius;
    real len = std::sqrt(q * q + xi[b.axis] * xi[b.axis]); d = len - (real)b.minor;
    for (int k = 0; k < D; ++k) g[k] = k == b.axis ? xi[k] / len : (q / len) * (xi[k] / rho);
  }
  m.d = d;
  if (d * d > fastd2) return m;                                // AutoBody.jl:31
  for (int k = 0; k < D; ++k) if (std::isnan(g[k])) return m;  // AutoBody.jl:33
  real mm = 0; for (int k = 0; k < D; ++k) mm += g[k] * g[k];
  mm = std::sqrt(mm);
  m.d = d / mm;
  for (int k = 0; k < D; ++k) {
    m.n[k] = g[k] / mm;
    m.V[k] = (real)b.vel[k] + 2 * (real)b.acc[k] * t + (real)b.amp[k] * (real)b.omega * std::cos((real)b.omega * t);
  }
  return m;
}
inline real sdf_raw(const wlo_body& b, int D, const real* x, real t) {
  return measure(b, D, x, t, (real)0).d;                       // sdf(body,x,t) AutoBody.jl:19
}
inline bool is_nobody(const wlo_body& b) { return b.kind == 0; }
inline real sdf_fill(const wlo_body& b, int D, const real* x, real t, real) { return sdf_raw(b, D, x, t); }

// ---- SetBody programs over RigidMap leaves ------------------------------------------------------------
struct Prog { int n; const wlo_node* nd; };

// rotation(θ) RigidMap.jl:45-49 (2D = the z rotation), row-major
void rotation(const wlo_node& b, int D, real R[3][3]) {
  const real t1 = D == 3 ? (real)b.theta[0] : 0, t2 = D == 3 ? (real)b.theta[1] : 0, t3 = (real)b.theta[2];
  const real c1 = std::cos(t1), s1 = std::sin(t1), c2 = std::cos(t2), s2 = std::sin(t2), c3 = std::cos(t3), s3 = std::sin(t3);
  if (D == 2) { R[0][0] = c3; R[0][1] = s3; R[1][0] = -s3; R[1][1] = c3; R[0][2] = R[1][2] = R[2][0] = R[2][1] = 0; R[2][2] = 1; return; }
  R[0][0] = c3 * c2;  R[0][1] = c3 * s2 * s1 + s3 * c1;   R[0][2] = -c3 * s2 * c1 + s3 * s1;
  R[1][0] = -s3 * c2; R[1][1] = -s3 * s2 * s1 + c3 * c1;  R[1][2] = s3 * s2 * c1 + c3 * s1;
  R[2][0] = s2;       R[2][1] = -c2 * s1;                 R[2][2] = c2 * c1;
}

// measure(body::AutoBody,x,t;fastd²) AutoBody.jl:29-37 for a library sdf under a RigidMap:
// ξ = R̂(x−x₀−xₚ)+xₚ (RigidMap.jl:38), J = R̂ (:40), n = J'∇sdf(ξ), V = −J\∂ₜmap = V + ω×(x−x₀−xₚ) (:41; the
// reference's LU solve of R̂·V' = R̂·(V+ω×r) differs from this at rounding level only).
Meas measure_leaf(const wlo_node& b, int D, const real* x, real fastd2) {
  Meas m; m.d = std::numeric_limits<real>::infinity();
  for (int k = 0; k < 3; ++k) m.n[k] = m.V[k] = 0;
  if (b.kind == 0) return m;
  real R[3][3]; rotation(b, D, R);
  real r[3] = {0, 0, 0}, xi[3] = {0, 0, 0}, g[3] = {0, 0, 0};
  for (int k = 0; k < D; ++k) r[k] = x[k] - (real)b.x0[k] - (real)b.xp[k];
  for (int k = 0; k < D; ++k) {
    real a = R[k][0] * r[0];
    for (int q = 1; q < D; ++q) a += R[k][q] * r[q];
    xi[k] = a + (real)b.xp[k];
  }
  real d;
  if (b.kind == 1) {
    real s = 0; for (int k = 0; k < D; ++k) s += xi[k] * xi[k];
    real rt = std::sqrt(s); d = rt - (real)b.radius;
    for (int k = 0; k < D; ++k) g[k] = xi[k] / rt;
  } else if (b.kind == 2) {
    real s = 0; for (int k = 0; k < D; ++k) if (k != b.axis) s += xi[k] * xi[k];
    real rt = std::sqrt(s); d = rt - (real)b.radius;
    for (int k = 0; k < D; ++k) g[k] = k == b.axis ? (real)0 : xi[k] / rt;
  } else if (b.kind == 3) {
    real s = 0; for (int k = 0; k < D; ++k) if (k != b.axis) s += xi[k] * xi[k];
    real rho = std::sqrt(s), q = rho - (real)b.radius;
    real len = std::sqrt(q * q + xi[b.axis] * xi[b.axis]); d = len - (real)b.minor;
    for (int k = 0; k < D; ++k) g[k] = k == b.axis ? xi[k] / len : (q / len) * (xi[k] / rho);
  } else {                                                     // capsule: √Σ(ξ − clamp(ξ_axis,±h)ê)² − r
    real y[3] = {xi[0], xi[1], xi[2]};
    const real h = (real)b.minor;
    y[b.axis] = xi[b.axis] - std::max(-h, std::min(xi[b.axis], h));
    real s = 0; for (int k = 0; k < D; ++k) s += y[k] * y[k];
    real len = std::sqrt(s); d = len - (real)b.radius;
    for (int k = 0; k < D; ++k) g[k] = y[k] / len;
  }
  m.d = d;
  if (d * d > fastd2) return m;                                // AutoBody.jl:31
  for (int k = 0; k < D; ++k) if (std::isnan(g[k])) return m;  // AutoBody.jl:33
  real n[3] = {0, 0, 0};
  for (int k = 0; k < D; ++k) {                                // n = J'n AutoBody.jl:35
    real a = R[0][k] * g[0];
    for (int q = 1; q < D; ++q) a += R[q][k] * g[q];
    n[k] = a;
  }
  real mm = 0; for (int k = 0; k < D; ++k) mm += n[k] * n[k];
  mm = std::sqrt(mm);
  m.d = d / mm;
  for (int k = 0; k < D; ++k) m.n[k] = n[k] / mm;
  const float* w = b.omega;
  real cr[3];
  if (D == 2) { cr[0] = (real)w[2] * -r[1]; cr[1] = (real)w[2] * r[0]; cr[2] = 0; }          // ×(a::Number,b) RigidMap.jl:44
  else { cr[0] = (real)w[1] * r[2] - (real)w[2] * r[1]; cr[1] = (real)w[2] * r[0] - (real)w[0] * r[2]; cr[2] = (real)w[0] * r[1] - (real)w[1] * r[0]; }
  for (int k = 0; k < D; ++k) m.V[k] = (real)b.V[k] + cr[k];
  return m;
}

// isless on floats / tuples / vectors (Base): NaN is greatest, −0.0 < 0.0, lexicographic
inline bool fless(real a, real b) { return !std::isnan(a) && (std::isnan(b) || (std::signbit(a) && !std::signbit(b) && a == b) || a < b); }
inline bool fequal(real a, real b) { return (std::isnan(a) && std::isnan(b)) || (a == b && std::signbit(a) == std::signbit(b)); }
bool meas_less(const Meas& a, const Meas& b, int D) {
  if (fless(a.d, b.d)) return true;
  if (!fequal(a.d, b.d)) return false;
  for (int k = 0; k < D; ++k) { if (fless(a.n[k], b.n[k])) return true; if (fless(b.n[k], a.n[k])) return false; }
  for (int k = 0; k < D; ++k) { if (fless(a.V[k], b.V[k])) return true; if (fless(b.V[k], a.V[k])) return false; }
  return false;
}

// measure(body::SetBody,x,t;fastd²) Body.jl:104-107: min / max of the (d,n,V) tuples, (−d,−n,V) for the complement
Meas measure(const Prog& p, int D, const real* x, real, real fastd2) {
  Meas st[8]; int sp = 0;
  for (int i = 0; i < p.n; ++i) {
    const wlo_node& b = p.nd[i];
    if (b.op == 0) st[sp++] = measure_leaf(b, D, x, fastd2);
    else if (b.op == 3) { Meas& a = st[sp - 1]; a.d = -a.d; for (int k = 0; k < D; ++k) a.n[k] = -a.n[k]; }
    else {
      const Meas bb = st[--sp]; Meas& a = st[sp - 1];
      if (b.op == 1) { if (meas_less(bb, a, D)) a = bb; }      // min(x,y) = isless(y,x) ? y : x
      else           { if (meas_less(a, bb, D)) a = bb; }      // max(x,y) = isless(x,y) ? y : x
    }
  }
  return st[0];
}
inline bool is_nobody(const Prog& p) { return p.n == 0 || (p.n == 1 && p.nd[0].kind == 0); }
// measure_sdf!(σ,body,t;fastd²) Body.jl:74: a lone AutoBody gives the raw sdf (AutoBody.jl:19), a SetBody goes
// through measure(...)[1] (Body.jl:66), i.e. the gradient-normalised distance inside the band.
inline real sdf_fill(const Prog& p, int D, const real* x, real t, real fastd2) {
  if (p.n == 1) return measure_leaf(p.nd[0], D, x, (real)0).d;
  return measure(p, D, x, t, fastd2).d;
}

const real PI_R = (real)3.14159265358979323846;
// kernels Body.jl:54-60
// Julia's sinpi/cospi are accurate to the last bit; evaluate in double and round once.
inline real sinpi_r(real d) { return (real)std::sin(3.14159265358979323846 * (double)d); }
inline real cospi_r(real d) { return (real)std::cos(3.14159265358979323846 * (double)d); }
inline real kern(real d) { return (1 + cospi_r(d)) / 2; }
inline real kern0(real d) { return (1 + d + sinpi_r(d) / PI_R) / 2; }
inline real kern1(real d) { return (1 - d * d) / 4 - (d * sinpi_r(d) + (1 + cospi_r(d)) / PI_R) / (2 * PI_R); }
inline real mu0f(real d, real e) {
  return d / e < -1 + std::sqrt(std::nextafter(std::fabs(d), std::numeric_limits<real>::infinity()) - std::fabs(d))
             ? (real)0 : kern0(std::min(d / e, (real)1));
}
inline real mu1f(real d, real e) { return e * kern1(std::max((real)-1, std::min(d / e, (real)1))); }

// ---------------------------------------------------------------- Flow + Simulation

struct Sim {
  wlo_config cfg;
  Grid g;
  int per[3];
  std::vector<real> u, u0, f, p, sigma, V, mu0, mu1;
  std::vector<real> dts;                  // flow.Δt
  MLP ml;
  real eps = 1;
  Trace last_trace[2];
};

// `measure!(a,body;t,ϵ)` Body.jl:28-51
template <class B>
void measure_flow(Sim& s, const B& b, real t, real eps) {
  if (is_nobody(b)) return;                                    // Body.jl:83
  const Grid& g = s.g; const int D = g.D;
  std::fill(s.V.begin(), s.V.end(), (real)0);
  std::fill(s.mu0.begin(), s.mu0.end(), (real)1);
  std::fill(s.mu1.begin(), s.mu1.end(), (real)0);
  const real d2 = (2 + eps) * (2 + eps);
  real* sg = s.sigma.data();
  loop(g, inside(g), [&](long I, int i, int j, int k) {       // measure_sdf! Body.jl:74
    real x[3] = {(real)i - (real)0.5, (real)j - (real)0.5, (real)k - (real)0.5};
    sg[I] = sdf_fill(b, D, x, t, d2);
  });
  loop(g, inside(g), [&](long I, int i, int j, int k) {
    if (sg[I] * sg[I] < d2) {
      for (int c = 0; c < D; ++c) {
        real x[3] = {(real)i - (real)0.5, (real)j - (real)0.5, (real)k - (real)0.5};
        x[c] -= (real)0.5;                                     // loc(i,I) core.jl:177
        Meas m = measure(b, D, x, t, d2);
        real di = std::fabs(m.d) <= (real)0.5 ? m.d : std::copysign(m.d, sg[I]);
        s.V[g.cells * c + I] = m.V[c];
        s.mu0[g.cells * c + I] = mu0f(di, eps);
        for (int jj = 0; jj < D; ++jj) s.mu1[I + g.cells * (c + (long)D * jj)] = mu1f(di, eps) * m.n[jj];
      }
    } else if (sg[I] < 0) {
      for (int c = 0; c < D; ++c) s.mu0[g.cells * c + I] = 0;
    }
  });
  real zero[3] = {0, 0, 0};
  BC(g, s.mu0.data(), zero, false, s.per);
  BC(g, s.V.data(), zero, s.cfg.exitBC != 0, s.per);
}

// `pressure_force(p,df,body,t)` Metrics.jl:116-133 (clobbers f, returns Float64 sums)
template <class B>
void pressure_force(Sim& s, const B& b, real t, double* out) {
  const Grid& g = s.g; const int D = g.D;
  std::fill(s.f.begin(), s.f.end(), (real)0);
  loop(g, inside(g), [&](long I, int i, int j, int k) {
    real x[3] = {(real)i - (real)0.5, (real)j - (real)0.5, (real)k - (real)0.5};
    Meas m = measure(b, D, x, t, (real)1);
    real kq = kern(std::max((real)-1, std::min(m.d, (real)1)));
    for (int c = 0; c < D; ++c) s.f[g.cells * c + I] = s.p[I] * (m.n[c] * kq);
  });
  for (int c = 0; c < 3; ++c) out[c] = 0;
  for (int c = 0; c < D; ++c) {
    const real* fc = s.f.data() + g.cells * c;
    out[c] = sum_box(g, whole(g), [&](long I) { return fc[I]; });
  }
}

// `∂(i,j,I,u)` Metrics.jl:42-45: ∂uᵢ/∂xⱼ at the centre of cell I
inline real dudx(const Grid& g, const real* u, int i, int j, long I) {
  const real* ui = u + g.cells * i;
  if (i == j) return ui[I + g.s[i]] - ui[I];
  return (ui[I + g.s[j]] + ui[I + g.s[j] + g.s[i]] - ui[I - g.s[j]] - ui[I - g.s[j] + g.s[i]]) / 4;
}
// `viscous_force(u,ν,df,body,t)` Metrics.jl:140-152: df[I,:] = -2ν·S(I,u)·nds over inside_u(u) (clobbers f)
template <class B>
void viscous_force(Sim& s, const B& b, real t, double* out) {
  const Grid& g = s.g; const int D = g.D;
  std::fill(s.f.begin(), s.f.end(), (real)0);
  const real m2nu = -2 * (real)s.cfg.nu;
  const real* u = s.u.data();
  loop(g, inside(g), [&](long I, int i, int j, int k) {
    real x[3] = {(real)i - (real)0.5, (real)j - (real)0.5, (real)k - (real)0.5};
    Meas m = measure(b, D, x, t, (real)1);
    real kq = kern(std::max((real)-1, std::min(m.d, (real)1)));
    real nds[3];
    for (int c = 0; c < D; ++c) nds[c] = m.n[c] * kq;
    for (int a = 0; a < D; ++a) {
      real acc = 0;
      for (int c = 0; c < D; ++c) {
        real S = (dudx(g, u, a, c, I) + dudx(g, u, c, a, I)) / 2;          // S(I,u) Metrics.jl:134
        acc += (m2nu * S) * nds[c];
      }
      s.f[g.cells * a + I] = acc;
    }
  });
  for (int c = 0; c < 3; ++c) out[c] = 0;
  for (int c = 0; c < D; ++c) {
    const real* fc = s.f.data() + g.cells * c;
    out[c] = sum_box(g, whole(g), [&](long I) { return fc[I]; });
  }
}

// `pressure_moment(x₀,p,df,body,t)` / `viscous_moment(x₀,u,ν,df,body,t)` Metrics.jl:168-192 with x₀ in T.
// 3D: cross(a,b); 2D: the scalar a₁b₂-a₂b₁ broadcast to both components of df (StaticArrays).
template <class B>
void force_moment(Sim& s, const B& b, real t, const real* x0, int viscous, double* out) {
  const Grid& g = s.g; const int D = g.D;
  std::fill(s.f.begin(), s.f.end(), (real)0);
  const real m2nu = -2 * (real)s.cfg.nu;
  const real* u = s.u.data();
  loop(g, inside(g), [&](long I, int i, int j, int k) {
    real x[3] = {(real)i - (real)0.5, (real)j - (real)0.5, (real)k - (real)0.5};
    Meas m = measure(b, D, x, t, (real)1);
    real kq = kern(std::max((real)-1, std::min(m.d, (real)1)));
    real v[3] = {0, 0, 0}, r[3] = {0, 0, 0};
    for (int c = 0; c < D; ++c) { v[c] = m.n[c] * kq; r[c] = x[c] - x0[c]; }
    if (viscous) {                                                   // S(I,u)*nds
      real w[3] = {0, 0, 0};
      for (int a = 0; a < D; ++a) {
        real acc = 0;
        for (int c = 0; c < D; ++c) acc += ((dudx(g, u, a, c, I) + dudx(g, u, c, a, I)) / 2) * v[c];
        w[a] = acc;
      }
      for (int c = 0; c < D; ++c) v[c] = w[c];
    }
    const real sc = viscous ? m2nu : s.p[I];
    if (D == 3) {
      s.f[I] = sc * (r[1] * v[2] - r[2] * v[1]);
      s.f[g.cells + I] = sc * (r[2] * v[0] - r[0] * v[2]);
      s.f[2 * g.cells + I] = sc * (r[0] * v[1] - r[1] * v[0]);
    } else {
      const real cz = sc * (r[0] * v[1] - r[1] * v[0]);
      s.f[I] = cz; s.f[g.cells + I] = cz;
    }
  });
  for (int c = 0; c < 3; ++c) out[c] = 0;
  for (int c = 0; c < D; ++c) {
    const real* fc = s.f.data() + g.cells * c;
    out[c] = sum_box(g, whole(g), [&](long I) { return fc[I]; });
  }
}

// `accelerate!(r,t,g,U)` Flow.jl:69-73 restricted to space-uniform g(i,t)+dU(i,t)/dt
void accelerate(const Grid& g, real* r, const real* acc) {
  if (!acc) return;
  for (int i = 0; i < g.D; ++i) {
    if (acc[i] == 0) continue;
    real* ri = r + g.cells * i; const real a = acc[i];
#pragma omp parallel for schedule(static)
    for (long I = 0; I < g.cells; ++I) ri[I] += a;
  }
}

// `accelerate!(r,t,g::Function,U::Function)` Flow.jl:69-73: r[Ii] += g(i,loc(Ii),t) + ∂ₜU(i,loc(Ii),t) over ALL of r;
// the caller supplies ∂ₜU in closed form (the reference differentiates U with ForwardDiff).
void accelerate_fn(const Grid& g, real* r, real t, wlo_fn gf, wlo_fn dUdt) {
  if (!gf && !dUdt) return;
  for (int i = 0; i < g.D; ++i) {
    real* ri = r + g.cells * i;
    loop(g, whole(g), [&](long I, int x, int y, int z) {
      real a = gf ? at_loc(gf, i, x, y, z, t) : (real)0;
      if (gf && dUdt) a = a + at_loc(dUdt, i, x, y, z, t); else if (dUdt) a = at_loc(dUdt, i, x, y, z, t);
      ri[I] += a;
    });
  }
}

// `mom_project!(a,b,w,t)` Flow.jl:223-232
void mom_project(Sim& s, real w, const BCSpec& U, Trace* tr) {
  const Grid& g = s.g; const int D = g.D;
  real dt = w * s.dts.back();
  real *x = s.p.data(), *z = s.sigma.data(), *u = s.u.data(); const real* L = s.mu0.data();
  loop(g, inside(g), [&](long I, int, int, int) { z[I] = divu(g, u, I); });
#pragma omp parallel for schedule(static)
  for (long I = 0; I < g.cells; ++I) x[I] *= dt;
  mlp_solver(s.ml, s.cfg.tol, s.cfg.itmx, tr);
  for (int i = 0; i < D; ++i)
    loop(g, inside(g), [&](long I, int, int, int) {
      u[g.cells * i + I] -= L[g.cells * i + I] * (x[I] - x[I - g.s[i]]);
    });
#pragma omp parallel for schedule(static)
  for (long I = 0; I < g.cells; ++I) x[I] /= dt;
  BC(g, u, U, s.cfg.exitBC != 0, s.per);
}

// `CFL(a)` Flow.jl:234-237 -- note `maximum` runs over the whole σ array, ghosts included.
real CFL(Sim& s) {
  const Grid& g = s.g;
  real* sg = s.sigma.data(); const real* u = s.u.data();
  loop(g, inside(g), [&](long I, int, int, int) { sg[I] = flux_out(g, u, I); });
  real m = max_box(g, whole(g), [&](long I) { return sg[I]; });
  return std::min((real)10, 1 / (m + 5 * s.cfg.nu));
}

// `mom_step!(a,b)` Flow.jl:156-167 with mom_predict! :190-196 and mom_correct! :205-210.
// U1 = uBC at t₁ (space-uniform); acc0/acc1 = g+dU/dt at t₀/t₁ (may be null).
void mom_step(Sim& s, const real* U1, const real* acc0, const real* acc1, wlo_fn ub = nullptr, wlo_fn gf = nullptr,
              wlo_fn dUdt = nullptr) {
  const Grid& g = s.g;
  const real nu = s.cfg.nu; const int sch = s.cfg.scheme; const bool ex = s.cfg.exitBC != 0;
  s.u0 = s.u; scale_u(g, s.u.data(), 0);
  const real dt = s.dts.back();
  real t1 = 0; for (real d : s.dts) t1 += d;              // t₁ = sum(a.Δt); t₀ = t₁ − a.Δt[end]  Flow.jl:157
  const real t0 = t1 - dt;
  const BCSpec bc{U1, ub, t1};
  // predictor
  conv_diff(g, s.f.data(), s.u0.data(), s.sigma.data(), sch, nu, s.per);
  accelerate(g, s.f.data(), acc0);
  accelerate_fn(g, s.f.data(), t0, gf, dUdt);
  BDIM(g, s.u.data(), s.u0.data(), s.f.data(), s.V.data(), s.mu0.data(), s.mu1.data(), dt);
  BC(g, s.u.data(), bc, ex, s.per);
  if (ex) exitBC(g, s.u.data(), s.u0.data(), dt);
  mom_project(s, 1, bc, &s.last_trace[0]);
  // corrector
  conv_diff(g, s.f.data(), s.u.data(), s.sigma.data(), sch, nu, s.per);
  accelerate(g, s.f.data(), acc1);
  accelerate_fn(g, s.f.data(), t1, gf, dUdt);
  BDIM(g, s.u.data(), s.u0.data(), s.f.data(), s.V.data(), s.mu0.data(), s.mu1.data(), dt);
  scale_u(g, s.u.data(), (real)0.5);
  BC(g, s.u.data(), bc, ex, s.per);
  mom_project(s, (real)0.5, bc, &s.last_trace[1]);
  s.dts.push_back(CFL(s));
}

}  // namespace

// ---------------------------------------------------------------- C exports (ctypes)
extern "C" {

int wlo_sizeof_real() { return (int)sizeof(real); }

// `Flow(N,uBC;...)` constructor Flow.jl:133-147 (default IC = uBC tuple)
void* wlo_create(const wlo_config* c) {
  Sim* s = new Sim();
  s->cfg = *c;
  int ng[3];
  for (int d = 0; d < 3; ++d) { ng[d] = c->N[d] + 2; s->per[d] = d < c->D ? c->perdir[d] : 0; }
  s->g.init(c->D, ng);
  const Grid& g = s->g; const int D = g.D;
  s->u.assign(g.cells * D, 0);
  for (int i = 0; i < D; ++i) std::fill(s->u.begin() + g.cells * i, s->u.begin() + g.cells * (i + 1), (real)c->uBC[i]);
  s->f.assign(g.cells * D, 0); s->p.assign(g.cells, 0); s->sigma.assign(g.cells, 0);
  s->V.assign(g.cells * D, 0); s->mu0.assign(g.cells * D, 1); s->mu1.assign(g.cells * D * D, 0);
  s->dts.assign(1, (real)c->dt0);
  real U[3] = {(real)c->uBC[0], (real)c->uBC[1], (real)c->uBC[2]};
  BC(g, s->u.data(), U, c->exitBC != 0, s->per);
  exitBC(g, s->u.data(), s->u.data(), 0);
  s->u0 = s->u;
  real zero[3] = {0, 0, 0};
  BC(g, s->mu0.data(), zero, false, s->per);
  return s;
}
void wlo_destroy(void* h) { delete (Sim*)h; }

// after the caller has written an initial condition into u (whole array, `apply!` Flow.jl:81-83)
void wlo_finish_ic(void* h, const real* U) {
  Sim* s = (Sim*)h;
  BC(s->g, s->u.data(), U, s->cfg.exitBC != 0, s->per);
  exitBC(s->g, s->u.data(), s->u.data(), 0);
  s->u0 = s->u;
}
void wlo_measure(void* h, const wlo_body* b, double t, double eps) {
  Sim* s = (Sim*)h; s->eps = (real)eps;
  measure_flow(*s, *b, (real)t, (real)eps);
}
// `pois_ctor(flow)` WaterLily.jl:97 ; returns number of levels, or -1 if the size rule fails
int wlo_build_poisson(void* h) {
  Sim* s = (Sim*)h;
  bool ok = mlp_build(s->ml, s->g, s->p.data(), s->mu0.data(), s->sigma.data(), s->per, s->cfg.maxlevels);
  return ok ? (int)s->ml.lv.size() : -1;
}
void wlo_update_poisson(void* h) { mlp_update(((Sim*)h)->ml); }

real* wlo_field(void* h, int field, int level, long* count, int* dims) {
  Sim* s = (Sim*)h;
  const Grid* g = &s->g; real* p = nullptr; long mult = 1;
  if (field < 10) {
    const int D = g->D;
    switch (field) {
      case 0: p = s->u.data(); mult = D; break;
      case 1: p = s->u0.data(); mult = D; break;
      case 2: p = s->f.data(); mult = D; break;
      case 3: p = s->p.data(); break;
      case 4: p = s->sigma.data(); break;
      case 5: p = s->V.data(); mult = D; break;
      case 6: p = s->mu0.data(); mult = D; break;
      case 7: p = s->mu1.data(); mult = D * D; break;
    }
  } else {
    if (level < 0 || level >= (int)s->ml.lv.size()) return nullptr;
    Level& l = s->ml.lv[level]; g = &l.g;
    switch (field) {
      case 10: p = l.x; break;
      case 11: p = l.r.data(); break;
      case 12: p = l.eps.data(); break;
      case 13: p = l.z; break;
      case 14: p = l.L; mult = g->D; break;
      case 15: p = l.D.data(); break;
      case 16: p = l.iD.data(); break;
    }
  }
  if (count) *count = g->cells * mult;
  if (dims) { dims[0] = g->n[0]; dims[1] = g->n[1]; dims[2] = g->n[2]; }
  return p;
}
int wlo_nlevels(void* h) { return (int)((Sim*)h)->ml.lv.size(); }

real wlo_dt(void* h) { return ((Sim*)h)->dts.back(); }
void wlo_set_dt(void* h, real dt) { ((Sim*)h)->dts.back() = dt; }
int wlo_nsteps(void* h) { return (int)((Sim*)h)->dts.size() - 1; }

// one `mom_step!`; out_n[2] receives the two Poisson iteration counts, returns the new Δt
real wlo_mom_step(void* h, const real* U1, const real* acc0, const real* acc1, int* out_n) {
  Sim* s = (Sim*)h;
  real U[3] = {(real)s->cfg.uBC[0], (real)s->cfg.uBC[1], (real)s->cfg.uBC[2]};
  mom_step(*s, U1 ? U1 : U, acc0, acc1);
  if (out_n) { out_n[0] = s->last_trace[0].n; out_n[1] = s->last_trace[1].n; }
  return s->dts.back();
}
// `mom_step!` with Function uBC / g (Flow.jl:156-167 with :69-73 and core.jl:201-219); null pointers fall back to
// the constant uBC of the configuration / no acceleration.
real wlo_mom_step_fn(void* h, wlo_fn ub, wlo_fn gf, wlo_fn dUdt, int* out_n) {
  Sim* s = (Sim*)h;
  real U[3] = {(real)s->cfg.uBC[0], (real)s->cfg.uBC[1], (real)s->cfg.uBC[2]};
  mom_step(*s, U, nullptr, nullptr, ub, gf, dUdt);
  if (out_n) { out_n[0] = s->last_trace[0].n; out_n[1] = s->last_trace[1].n; }
  return s->dts.back();
}
// Flow constructor's `BC!(u,uBC,exitBC,perdir); exitBC!(u,u,0); u⁰ = copy(u)` (Flow.jl:141-142) for a Function uBC at t = 0
void wlo_finish_ic_fn(void* h, wlo_fn ub) {
  Sim* s = (Sim*)h;
  BC_fn(s->g, s->u.data(), ub, s->cfg.exitBC != 0, s->per, (real)0);
  exitBC(s->g, s->u.data(), s->u.data(), 0);
  s->u0 = s->u;
}
// `apply!((i,x)->f(i,x,t), u)` core.jl:184-187: every index of every component
void wlo_apply_fn(real* u, int D, const int* ng, wlo_fn f, real t) {
  Grid g; g.init(D, ng);
  for (int i = 0; i < D; ++i) {
    real* ui = u + g.cells * i;
    loop(g, whole(g), [&](long I, int x, int y, int z) { ui[I] = at_loc(f, i, x, y, z, t); });
  }
}
void wlo_BC_fn(real* a, int D, const int* ng, wlo_fn ub, int saveexit, const int* per, real t) {
  Grid g; g.init(D, ng); BC_fn(g, a, ub, saveexit != 0, per, t);
}
void wlo_accelerate_fn(real* r, int D, const int* ng, real t, wlo_fn gf, wlo_fn dUdt) {
  Grid g; g.init(D, ng); accelerate_fn(g, r, t, gf, dUdt);
}
void wlo_get_trace(void* h, int which, int* n, float* rinf, float* r1, float* omega) {
  Trace& t = ((Sim*)h)->last_trace[which];
  *n = t.n;
  for (int i = 0; i <= t.n && i < 64; ++i) { rinf[i] = t.rinf[i]; r1[i] = t.r1[i]; omega[i] = t.omega[i]; }
}
int wlo_solver(void* h) {   // stand-alone MultiLevelPoisson solver! on the current x, z
  Sim* s = (Sim*)h;
  return mlp_solver(s->ml, s->cfg.tol, s->cfg.itmx, &s->last_trace[0]);
}
int wlo_poisson_solver(void* h, int itmx) {   // stand-alone single-level pcg solver! (Poisson.jl:212)
  Sim* s = (Sim*)h;
  return poisson_solver(s->ml.lv[0], s->cfg.tol, itmx);
}
void wlo_mult(void* h, real* x) { mult_into_z(((Sim*)h)->ml.lv[0], x); }   // z = A x  (mult!)
real wlo_cfl(void* h) { return CFL(*(Sim*)h); }
void wlo_pressure_force(void* h, const wlo_body* b, double t, double* out) { pressure_force(*(Sim*)h, *b, (real)t, out); }
void wlo_viscous_force(void* h, const wlo_body* b, double t, double* out) { viscous_force(*(Sim*)h, *b, (real)t, out); }
// `update!(meanflow,flow)` Metrics.jl:239-249: the three running-average loops over ALL indices, ε from the caller
void wlo_meanflow_update(void* h, real* P, real* U, real* UU, real eps) {
  Sim* s = (Sim*)h; const Grid& g = s->g; const int D = g.D;
  const real om = 1 - eps;
  for (long I = 0; I < g.cells; ++I) P[I] = eps * s->p[I] + om * P[I];
  for (long I = 0; I < g.cells * D; ++I) U[I] = eps * s->u[I] + om * U[I];
  if (UU)
    for (int i = 0; i < D; ++i)
      for (int j = 0; j < D; ++j)
        for (long I = 0; I < g.cells; ++I) {
          real& q = UU[I + g.cells * (i + (long)D * j)];
          q = eps * (s->u[g.cells * i + I] * s->u[g.cells * j + I]) + om * q;
        }
}
void wlo_measure_csg(void* h, const wlo_node* nd, int n, double eps) {
  Sim* s = (Sim*)h; s->eps = (real)eps;
  measure_flow(*s, Prog{n, nd}, (real)0, (real)eps);
}
void wlo_pressure_force_csg(void* h, const wlo_node* nd, int n, double* out) { pressure_force(*(Sim*)h, Prog{n, nd}, (real)0, out); }
void wlo_viscous_force_csg(void* h, const wlo_node* nd, int n, double* out) { viscous_force(*(Sim*)h, Prog{n, nd}, (real)0, out); }
void wlo_force_moment_csg(void* h, const wlo_node* nd, int n, const double* x0, int viscous, double* out) {
  real x[3] = {(real)x0[0], (real)x0[1], (real)x0[2]};
  force_moment(*(Sim*)h, Prog{n, nd}, (real)0, x, viscous, out);
}
void wlo_measure_point_csg(const wlo_node* nd, int n, int D, const real* x, real fastd2, real* d, real* nn, real* V) {
  Meas m = measure(Prog{n, nd}, D, x, (real)0, fastd2);
  *d = m.d; for (int k = 0; k < 3; ++k) { nn[k] = m.n[k]; V[k] = m.V[k]; }
}
void wlo_force_moment(void* h, const wlo_body* b, double t, const double* x0, int viscous, double* out) {
  real x[3] = {(real)x0[0], (real)x0[1], (real)x0[2]};
  force_moment(*(Sim*)h, *b, (real)t, x, viscous, out);
}

// ---- single operations, for kernel-level parity tests
void wlo_op_conv_diff(void* h, int use_u0) {
  Sim* s = (Sim*)h;
  conv_diff(s->g, s->f.data(), use_u0 ? s->u0.data() : s->u.data(), s->sigma.data(), s->cfg.scheme, s->cfg.nu, s->per);
}
void wlo_op_bdim(void* h) {
  Sim* s = (Sim*)h;
  BDIM(s->g, s->u.data(), s->u0.data(), s->f.data(), s->V.data(), s->mu0.data(), s->mu1.data(), s->dts.back());
}
void wlo_op_scale_u(void* h, real sc) { Sim* s = (Sim*)h; scale_u(s->g, s->u.data(), sc); }
void wlo_op_project(void* h, real w) {
  Sim* s = (Sim*)h;
  real U[3] = {(real)s->cfg.uBC[0], (real)s->cfg.uBC[1], (real)s->cfg.uBC[2]};
  mom_project(*s, w, BCSpec{U, nullptr, 0}, &s->last_trace[0]);
}
void wlo_op_residual(void* h, int l) { residual(((Sim*)h)->ml.lv[l]); }
void wlo_op_jacobi(void* h, int l) { Jacobi(((Sim*)h)->ml.lv[l]); }
void wlo_op_smooth(void* h, int l, real w) { GaussSeidelRB(((Sim*)h)->ml.lv[l], 4, w); }
void wlo_op_increment(void* h, int l, real w) { increment(((Sim*)h)->ml.lv[l], w); }
void wlo_op_restrict(void* h, int l) { Sim* s = (Sim*)h; restrict_r(s->ml.lv[l + 1], s->ml.lv[l]); }
void wlo_op_prolongate(void* h, int l) { Sim* s = (Sim*)h; prolongate(s->ml.lv[l], s->ml.lv[l + 1]); }
void wlo_op_pcg(void* h, int l, int it) { pcg(((Sim*)h)->ml.lv[l], it); }
void wlo_op_vcycle(void* h, real w) { Vcycle(((Sim*)h)->ml, 0, w); }
real wlo_op_L1(void* h, int l) { return L1(((Sim*)h)->ml.lv[l]); }
real wlo_op_Linf(void* h, int l) { return Linf(((Sim*)h)->ml.lv[l]); }

// ---- free functions on caller-owned arrays (mirror test/test_core.jl, test/test_flow.jl)
void wlo_BC(real* a, int D, const int* ng, const real* U, int saveexit, const int* per) {
  Grid g; g.init(D, ng); BC(g, a, U, saveexit != 0, per);
}
void wlo_exitBC(real* u, const real* u0, int D, const int* ng, real dt) { Grid g; g.init(D, ng); exitBC(g, u, u0, dt); }
void wlo_perBC(real* a, int D, const int* ng, const int* per) { Grid g; g.init(D, ng); perBC(g, a, per); }
real wlo_scheme(int scheme, real u, real c, real d) { return lam(scheme, u, c, d); }
real wlo_mu0(real d, real e) { return mu0f(d, e); }
real wlo_mu1(real d, real e) { return mu1f(d, e); }
void wlo_measure_point(const wlo_body* b, int D, const real* x, real t, real fastd2, real* d, real* n, real* V) {
  Meas m = measure(*b, D, x, t, fastd2);
  *d = m.d; for (int k = 0; k < 3; ++k) { n[k] = m.n[k]; V[k] = m.V[k]; }
}

}  // extern "C"
