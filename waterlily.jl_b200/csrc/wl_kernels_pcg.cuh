// wl_kernels_pcg.cuh -- stand-alone `Poisson` solver: `pcg!(p;it=6)` and `solver!(p::Poisson)`
// (reference src/Poisson.jl:150-223).
//
// Jacobi-preconditioned conjugate gradient on ONE level.  The reference interleaves `@inside` loops
// with three dot products per iteration whose values steer the host control flow (early exits on
// |ρ| < 10eps, α outside [1e-2,1e2]); here every loop that feeds a dot product also produces its
// block partials, so an iteration is three kernels and three scalar read-backs:
//   K-PCG-MULT   (after perBC!(ϵ)) z = Aϵ ; Σ z·ϵ                               (Poisson.jl:173-175)
//   K-PCG-UPD    x += αϵ ; r -= αz ; [z = r·iD ; Σ r·z]                          (:177-181)
//   K-PCG-DIR    ϵ = βϵ + z                                                      (:184)
// Dot products accumulate in double in block order (block_reduce_finish), like every other
// reduction of the library: the reference's `⋅` is BLAS/pairwise and not reproducible bit for bit.
// `r⋅z` runs over the whole array in the reference; r's ghosts are zero (Poisson.jl:190), so the
// interior sum is the same number.
#pragma once
#include "wl_kernels_mg2.cuh"

namespace wl {

struct PcgArgs {
  Lay l;
  const float* L;
  const float* iD;
  float* x;
  float* r;
  float* eps;
  float* z;
  float coef;        // α (K-PCG-UPD) or β (K-PCG-DIR)
  int more;          // K-PCG-UPD: also z = r·iD and Σ r·z (not the last iteration)
  RedWork w;
  Red* out;
};

// z = ϵ = r·iD inside; Σ r·z  (Poisson.jl:168-169)
template <int D>
__global__ void __launch_bounds__(256) k_pcg_init(const PcgArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  double s = 0.0;
  if (inb) {
    const int o = l.off(i, j, k);
    const float r = a.r[o], z = r * a.iD[o];
    a.z[o] = z; a.eps[o] = z;
    s = (double)(r * z);
  }
  block_reduce_finish(s, -INFINITY, a.w, a.out);
}

// z = mult(I,L,D,ϵ) inside; Σ z·ϵ  (Poisson.jl:174-175).  The caller has applied perBC!(ϵ) (k_perbc), so
// ϵ's ghost cells hold what the reference leaves there.
template <int D>
__global__ void __launch_bounds__(256) k_pcg_mult(const PcgArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  double s = 0.0;
  if (inb) {
    const int o = l.off(i, j, k);
    const float z = multA<D>(l, a.L, a.eps, o);
    a.z[o] = z;
    s = (double)(z * a.eps[o]);
  }
  block_reduce_finish(s, -INFINITY, a.w, a.out);
}

// x += αϵ ; r -= αz ; then (more) z = r·iD and Σ r·z  (Poisson.jl:177-181)
template <int D>
__global__ void __launch_bounds__(256) k_pcg_update(const PcgArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  double s = 0.0;
  if (inb) {
    const int o = l.off(i, j, k);
    a.x[o] = a.x[o] + a.coef * a.eps[o];
    const float r = a.r[o] - a.coef * a.z[o];
    a.r[o] = r;
    if (a.more) {
      const float z = r * a.iD[o];
      a.z[o] = z;
      s = (double)(r * z);
    }
  }
  if (a.more) block_reduce_finish(s, -INFINITY, a.w, a.out);
}

// ϵ = βϵ + z  (Poisson.jl:184)
template <int D>
__global__ void __launch_bounds__(256) k_pcg_dir(const PcgArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  WL_INTERIOR_IJK(D, l)
  if (!inb) return;
  const int o = l.off(i, j, k);
  a.eps[o] = a.coef * a.eps[o] + a.z[o];
}

}  // namespace wl
