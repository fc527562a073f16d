// wl_kernels_body.cuh -- device-side measure! for parametric bodies and pressure_force.
// Reference: src/Body.jl:28-60, src/AutoBody.jl:29-37, src/Metrics.jl:116-133.
#pragma once
#include "wl_common.cuh"

namespace wl {

struct BodyDev {
  int kind, axis;
  float c[3];      // centre(t), evaluated on the host in Float32
  float Vb[3];     // d centre / dt
  float radius, minor;
};

struct Meas { float d; float n[3]; float V[3]; };

// measure(body::AutoBody,x,t;fastd²) AutoBody.jl:29-37 with J = I and the closed-form gradient.
template <int D>
__device__ __forceinline__ Meas measure_pt(const BodyDev& b, const float* x, float fastd2) {
  Meas m;
  m.d = INFINITY;
#pragma unroll
  for (int q = 0; q < 3; ++q) m.n[q] = m.V[q] = 0.f;
  if (b.kind == 0) return m;
  float xi[3] = {0.f, 0.f, 0.f}, g[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < D; ++q) xi[q] = x[q] - b.c[q];
  float d;
  if (b.kind == 1) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < D; ++q) s += xi[q] * xi[q];
    const float rt = sqrtf(s);
    d = rt - b.radius;
#pragma unroll
    for (int q = 0; q < D; ++q) g[q] = xi[q] / rt;
  } else if (b.kind == 2) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < D; ++q) if (q != b.axis) s += xi[q] * xi[q];
    const float rt = sqrtf(s);
    d = rt - b.radius;
#pragma unroll
    for (int q = 0; q < D; ++q) g[q] = q == b.axis ? 0.f : xi[q] / rt;
  } else {
    float s = 0.f, xa = 0.f;
#pragma unroll
    for (int q = 0; q < D; ++q) { if (q != b.axis) s += xi[q] * xi[q]; else xa = xi[q]; }
    const float rho = sqrtf(s), qq = rho - b.radius;
    const float len = sqrtf(qq * qq + xa * xa);
    d = len - b.minor;
#pragma unroll
    for (int q = 0; q < D; ++q) g[q] = q == b.axis ? xi[q] / len : (qq / len) * (xi[q] / rho);
  }
  m.d = d;
  if (d * d > fastd2) return m;
  bool nan = false;
#pragma unroll
  for (int q = 0; q < D; ++q) nan |= isnan(g[q]);
  if (nan) return m;
  float mm = 0.f;
#pragma unroll
  for (int q = 0; q < D; ++q) mm += g[q] * g[q];
  mm = sqrtf(mm);
  m.d = d / mm;
#pragma unroll
  for (int q = 0; q < D; ++q) { m.n[q] = g[q] / mm; m.V[q] = b.Vb[q]; }
  return m;
}

template <int D>
__device__ __forceinline__ float sdf_fill(const BodyDev& b, const float* x, float) { return measure_pt<D>(b, x, 0.f).d; }

// ---- SetBody programs over RigidMap leaves (Body.jl:91-107, RigidMap.jl:28-50) ------------------------------
constexpr int CSG_MAX = 8;
struct NodeDev {
  int op, kind, axis;          // op: 0 leaf, 1 min (∪,+), 2 max (∩), 3 complement
  float radius, minor;
  float x0[3], xp[3], R[9], V[3], w[3];   // R̂ = rotation(θ), row-major, formed on the host (RigidMap.jl:45-49)
};
struct BodyProg { int n; NodeDev nd[CSG_MAX]; };

// measure(AutoBody(sdf,RigidMap),x,t;fastd²): ξ = R̂(x−x₀−xₚ)+xₚ, n = R̂ᵀ∇sdf(ξ) normalised, V = V + ω×(x−x₀−xₚ)
template <int D>
__device__ __forceinline__ Meas measure_leaf(const NodeDev& b, const float* x, float fastd2) {
  Meas m;
  m.d = INFINITY;
#pragma unroll
  for (int q = 0; q < 3; ++q) m.n[q] = m.V[q] = 0.f;
  if (b.kind == 0) return m;
  float r[3] = {0.f, 0.f, 0.f}, xi[3] = {0.f, 0.f, 0.f}, g[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < D; ++q) r[q] = x[q] - b.x0[q] - b.xp[q];
#pragma unroll
  for (int k = 0; k < D; ++k) {
    float a = b.R[3 * k] * r[0];
#pragma unroll
    for (int q = 1; q < D; ++q) a += b.R[3 * k + q] * r[q];
    xi[k] = a + b.xp[k];
  }
  float d;
  if (b.kind == 1) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < D; ++q) s += xi[q] * xi[q];
    const float rt = sqrtf(s);
    d = rt - b.radius;
#pragma unroll
    for (int q = 0; q < D; ++q) g[q] = xi[q] / rt;
  } else if (b.kind == 2) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < D; ++q) if (q != b.axis) s += xi[q] * xi[q];
    const float rt = sqrtf(s);
    d = rt - b.radius;
#pragma unroll
    for (int q = 0; q < D; ++q) g[q] = q == b.axis ? 0.f : xi[q] / rt;
  } else if (b.kind == 3) {
    float s = 0.f, xa = 0.f;
#pragma unroll
    for (int q = 0; q < D; ++q) { if (q != b.axis) s += xi[q] * xi[q]; else xa = xi[q]; }
    const float rho = sqrtf(s), qq = rho - b.radius;
    const float len = sqrtf(qq * qq + xa * xa);
    d = len - b.minor;
#pragma unroll
    for (int q = 0; q < D; ++q) g[q] = q == b.axis ? xi[q] / len : (qq / len) * (xi[q] / rho);
  } else {                                                   // capsule
    float y[3] = {xi[0], xi[1], xi[2]};
    const float h = b.minor;
#pragma unroll
    for (int q = 0; q < D; ++q) if (q == b.axis) y[q] = xi[q] - fmaxf(-h, fminf(xi[q], h));
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < D; ++q) s += y[q] * y[q];
    const float len = sqrtf(s);
    d = len - b.radius;
#pragma unroll
    for (int q = 0; q < D; ++q) g[q] = y[q] / len;
  }
  m.d = d;
  if (d * d > fastd2) return m;
  bool nan = false;
#pragma unroll
  for (int q = 0; q < D; ++q) nan |= isnan(g[q]);
  if (nan) return m;
  float nn[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < D; ++k) {
    float a = b.R[k] * g[0];
#pragma unroll
    for (int q = 1; q < D; ++q) a += b.R[3 * q + k] * g[q];
    nn[k] = a;
  }
  float mm = 0.f;
#pragma unroll
  for (int q = 0; q < D; ++q) mm += nn[q] * nn[q];
  mm = sqrtf(mm);
  m.d = d / mm;
  float cr[3];
  if (D == 2) { cr[0] = b.w[2] * -r[1]; cr[1] = b.w[2] * r[0]; cr[2] = 0.f; }
  else { cr[0] = b.w[1] * r[2] - b.w[2] * r[1]; cr[1] = b.w[2] * r[0] - b.w[0] * r[2]; cr[2] = b.w[0] * r[1] - b.w[1] * r[0]; }
#pragma unroll
  for (int q = 0; q < D; ++q) { m.n[q] = nn[q] / mm; m.V[q] = b.V[q] + cr[q]; }
  return m;
}

// Base.isless on floats and on (d, n, V) tuples: NaN greatest, −0.0 < 0.0, lexicographic
__device__ __forceinline__ bool fless(float a, float b) {
  return !isnan(a) && (isnan(b) || (signbit(a) && !signbit(b) && a == b) || a < b);
}
__device__ __forceinline__ bool fequal(float a, float b) { return (isnan(a) && isnan(b)) || (a == b && signbit(a) == signbit(b)); }
template <int D>
__device__ __forceinline__ bool meas_less(const Meas& a, const Meas& b) {
  if (fless(a.d, b.d)) return true;
  if (!fequal(a.d, b.d)) return false;
#pragma unroll
  for (int k = 0; k < D; ++k) { if (fless(a.n[k], b.n[k])) return true; if (fless(b.n[k], a.n[k])) return false; }
#pragma unroll
  for (int k = 0; k < D; ++k) { if (fless(a.V[k], b.V[k])) return true; if (fless(b.V[k], a.V[k])) return false; }
  return false;
}

// measure(body::SetBody,x,t;fastd²) Body.jl:104-107 over the postfix program (validated on the host)
template <int D>
__device__ __forceinline__ Meas measure_pt(const BodyProg& p, const float* x, float fastd2) {
  Meas st[CSG_MAX];
  int sp = 0;
  for (int i = 0; i < p.n; ++i) {
    const NodeDev& b = p.nd[i];
    if (b.op == 0) st[sp++] = measure_leaf<D>(b, x, fastd2);
    else if (b.op == 3) {
      Meas& a = st[sp - 1];
      a.d = -a.d;
#pragma unroll
      for (int k = 0; k < D; ++k) a.n[k] = -a.n[k];
    } else {
      const Meas bb = st[--sp];
      Meas& a = st[sp - 1];
      if (b.op == 1) { if (meas_less<D>(bb, a)) a = bb; }
      else           { if (meas_less<D>(a, bb)) a = bb; }
    }
  }
  return st[0];
}
// measure_sdf!(σ,body,t;fastd²): raw sdf for a lone AutoBody (AutoBody.jl:19), measure(...)[1] for a SetBody (Body.jl:66)
template <int D>
__device__ __forceinline__ float sdf_fill(const BodyProg& p, const float* x, float fastd2) {
  if (p.n == 1) return measure_leaf<D>(p.nd[0], x, 0.f).d;
  return measure_pt<D>(p, x, fastd2).d;
}

// kernels Body.jl:54-60
// Julia's sinpi/cospi are accurate to the last bit: evaluate in double and round once (the CPU
// checker does the same), so μ₀/μ₁ agree bit for bit.  Only band cells reach these, the cost is nil.
__device__ __forceinline__ float sinpi_r(float d) { return (float)sin(3.14159265358979323846 * (double)d); }
__device__ __forceinline__ float cospi_r(float d) { return (float)cos(3.14159265358979323846 * (double)d); }
__device__ __forceinline__ float kern(float d) { return (1.f + cospi_r(d)) / 2.f; }
__device__ __forceinline__ float kern0(float d) { return (1.f + d + sinpi_r(d) / 3.14159265358979323846f) / 2.f; }
__device__ __forceinline__ float kern1(float d) {
  const float pi = 3.14159265358979323846f;
  return (1.f - d * d) / 4.f - (d * sinpi_r(d) + (1.f + cospi_r(d)) / pi) / (2.f * pi);
}
__device__ __forceinline__ float mu0f(float d, float e) {
  const float ad = fabsf(d);
  const float ulp = nextafterf(ad, INFINITY) - ad;        // eps(d)
  return d / e < -1.f + sqrtf(ulp) ? 0.f : kern0(fminf(d / e, 1.f));
}
__device__ __forceinline__ float mu1f(float d, float e) { return e * kern1(fmaxf(-1.f, fminf(d / e, 1.f))); }

template <class B>
struct MeasArgsT {
  Lay l;
  B b;
  float eps;
  int kz0;           // global index of local plane 0 (slab decomposition)
  float* sigma; float* V; float* mu0; float* mu1;
  int* band_list; int* band_count; int band_cap;   // interior cells with any μ₁ ≠ 0, appended here (K-MOM's band patch)
};

// measure! (Body.jl:28-48) in one pass over ALL cells: the three fills of :29, measure_sdf! :30
// and the band loop :31-48.  Ghost cells get V=0, μ₀=1, μ₁=0 (σ untouched); BC! follows.
using MeasArgs = MeasArgsT<BodyDev>;
template <int D, class B>
__global__ void __launch_bounds__(256) k_measure(const MeasArgsT<B> a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  const int o = l.off(i, j, k);
  const bool in = i >= 1 && i <= l.n0 - 2 && j >= 1 && j <= l.n1 - 2 && (D == 2 || (k >= 1 && k <= l.n2 - 2));
  float Vv[3] = {0.f, 0.f, 0.f}, m0[3] = {1.f, 1.f, 1.f}, m1[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (in) {
    const float d2 = (2.f + a.eps) * (2.f + a.eps);
    float x[3] = {(float)i - 0.5f, (float)j - 0.5f, (float)(k + a.kz0) - 0.5f};
    const float sd = sdf_fill<D>(a.b, x, d2);               // measure_sdf!(a.σ,body,t;fastd²=d²)
    a.sigma[o] = sd;
    if (sd * sd < d2) {
#pragma unroll
      for (int ci = 0; ci < D; ++ci) {
        float xf[3] = {x[0], x[1], x[2]};
        xf[ci] -= 0.5f;                                       // loc(i,I) core.jl:177
        const Meas m = measure_pt<D>(a.b, xf, d2);
        const float di = fabsf(m.d) <= 0.5f ? m.d : copysignf(m.d, sd);
        Vv[ci] = m.V[ci];
        m0[ci] = mu0f(di, a.eps);
        const float w1 = mu1f(di, a.eps);
#pragma unroll
        for (int dj = 0; dj < D; ++dj) m1[ci + D * dj] = w1 * m.n[dj];
      }
    } else if (sd < 0.f) {
#pragma unroll
      for (int ci = 0; ci < D; ++ci) m0[ci] = 0.f;
    }
  }
#pragma unroll
  for (int ci = 0; ci < D; ++ci) {
    a.V[l.cstride * ci + o] = Vv[ci];
    a.mu0[l.cstride * ci + o] = m0[ci];
  }
#pragma unroll
  for (int q = 0; q < D * D; ++q) a.mu1[l.cstride * q + o] = m1[q];
  if (a.band_count && in) {                                   // same predicate as k_band_scan, on the values just written
    bool any = false;
#pragma unroll
    for (int q = 0; q < D * D; ++q) any |= m1[q] != 0.f;
    if (any) {
      const int idx = atomicAdd(a.band_count, 1);
      if (idx < a.band_cap) a.band_list[idx] = o;
    }
  }
}

// pressure_force (Metrics.jl:126-133): df = p·nds over inside (df ≡ flow.f, zero elsewhere).
template <class B> struct ForceArgsT { Lay l; B b; const float* p; float* f; int kz0; };
using ForceArgs = ForceArgsT<BodyDev>;
template <int D, class B>
__global__ void __launch_bounds__(256) k_force_df(const ForceArgsT<B> a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  const int o = l.off(i, j, k);
  const bool in = i >= 1 && i <= l.n0 - 2 && j >= 1 && j <= l.n1 - 2 && (D == 2 || (k >= 1 && k <= l.n2 - 2));
  float df[3] = {0.f, 0.f, 0.f};
  if (in) {
    const float x[3] = {(float)i - 0.5f, (float)j - 0.5f, (float)(k + a.kz0) - 0.5f};
    const Meas m = measure_pt<D>(a.b, x, 1.f);
    const float kq = kern(fmaxf(-1.f, fminf(m.d, 1.f)));
    const float pv = a.p[o];
#pragma unroll
    for (int ci = 0; ci < D; ++ci) df[ci] = pv * (m.n[ci] * kq);
  }
#pragma unroll
  for (int ci = 0; ci < D; ++ci) a.f[l.cstride * ci + o] = df[ci];
}

// viscous_force (Metrics.jl:140-152): df = -2ν·S(I,u)·nds over inside_u(u), S the rate-of-strain tensor
// (:134) from the staggered differences ∂(i,j,I,u) (:42-45); summed per component by k_sum_all.
template <class B> struct ViscArgsT { Lay l; B b; const float* u; float* f; float nu; int kz0; };
using ViscArgs = ViscArgsT<BodyDev>;
template <int D, class B>
__global__ void __launch_bounds__(256) k_visc_df(const ViscArgsT<B> a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  const int o = l.off(i, j, k);
  const bool in = i >= 1 && i <= l.n0 - 2 && j >= 1 && j <= l.n1 - 2 && (D == 2 || (k >= 1 && k <= l.n2 - 2));
  float df[3] = {0.f, 0.f, 0.f};
  if (in) {
    const float x[3] = {(float)i - 0.5f, (float)j - 0.5f, (float)(k + a.kz0) - 0.5f};
    const Meas m = measure_pt<D>(a.b, x, 1.f);
    const float kq = kern(fmaxf(-1.f, fminf(m.d, 1.f)));
    const int st[3] = {1, l.pitch, l.plane};
    float g[3][3];                                           // g[c][d] = ∂u_c/∂x_d
#pragma unroll
    for (int c = 0; c < D; ++c) {
      const float* uc = a.u + l.cstride * c;
#pragma unroll
      for (int d = 0; d < D; ++d)
        g[c][d] = c == d ? uc[o + st[c]] - uc[o]
                         : (uc[o + st[d]] + uc[o + st[d] + st[c]] - uc[o - st[d]] - uc[o - st[d] + st[c]]) / 4.f;
    }
    const float m2nu = -2.f * a.nu;
#pragma unroll
    for (int c = 0; c < D; ++c) {
      float acc = 0.f;
#pragma unroll
      for (int d = 0; d < D; ++d) acc += (m2nu * ((g[c][d] + g[d][c]) / 2.f)) * (m.n[d] * kq);
      df[c] = acc;
    }
  }
#pragma unroll
  for (int ci = 0; ci < D; ++ci) a.f[l.cstride * ci + o] = df[ci];
}

// pressure_moment / viscous_moment (Metrics.jl:168-192): df = p·cross(x-x₀, nds) or -2ν·cross(x-x₀, S·nds);
// in 2D the cross product is the scalar a₁b₂-a₂b₁, stored in both components like the reference's broadcast.
template <class B> struct MomentArgsT { Lay l; B b; const float* p; const float* u; float* f; float nu; int kz0; int viscous; float x0[3]; };
using MomentArgs = MomentArgsT<BodyDev>;
template <int D, class B>
__global__ void __launch_bounds__(256) k_moment_df(const MomentArgsT<B> a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  const int o = l.off(i, j, k);
  const bool in = i >= 1 && i <= l.n0 - 2 && j >= 1 && j <= l.n1 - 2 && (D == 2 || (k >= 1 && k <= l.n2 - 2));
  float df[3] = {0.f, 0.f, 0.f};
  if (in) {
    const float x[3] = {(float)i - 0.5f, (float)j - 0.5f, (float)(k + a.kz0) - 0.5f};
    const Meas m = measure_pt<D>(a.b, x, 1.f);
    const float kq = kern(fmaxf(-1.f, fminf(m.d, 1.f)));
    float v[3] = {0.f, 0.f, 0.f}, r[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < D; ++c) { v[c] = m.n[c] * kq; r[c] = x[c] - a.x0[c]; }
    if (a.viscous) {
      const int st[3] = {1, l.pitch, l.plane};
      float g[3][3];
#pragma unroll
      for (int c = 0; c < D; ++c) {
        const float* uc = a.u + l.cstride * c;
#pragma unroll
        for (int d = 0; d < D; ++d)
          g[c][d] = c == d ? uc[o + st[c]] - uc[o]
                           : (uc[o + st[d]] + uc[o + st[d] + st[c]] - uc[o - st[d]] - uc[o - st[d] + st[c]]) / 4.f;
      }
      float w[3] = {0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < D; ++c) {
        float acc = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) acc += ((g[c][d] + g[d][c]) / 2.f) * v[d];
        w[c] = acc;
      }
#pragma unroll
      for (int c = 0; c < D; ++c) v[c] = w[c];
    }
    const float sc = a.viscous ? -2.f * a.nu : a.p[o];
    if (D == 3) {
      df[0] = sc * (r[1] * v[2] - r[2] * v[1]);
      df[1] = sc * (r[2] * v[0] - r[0] * v[2]);
      df[2] = sc * (r[0] * v[1] - r[1] * v[0]);
    } else {
      df[0] = df[1] = sc * (r[0] * v[1] - r[1] * v[0]);
    }
  }
#pragma unroll
  for (int ci = 0; ci < D; ++ci) a.f[l.cstride * ci + o] = df[ci];
}

// update!(meanflow,flow) (Metrics.jl:239-249): P, U and (optionally) the u⊗u tensor as running averages over
// ALL indices, ghosts included; ε = dt/(dt + T_avg + eps(T)) is formed by the caller, who owns the time vector.
struct MeanArgs { Lay l; const float* p; const float* u; float* P; float* U; float* UU; float eps; };
template <int D>
__global__ void __launch_bounds__(256) k_mean_update(const MeanArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  if (i < 0 || i >= l.n0 || j >= l.n1 || k >= l.n2) return;
  const int o = l.off(i, j, k);
  const float e = a.eps, om = 1.f - a.eps;
  a.P[o] = e * a.p[o] + om * a.P[o];
  float uc[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int c = 0; c < D; ++c) {
    uc[c] = a.u[l.cstride * c + o];
    a.U[l.cstride * c + o] = e * uc[c] + om * a.U[l.cstride * c + o];
  }
  if (a.UU) {
#pragma unroll
    for (int c = 0; c < D; ++c)
#pragma unroll
      for (int d = 0; d < D; ++d) {
        float* q = a.UU + l.cstride * (c + D * d) + o;
        *q = e * (uc[c] * uc[d]) + om * *q;
      }
  }
}

// Σ over the whole array of one scalar, in Float64 (sum(To,df,dims=...) Metrics.jl:132)
struct SumArgs { Lay l; const float* a; RedWork w; Red* out; };
template <int D>
__global__ void __launch_bounds__(256) k_sum_all(const SumArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const int i = blockIdx.x * 32 + threadIdx.x - LPAD;
  const int j = blockIdx.y * blockDim.y + threadIdx.y;
  const int k = D == 3 ? blockIdx.z * blockDim.z + threadIdx.z : 0;
  double s = 0.0;
  if (i >= 0 && i < l.n0 && j < l.n1 && k < l.n2) s = (double)a.a[l.off(i, j, k)];
  block_reduce_finish(s, -INFINITY, a.w, a.out);
}

// pressure_force for a HOST-measured body (an arbitrary AutoBody closure cannot cross the C ABI): the caller
// evaluates nds(body,loc(0,I),t) = n·kern(clamp(d,-1,1)) (src/Metrics.jl:110-113) on the cells with |d| < 1 and
// uploads them as a sparse list; df[I,c] = p[I]·w[I,c] is formed in Float32 as the reference's `@loop` does
// (:130) and summed in Float64 (:131).  One launch per component (the block reduction owns the tickets).
struct ForceWArgs { const float* p; const int* off; const float* w; int nnz, D, comp; RedWork rw; Red* out; };
__global__ void __launch_bounds__(256) k_force_w(const ForceWArgs a) {
  pdl_entry();
  double s = 0.0;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < a.nnz; q += gridDim.x * blockDim.x) {
    const float df = a.p[a.off[q]] * a.w[(size_t)q * a.D + a.comp];
    s += (double)df;
  }
  block_reduce_finish(s, -INFINITY, a.rw, a.out);
}

// ---------------------------------------------------------------- host <-> device repacking
// The C ABI moves fields in the reference's dense layout (n0,n1,n2[,comp]); one contiguous DMA to or
// from a packed device buffer plus this kernel replaces a strided cudaMemcpy2D per component (whose
// 520-byte rows ran at a fraction of the PCIe rate).
struct PackArgs { Lay l; float* field; float* packed; int ncomp; int to_field; };
__global__ void __launch_bounds__(256) k_repack(const PackArgs a) {
  pdl_entry();
  const Lay& l = a.l;
  const long long per = (long long)l.n0 * l.n1 * l.n2, tot = per * a.ncomp;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < tot; q += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(q / per);
    long long r = q - c * per;
    const int k = (int)(r / ((long long)l.n0 * l.n1));
    r -= (long long)k * l.n0 * l.n1;
    const int j = (int)(r / l.n0), i = (int)(r - (long long)j * l.n0);
    float* f = a.field + l.cstride * c + l.off(i, j, k);
    if (a.to_field) *f = a.packed[q]; else a.packed[q] = *f;
  }
}

}  // namespace wl
